"""GPU parity: each CUDA entry point (through the C ABI) against the CPU oracle.

Bars (BASELINE.json north_star): event binning bit-exact; LIF/LI forward bit-exact against the fp32
restatement (same op order, no FMA contraction); conv within 2e-6 relative of an fp64 reference
(fp32 products, fp32 accumulation); pooling / merges bit-exact.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import golden
from oracle import binning as ob
from oracle import norse_restated as nr

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _pkg():
    import snn_for_object_detection_b200 as pkg

    return pkg


def _unpack(bits, shape):
    return np.unpackbits(bits)[: int(np.prod(shape))].reshape(shape).astype(np.float32)


# ----------------------------------------------------------------------------------------------
# event binning
# ----------------------------------------------------------------------------------------------
def _bin(samples, t0, step, T, H, W, clip_x=True, sorted_by_time=True, out=None, count_dropped=False):
    pkg = _pkg()
    ev, offs = pkg.events.pack_samples(samples)
    nd = torch.zeros(1, dtype=torch.int32, device=DEV) if count_dropped else None
    f = pkg.ops.bin_events(ev.to(DEV), offs.to(DEV), torch.tensor(t0, dtype=torch.int64, device=DEV), step, T, H, W,
                           clip_x=clip_x, sorted_by_time=sorted_by_time, out=out, n_dropped=nd)
    return (f, int(nd.item())) if count_dropped else f


def test_bin_golden_st_bit_exact():
    g = golden("binning_st.npz")
    T, H, W = int(g["T"]), int(g["H"]), int(g["W"])
    ref = _unpack(g["features_bits"], (T, 2, H, W))
    for srt in (True, False):
        f = _bin([g["records"]], [int(g["first_event_time_us"])], int(g["step_us"]), T, H, W, True, srt)
        assert np.array_equal(f[:, 0].cpu().numpy(), ref)


def test_bin_golden_mt_bit_exact():
    g = golden("binning_mt.npz")
    T, H, W, step = int(g["T"]), int(g["H"]), int(g["W"]), int(g["step_us"])
    ref = _unpack(g["features_bits"], (2, T, 2, H, W))
    # both windows as one batch of two samples reading the same record stream
    f = _bin([g["records"], g["records"]], [0, T * step], step, T, H, W, clip_x=False)
    assert np.array_equal(f[:, 0].cpu().numpy(), ref[0])
    assert np.array_equal(f[:, 1].cpu().numpy(), ref[1])


@pytest.mark.parametrize("H,W", [(240, 304), (37, 51), (8, 4), (720, 1280)])
@pytest.mark.parametrize("srt", [True, False])
def test_bin_random_ragged_batches(H, W, srt):
    rng = np.random.default_rng(H * 1000 + W)
    T, step = 5, 16000
    counts = [0, 1, 777, 20000, 3]
    t0 = [0, 100, 5000, 16000, 7]
    samples = [ob.make_records(rng, n, 0, (T + 2) * step, H, W) for n in counts]
    f = _bin(samples, t0, step, T, H, W, clip_x=True, sorted_by_time=srt)
    ref = ob.bin_batch(samples, t0, step, T, H, W, clip_x=True)
    assert np.array_equal(f.cpu().numpy(), ref)
    assert set(np.unique(ref)) <= {0.0, 1.0}


def test_bin_out_of_frame_events_clip_and_drop():
    rng = np.random.default_rng(9)
    T, step, H, W = 3, 16000, 40, 64
    rec = ob.make_records(rng, 5000, 0, T * step, H, W, x_overflow=5)
    f = _bin([rec], [0], step, T, H, W, clip_x=True)
    assert np.array_equal(f.cpu().numpy(), ob.bin_batch([rec], [0], step, T, H, W, clip_x=True))
    f2, nd = _bin([rec], [0], step, T, H, W, clip_x=False, count_dropped=True)
    assert np.array_equal(f2.cpu().numpy(), ob.bin_batch([rec], [0], step, T, H, W, clip_x=False))
    x = rec[:, 1] & 16383
    assert nd == int((x >= W).sum())


def test_bin_writes_every_element_and_strided_out():
    rng = np.random.default_rng(2)
    T, B, step, H, W = 4, 3, 16000, 24, 32
    samples = [ob.make_records(rng, 900, 0, T * step, H, W) for _ in range(B)]
    big = torch.full((T, B + 2, 2, H, W), 7.0, device=DEV)
    out = big[:, 1 : 1 + B]
    _bin(samples, [0] * B, step, T, H, W, out=out)
    assert np.array_equal(out.cpu().numpy(), ob.bin_batch(samples, [0] * B, step, T, H, W))
    assert float(big[:, 0].min()) == 7.0 and float(big[:, -1].min()) == 7.0


def test_bin_full_size_property_gen1_and_1mpx():
    """BASELINE sizes: ones == number of distinct (t-bin, p, y, x); idempotent on duplicated input."""
    for (H, W, n_per_step, B) in ((240, 304, 10000, 5), (720, 1280, 150000, 2)):
        rng = np.random.default_rng(H)
        T, step = 32, 16000
        samples = [ob.make_records(rng, n_per_step * T, 0, T * step, H, W) for _ in range(B)]
        f = _bin(samples, [0] * B, step, T, H, W)
        for b in range(B):
            t = samples[b][:, 0].astype(np.int64) // step
            key = (t << 29) | (samples[b][:, 1].astype(np.int64) & 0x1FFFFFFF)
            assert int(f[:, b].sum().item()) == np.unique(key).size
        dup = [np.concatenate([s, s])[np.argsort(np.concatenate([s[:, 0], s[:, 0]]), kind="stable")] for s in samples]
        f2 = _bin(dup, [0] * B, step, T, H, W)
        assert torch.equal(f, f2)


# ----------------------------------------------------------------------------------------------
# LIF / LI scans
# ----------------------------------------------------------------------------------------------
def _oracle_scan(cell, x, state=None):
    outs, vs, is_ = [], [], []
    for t in range(x.shape[0]):
        o, state = cell(x[t], state)
        outs.append(o), vs.append(state.v), is_.append(state.i)
    return torch.stack(outs), torch.stack(vs), torch.stack(is_), state


@pytest.mark.parametrize("n", [1, 5, 1024, 4099])
@pytest.mark.parametrize("with_state", [False, True])
def test_lif_scan_bit_exact(n, with_state):
    pkg = _pkg()
    torch.manual_seed(n)
    T = 9
    x = torch.randn(T, n) * 3
    state = nr.LIFFeedForwardState(torch.rand(n), torch.randn(n)) if with_state else None
    with torch.no_grad():
        zo, vo, io, so = _oracle_scan(nr.LIFCell(), x, state)
    xs = x.to(DEV).reshape(T * 1, n)
    v0 = state.v.to(DEV).reshape(1, n) if with_state else None
    i0 = state.i.to(DEV).reshape(1, n) if with_state else None
    z, v, i, vd, vt, it = pkg.ops.lif_scan(xs, T, v0, i0, want_v_dec=True, want_trace=True)
    assert n < 100 or zo.sum() > 0
    assert torch.equal(z.cpu().reshape(T, n), zo)
    assert torch.equal(vt.cpu().reshape(T, n), vo) and torch.equal(it.cpu().reshape(T, n), io)
    assert torch.equal(v.cpu().reshape(n), so.v.detach()) and torch.equal(i.cpu().reshape(n), so.i)


@pytest.mark.parametrize("tanh", [False, True])
def test_li_scan(tanh):
    pkg = _pkg()
    torch.manual_seed(1)
    T, n = 8, 1003
    x = torch.randn(T, n)
    with torch.no_grad():
        vo, _, io, so = _oracle_scan(nr.LICell(), x)
    out, v, i, it = pkg.ops.li_scan(x.to(DEV), T, apply_tanh=tanh, want_trace=True)
    assert torch.equal(it.cpu(), io) and torch.equal(v.cpu().reshape(n), so.v.detach().reshape(n))
    if tanh:
        assert torch.allclose(out.cpu(), torch.tanh(vo), rtol=0, atol=2e-7)
    else:
        assert torch.equal(out.cpu(), vo)


def test_li_and_subthreshold_lif_kernels_pinned_to_the_reference_sli_cell():
    """Rows a8/a9: soda_li_fwd, and soda_lif_fwd below threshold, are bit-identical to T = 16 steps of the
    reference's OWN sli_feed_forward_step with v_st = 1e4 (= the LI update; tests/golden/li_from_sli.npz,
    models/modules/sli.py:110-126), also when the scan is cut in two with the state carried through HBM."""
    from helpers import golden

    pkg = _pkg()
    g = golden("li_from_sli.npz")
    x = torch.from_numpy(g["x_big"]).to(DEV)
    T = x.shape[0]
    v_ref, i_ref = torch.from_numpy(g["v_big"]), torch.from_numpy(g["i_big"])
    out, v, i, it = pkg.ops.li_scan(x.reshape(T, -1), T, want_trace=True)
    assert torch.equal(out.cpu().reshape(v_ref.shape), v_ref) and torch.equal(it.cpu().reshape(i_ref.shape), i_ref)
    assert torch.equal(v.cpu().reshape(v_ref[-1].shape), v_ref[-1])
    h = 7
    _, v1, i1, _ = pkg.ops.li_scan(x[:h].reshape(h, -1), h)
    out2, v2, i2, _ = pkg.ops.li_scan(x[h:].reshape(T - h, -1), T - h, v0=v1, i0=i1)
    assert torch.equal(out2.cpu().reshape(v_ref[h:].shape), v_ref[h:])
    assert torch.equal(i2.cpu().reshape(i_ref[-1].shape), i_ref[-1])
    # LIF below threshold: never spikes, never resets => the same (v, i) trajectory
    x = torch.from_numpy(g["x_sub"]).to(DEV)
    v_ref, i_ref = torch.from_numpy(g["v_sub"]), torch.from_numpy(g["i_sub"])
    z, v, i, _, vt, it = pkg.ops.lif_scan(x.reshape(T, -1), T, want_trace=True)
    assert float(z.abs().max()) == 0.0
    assert torch.equal(vt.cpu().reshape(v_ref.shape), v_ref) and torch.equal(it.cpu().reshape(i_ref.shape), i_ref)


def test_lif_decay_first_order_switch():
    pkg = _pkg()
    torch.manual_seed(4)
    T, n = 6, 300
    x = torch.randn(T, n) * 3
    nr.ORDER = "decay_first"
    try:
        with torch.no_grad():
            zo, vo, io, _ = _oracle_scan(nr.LIFCell(), x)
    finally:
        nr.ORDER = "jump_first"
    p = pkg.ops.LIF_DEFAULT._replace(jump_first=0)
    z, v, i, _, vt, it = pkg.ops.lif_scan(x.to(DEV), T, p=p, want_trace=True)
    assert torch.equal(z.cpu(), zo) and torch.equal(vt.cpu(), vo) and torch.equal(it.cpu(), io)


@pytest.mark.parametrize("detach", [False, True])
def test_lif_backward_matches_oracle_autograd(detach):
    """fp32 reverse scan vs fp64 autograd through the restated cell (SuperSpike, alpha=100)."""
    from snn_for_object_detection_b200.modules.cells import LIFCell, LIFFeedForwardState

    torch.manual_seed(3)
    T, n = 7, 515
    x = (torch.randn(T, n) * 2).double()
    v0, i0 = torch.rand(n).double(), torch.randn(n).double()
    gz = torch.randn(T, n).double()
    gv, gi = torch.randn(n).double(), torch.randn(n).double()
    xo, v0o, i0o = (t.clone().requires_grad_(True) for t in (x, v0, i0))
    nr.DETACH_RESET = detach
    try:
        state = nr.LIFFeedForwardState(v0o, i0o)
        zs = []
        for t in range(T):
            z, state = nr.lif_feed_forward_step(xo[t], state)
            zs.append(z)
        loss = (torch.stack(zs) * gz).sum() + (state.v * gv).sum() + (state.i * gi).sum()
        loss.backward()
    finally:
        nr.DETACH_RESET = False
    cell = LIFCell(detach_reset=detach)
    xc = x.float().to(DEV).requires_grad_(True)
    v0c = v0.float().to(DEV).reshape(1, n).requires_grad_(True)
    i0c = i0.float().to(DEV).reshape(1, n).requires_grad_(True)
    z, st = cell.forward_seq(xc.reshape(T, n), T, LIFFeedForwardState(v0c, i0c))
    ((z * gz.float().to(DEV)).sum() + (st.v * gv.float().to(DEV)).sum() + (st.i * gi.float().to(DEV)).sum()).backward()
    for got, want in ((xc.grad, xo.grad), (v0c.grad.reshape(n), v0o.grad), (i0c.grad.reshape(n), i0o.grad)):
        err = (got.cpu().double() - want).abs().max() / want.abs().max()
        assert float(err) < 2e-5, float(err)


# ----------------------------------------------------------------------------------------------
# convolution, pooling, merges
# ----------------------------------------------------------------------------------------------
CONV_CASES = [
    # N, Cin, H, W, Cout, k, stride, nchw_input
    (3, 2, 24, 32, 8, 3, 1, True),
    (2, 2, 30, 38, 64, 3, 2, True),
    (2, 32, 15, 19, 64, 3, 2, False),
    (2, 64, 15, 19, 27, 1, 1, False),
    (2, 64, 8, 10, 36, 1, 1, False),
    (1, 48, 9, 7, 40, 3, 1, False),
    (2, 8, 12, 16, 32, 3, 1, False),
    (1, 128, 8, 10, 256, 3, 2, False),
    (2, 256, 4, 5, 128, 1, 1, False),
]


def _conv_case(case, kind, exact_input=False):
    pkg = _pkg()
    N, Cin, H, W, Cout, k, s, nchw = case
    torch.manual_seed(Cin * 7 + Cout)
    x = torch.randint(0, 5, (N, Cin, H, W)).float() if exact_input else torch.randn(N, Cin, H, W)
    w = torch.randn(Cout, Cin, k, k) / (Cin * k * k) ** 0.5
    scale, shift = torch.rand(Cout) + 0.5, torch.randn(Cout)
    ref = F.conv2d(x.double(), (w * scale.view(-1, 1, 1, 1)).double(), stride=s, padding=k // 2) + shift.double().view(1, -1, 1, 1)
    xd = x.to(DEV) if nchw else x.to(DEV).contiguous(memory_format=torch.channels_last)
    if exact_input:
        pkg.ops.mark_exact(xd)
    pc = pkg.ops.pack_conv(w.to(DEV), s, scale.to(DEV), shift.to(DEV), kind=kind)
    y = pkg.ops.conv2d(xd, pc)
    assert y.shape == ref.shape
    err = float((y.cpu().double() - ref).abs().max() / ref.abs().max())
    # residual + channel-slice destination
    res = torch.randn_like(ref).float()
    buf = torch.zeros(N, Cout + 16, *ref.shape[2:], device=DEV).contiguous(memory_format=torch.channels_last)
    out = buf[:, 8 : 8 + Cout]
    pkg.ops.conv2d(xd, pc, residual=res.to(DEV).contiguous(memory_format=torch.channels_last), out=out)
    err2 = float((out.cpu().double() - (ref + res.double())).abs().max() / ref.abs().max())
    assert float(buf[:, :8].abs().max()) == 0.0 and float(buf[:, 8 + Cout :].abs().max()) == 0.0
    return max(err, err2)


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv2d_fp32_pipe_against_fp64(case):
    assert _conv_case(case, 0) < 2e-6


TC_CASES = [c for c in CONV_CASES if c[1] % 32 == 0 and not c[7]] + [
    (2, 64, 30, 38, 64, 3, 1, False),
    (3, 32, 17, 23, 32, 1, 1, False),
    (2, 96, 16, 16, 64, 1, 1, False),
    (1, 128, 30, 38, 128, 3, 1, False),
    (2, 64, 60, 76, 128, 3, 2, False),
    (2, 384, 8, 10, 256, 1, 1, False),
]


@pytest.mark.parametrize("exact_input", [False, True])
@pytest.mark.parametrize("case", TC_CASES)
def test_conv2d_tcgen05_against_fp64(case, exact_input):
    """3xTF32 split on the tensor pipe: fp32-level accuracy (bar: 1e-5 relative to the output scale)."""
    err = _conv_case(case, 1, exact_input)
    print(f"tcgen05 conv {case} exact_input={exact_input}: max rel err {err:.2e}")
    assert err < 1e-5, err


def test_conv2d_tcgen05_channel_slice_input():
    """Input is a channel slice of a wider concat buffer (pix stride > Cin)."""
    pkg = _pkg()
    torch.manual_seed(0)
    big = torch.randn(2, 96, 15, 19, device=DEV).contiguous(memory_format=torch.channels_last)
    x = big[:, 32:64]
    w = torch.randn(64, 32, 3, 3, device=DEV) * 0.1
    ref = F.conv2d(x.double(), w.double(), padding=1)
    y = pkg.ops.conv2d(x, pkg.ops.pack_conv(w, 1, kind=1))
    assert float((y.double() - ref).abs().max() / ref.abs().max()) < 1e-5


# ----------------------------------------------------------------------------------------------
# fused Conv -> BN -> LIF kernels and the residual / slice-store scan
# ----------------------------------------------------------------------------------------------
FUSED_CASES = [
    # T, B, Cin, H, W, Cout, k, s, residual
    (5, 2, 32, 30, 38, 32, 3, 1, True),
    (4, 3, 64, 17, 23, 64, 3, 1, True),
    (6, 1, 64, 32, 40, 128, 3, 2, False),
    (3, 2, 128, 16, 24, 256, 3, 2, False),
    (4, 2, 96, 15, 19, 64, 1, 1, False),
    (7, 1, 128, 8, 10, 128, 3, 1, True),
]


@pytest.mark.parametrize("case", FUSED_CASES)
def test_conv_lif_tcgen05_equals_conv_then_scan(case):
    """soda_conv_lif_fwd (state on-chip across T) == soda_conv2d_fwd + soda_lif_fwd (+ residual), bit for
    bit, including a channel-slice destination and a carried-in state; and against the CPU oracle."""
    pkg = _pkg()
    ops = pkg.ops
    T, B, Cin, H, W, Cout, k, s, with_res = case
    torch.manual_seed(T * 100 + Cout)
    x = (torch.randn(T * B, Cin, H, W) * 0.9).to(DEV).contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, k, k) / (Cin * k * k) ** 0.5 * 2.0
    scale, shift = torch.rand(Cout) + 0.5, torch.randn(Cout) * 0.2
    pc = ops.pack_conv(w.to(DEV), s, scale.to(DEV), shift.to(DEV), kind=1)
    res = x if with_res else None
    y0 = ops.conv2d(x, pc)
    z0, v0, i0, _, _, _ = ops.lif_scan(y0, T)
    ref = z0 + x if with_res else z0
    Ho, Wo = y0.shape[2:]
    buf = torch.zeros(T * B, Cout + 8, Ho, Wo, device=DEV).contiguous(memory_format=torch.channels_last)
    out = buf[:, 4 : 4 + Cout]
    z1, v1, i1 = ops.conv_lif(x, pc, T, ops.LIF_DEFAULT, residual=res, out=out)
    assert 0.01 < float(z0.mean()) < 0.9
    assert torch.equal(z1, ref) and torch.equal(v1, v0) and torch.equal(i1, i0)
    assert float(buf[:, :4].abs().max()) == 0.0 and float(buf[:, 4 + Cout :].abs().max()) == 0.0
    # streaming: first 2 steps, then the rest with the carried state
    za, va, ia = ops.conv_lif(x[: 2 * B], pc, 2, ops.LIF_DEFAULT, residual=None if res is None else x[: 2 * B])
    zb, vb, ib = ops.conv_lif(x[2 * B :], pc, T - 2, ops.LIF_DEFAULT, va, ia,
                              residual=None if res is None else x[2 * B :])
    assert torch.equal(torch.cat([za, zb]), ref) and torch.equal(vb, v0) and torch.equal(ib, i0)
    # oracle: fp64 conv -> fp32 -> restated LIFCell, time-major
    yr = (F.conv2d(x.cpu().double(), (w * scale.view(-1, 1, 1, 1)).double(), stride=s, padding=k // 2)
          + shift.double().view(1, -1, 1, 1)).float()
    cell, st, zs = nr.LIFCell(), None, []
    for t in range(T):
        z, st = cell(yr[t * B : (t + 1) * B], st)
        zs.append(z)
    zr = torch.cat(zs)
    assert float((z0.cpu() != zr).float().mean()) <= 1e-4


def test_conv_lif_direct_event_frames():
    """First layer of the detector: NCHW {0,1} event frames, Cin = 2, 3x3 stride 2 -> BN -> LIF in the direct
    FP32-pipe kernel == conv2d(fp32 kind) + lif_scan bit for bit; odd sizes, state carry, slice output."""
    pkg = _pkg()
    ops = pkg.ops
    for (T, B, H, W, Cout, k, s) in [(6, 2, 37, 51, 64, 3, 2), (4, 1, 24, 32, 16, 3, 1), (5, 3, 16, 20, 8, 1, 1)]:
        torch.manual_seed(T + Cout)
        x = (torch.rand(T * B, 2, H, W) < 0.3).float().to(DEV)
        w = torch.randn(Cout, 2, k, k) * 2.0
        scale, shift = torch.rand(Cout) + 1.0, torch.randn(Cout) * 0.1
        pc = ops.pack_conv(w.to(DEV), s, scale.to(DEV), shift.to(DEV), kind=0)
        assert ops.conv_lif_supported(x, pc)
        y0 = ops.conv2d(x, pc)
        z0, v0, i0, _, _, _ = ops.lif_scan(y0, T)
        Ho, Wo = y0.shape[2:]
        buf = torch.zeros(T * B, Cout + 8, Ho, Wo, device=DEV).contiguous(memory_format=torch.channels_last)
        z1, v1, i1 = ops.conv_lif(x, pc, T, ops.LIF_DEFAULT, out=buf[:, 4 : 4 + Cout])
        assert float(z0.mean()) > 0.005
        assert torch.equal(z1, z0) and torch.equal(v1, v0) and torch.equal(i1, i0)
        assert float(buf[:, :4].abs().max()) == 0.0 and float(buf[:, 4 + Cout :].abs().max()) == 0.0
        za, va, ia = ops.conv_lif(x[:B], pc, 1, ops.LIF_DEFAULT)
        zb, vb, ib = ops.conv_lif(x[B:], pc, T - 1, ops.LIF_DEFAULT, va, ia)
        assert torch.equal(torch.cat([za, zb]), z0) and torch.equal(vb, v0) and torch.equal(ib, i0)



def test_lif_backward_reads_a_concat_gradient_slice_in_place():
    """soda_lif_bwd_strided: grad_out as a channel slice of a wider pixel-major tensor (the gradient a Dense concat
    hands to one branch) gives bit-identical gradients to the dense call on a copy of that slice."""
    ops = _pkg().ops
    torch.manual_seed(5)
    T, B, C, H, W, Ct = 6, 2, 32, 9, 13, 96
    x = torch.randn(T * B, C, H, W, device=DEV).contiguous(memory_format=torch.channels_last) * 2
    z, v, i, vd, _, _ = ops.lif_scan(x, T, None, None, ops.LIF_DEFAULT, want_v_dec=True)
    wide = torch.randn(T * B, Ct, H, W, device=DEV).contiguous(memory_format=torch.channels_last)
    g_slice = wide[:, 32:64]
    assert not g_slice.is_contiguous(memory_format=torch.channels_last)
    gv = torch.randn_like(v)
    gi = torch.randn_like(i)
    got = ops.lif_scan_bwd(g_slice, gv, gi, vd, T, ops.LIF_DEFAULT)
    want = ops.lif_scan_bwd(g_slice.contiguous(memory_format=torch.channels_last), gv, gi, vd, T, ops.LIF_DEFAULT)
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    assert float(got[0].abs().max()) > 0

@pytest.mark.parametrize("with_res", [False, True])
def test_lif_scan_out_residual_and_slice_store(with_res):
    """soda_lif_res_fwd == soda_lif_fwd (+ residual) written into a concat slice."""
    ops = _pkg().ops
    T, B, Cc, H, W = 5, 2, 32, 15, 19
    torch.manual_seed(3)
    x = (torch.randn(T * B, Cc, H, W) * 3).to(DEV).contiguous(memory_format=torch.channels_last)
    big = torch.randn(T * B, 96, H, W, device=DEV).contiguous(memory_format=torch.channels_last)
    res = big[:, 64:96] if with_res else None
    z0, v0, i0, _, _, _ = ops.lif_scan(x, T)
    ref = z0 + res if with_res else z0
    buf = torch.zeros(T * B, 80, H, W, device=DEV).contiguous(memory_format=torch.channels_last)
    z1, v1, i1 = ops.lif_scan_out(x, T, residual=res, out=buf[:, 16:48])
    assert torch.equal(z1, ref) and torch.equal(v1, v0) and torch.equal(i1, i0)
    assert float(buf[:, :16].abs().max()) == 0.0 and float(buf[:, 48:].abs().max()) == 0.0
    za, va, ia = ops.lif_scan_out(x[: 2 * B], 2, residual=None if res is None else res[: 2 * B])
    zb, vb, ib = ops.lif_scan_out(x[2 * B :], T - 2, va, ia, residual=None if res is None else res[2 * B :])
    assert torch.equal(torch.cat([za, zb]), ref) and torch.equal(vb, v0)


@pytest.mark.parametrize("mode", ["S", "A", "M"])
@pytest.mark.parametrize("shape", [(2, 8, 24, 32), (1, 5, 9, 11), (3, 32, 15, 19)])
def test_pool2d(mode, shape):
    pkg = _pkg()
    torch.manual_seed(0)
    x = torch.randint(0, 2, shape).float() if mode == "S" else torch.randn(shape)
    if mode == "S":
        ref = F.avg_pool2d(x, 2, 2, 0) * 2 * 2
    elif mode == "A":
        ref = F.avg_pool2d(x, 2, 2)
    else:
        ref = F.max_pool2d(x, 2, 2)
    y = pkg.ops.pool2d(x.to(DEV).contiguous(memory_format=torch.channels_last), 2, 2, mode)
    if mode == "A":
        assert torch.allclose(y.cpu(), ref, rtol=1e-6, atol=1e-7)
    else:
        assert torch.equal(y.cpu(), ref)


def test_add_n_and_channel_affine():
    pkg = _pkg()
    torch.manual_seed(0)
    a, b = torch.randn(2, 12, 5, 7), torch.randn(2, 12, 5, 7)
    y = pkg.ops.add_n([a.to(DEV).contiguous(memory_format=torch.channels_last), b.to(DEV)])
    assert torch.equal(y.cpu(), torch.stack([a, b]).sum(0))
    sc, sh = torch.randn(12), torch.randn(12)
    z = pkg.ops.channel_affine(a.to(DEV), sc.to(DEV), sh.to(DEV))
    assert torch.equal(z.cpu(), a * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))


def test_cpu_tensor_is_rejected_not_computed():
    pkg = _pkg()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.lif_scan(torch.zeros(4, 8), 4)


# ----------------------------------------------------------------------------------------------
# detection decode + NMS
# ----------------------------------------------------------------------------------------------
def _det_close(got, ref):
    """Same rows in the same order: class and confidence exact, boxes to 1e-5 (GPU expf vs CPU exp)."""
    assert got.shape == ref.shape
    assert torch.equal(got[..., 0], ref[..., 0]), int((got[..., 0] != ref[..., 0]).sum())
    assert torch.equal(got[..., 1], ref[..., 1])
    assert torch.allclose(got[..., 2:], ref[..., 2:], atol=1e-5, rtol=1e-5)


def test_multibox_detection_matches_reference_golden():
    """soda_multibox_detection against the output of the reference's OWN utils/box.py:multibox_detection
    (tests/golden/box.npz, produced by oracle/make_golden.py from /root/reference)."""
    ops = _pkg().ops
    g = golden("box.npz")
    det = ops.multibox_detection(torch.from_numpy(g["cls_probs"]).to(DEV), torch.from_numpy(g["offsets"]).to(DEV),
                                 torch.from_numpy(g["anchors"]).to(DEV))
    ref = torch.from_numpy(g["det"])
    assert int((ref[..., 0] >= 0).sum()) > 0
    _det_close(det.cpu(), ref)


@pytest.mark.parametrize("A,C,B", [(13545, 3, 5), (1000, 4, 2), (1, 2, 1), (4097, 3, 3)])
def test_multibox_detection_random_against_oracle(A, C, B):
    """GEN1-sized anchor sets with many overlapping candidates per class; odd sizes; a single anchor."""
    from snn_for_object_detection_b200.utils import box

    ops = _pkg().ops
    torch.manual_seed(A + C)
    ctr = torch.rand(A, 2)
    wh = torch.rand(A, 2) * 0.2 + 0.02
    anchors = torch.cat([ctr - wh / 2, ctr + wh / 2], dim=1)
    probs = torch.softmax(torch.randn(B, A, C) * 2.0, dim=2)
    probs[:, ::7, 0] = 0.999  # background rows
    probs[0, : min(A, 5)] = torch.tensor([0.995] + [0.005 / (C - 1)] * (C - 1))  # conf below nothing; background wins
    offsets = torch.randn(B, A, 4) * 0.5
    ref = box.multibox_detection(probs, offsets, anchors)  # CPU restatement == reference (test_box_utils_match_reference)
    det = ops.multibox_detection(probs.to(DEV), offsets.to(DEV), anchors.to(DEV)).cpu()
    n_keep = int((ref[..., 0] >= 0).sum())
    assert A < 100 or n_keep > 10
    _det_close(det, ref)


def test_multibox_detection_low_confidence_rule_and_limits():
    ops = _pkg().ops
    # two identical boxes of class 0: the better one is kept, the other suppressed (class -1, conf kept);
    # a third, far away, with conf below pos_threshold: class -1 and conf = 1 - conf
    anchors = torch.tensor([[0.1, 0.1, 0.3, 0.3], [0.1, 0.1, 0.3, 0.3], [0.6, 0.6, 0.9, 0.9]], device=DEV)
    probs = torch.tensor([[[0.2, 0.8], [0.1, 0.9], [0.995, 0.005]]], device=DEV)
    det = ops.multibox_detection(probs, torch.zeros(1, 3, 4, device=DEV), anchors).cpu()[0]
    assert det[0, 0] == 0 and abs(float(det[0, 1]) - 0.9) < 1e-7
    assert det[1, 0] == -1 and abs(float(det[1, 1]) - 0.8) < 1e-7
    assert det[2, 0] == -1 and abs(float(det[2, 1]) - 0.995) < 1e-7
    with pytest.raises(_pkg()._lib.SodaError):
        ops.multibox_detection(torch.zeros(1, 262144, 3, device=DEV), torch.zeros(1, 262144, 4, device=DEV),
                               torch.zeros(262144, 4, device=DEV))


def test_multibox_detection_1mpx_anchor_count_against_oracle():
    """170 280 anchors (1Mpx): the sort keys live in the workspace instead of shared memory; rows equal the CPU
    restatement of the reference (mostly-background scores keep the oracle's Python NMS loop short)."""
    from snn_for_object_detection_b200.utils import box

    ops = _pkg().ops
    A, C, B = 170280, 3, 2
    torch.manual_seed(170)
    ctr = torch.rand(A, 2)
    wh = torch.rand(A, 2) * 0.05 + 0.01
    anchors = torch.cat([ctr - wh / 2, ctr + wh / 2], dim=1)
    logits = torch.randn(B, A, C)
    logits[:, :, 0] += 4.0                               # ~98 % background
    probs = torch.softmax(logits, dim=2)
    offsets = torch.randn(B, A, 4) * 0.3
    ref = box.multibox_detection(probs, offsets, anchors)
    det = ops.multibox_detection(probs.to(DEV), offsets.to(DEV), anchors.to(DEV)).cpu()
    assert int((ref[..., 0] >= 0).sum()) > 500
    _det_close(det, ref)


@pytest.mark.parametrize("mode", ["S", "A", "M"])
@pytest.mark.parametrize("hw", [(24, 32), (15, 19), (4, 7)])
def test_lif_pool_scan_equals_scan_then_pool(mode, hw):
    """soda_lif_pool_fwd == soda_lif_fwd followed by soda_pool2d_fwd (2x2, stride 2, floor mode) bit for bit,
    full-resolution state included (odd last row / column integrates but is not pooled), slice output, state carry."""
    ops = _pkg().ops
    H, W = hw
    T, B, Cc = 5, 2, 8
    torch.manual_seed(H * W)
    x = (torch.randn(T * B, Cc, H, W) * 3).to(DEV).contiguous(memory_format=torch.channels_last)
    z0, v0, i0, _, _, _ = ops.lif_scan(x, T)
    ref = ops.pool2d(z0, 2, 2, mode)
    buf = torch.zeros(T * B, Cc + 8, H // 2, W // 2, device=DEV).contiguous(memory_format=torch.channels_last)
    y, v1, i1 = ops.lif_pool_scan(x, T, mode, out=buf[:, 4 : 4 + Cc])
    assert float(z0.mean()) > 0.05
    assert torch.equal(y, ref) and torch.equal(v1, v0) and torch.equal(i1, i0)
    assert float(buf[:, :4].abs().max()) == 0.0 and float(buf[:, 4 + Cc :].abs().max()) == 0.0
    ya, va, ia = ops.lif_pool_scan(x[: 2 * B], 2, mode)
    yb, vb, ib = ops.lif_pool_scan(x[2 * B :], T - 2, mode, va, ia)
    assert torch.equal(torch.cat([ya, yb]), ref) and torch.equal(vb, v0) and torch.equal(ib, i0)
    # and against the CPU oracle: restated LIFCell + torch pooling
    cell, st, zs = nr.LIFCell(), None, []
    xc = x.cpu().contiguous()
    for t in range(T):
        z, st = cell(xc[t * B : (t + 1) * B], st)
        zs.append(z)
    zc = torch.cat(zs)
    pc = {"S": F.avg_pool2d(zc, 2, 2) * 4, "A": F.avg_pool2d(zc, 2, 2), "M": F.max_pool2d(zc, 2, 2)}[mode]
    assert torch.equal(y.cpu(), pc)


@pytest.mark.gpu
@pytest.mark.parametrize("case", [(2, 64, 30, 38, 64, 3, 1), (2, 64, 60, 76, 128, 3, 2), (3, 96, 17, 23, 40, 1, 1),
                                  (2, 384, 8, 10, 256, 1, 1)])
@pytest.mark.parametrize("bound", [1, 256, 257])
def test_conv2d_spike_count_inputs_bf16_only_scheme(case, bound, monkeypatch):
    """Inputs tagged as spike counts <= 256 take the bf16-only operand scheme (x exact in bf16 times
    [bf16 rn(w) | bf16 rn(w - rn(w))]); 257 falls back to tf32 + correction.  Both within the conv bar against fp64,
    and equal to each other to 1e-5."""
    pkg = _pkg()
    N, Cin, H, W, Cout, k, s = case
    torch.manual_seed(bound)
    x = torch.randint(0, bound + 1, (N, Cin, H, W)).float()
    x[0, 0, 0, 0] = bound
    w = torch.randn(Cout, Cin, k, k) / (Cin * k * k) ** 0.5
    scale, shift = torch.rand(Cout) + 0.5, torch.randn(Cout)
    ref = F.conv2d(x.double(), (w * scale.view(-1, 1, 1, 1)).double(), stride=s, padding=k // 2) + shift.double().view(1, -1, 1, 1)
    xd = pkg.ops.mark_exact(x.to(DEV).contiguous(memory_format=torch.channels_last), bound)
    pc = pkg.ops.pack_conv(w.to(DEV), s, scale.to(DEV), shift.to(DEV), kind=1)
    y = pkg.ops.conv2d(xd, pc).cpu().double()
    monkeypatch.setenv("SODA_TC_SPK", "0")
    y0 = pkg.ops.conv2d(xd, pc).cpu().double()
    sc = float(ref.abs().max())
    assert float((y - ref).abs().max()) / sc < 1e-5
    assert float((y0 - ref).abs().max()) / sc < 1e-5
    assert float((y - y0).abs().max()) / sc < 1e-5
    if bound == 257:
        assert torch.equal(y, y0)   # not bf16-exact: the flag is not set, same kernel both times


@pytest.mark.gpu
def test_fused_conv_lif_spike_inputs_bf16_only_scheme(monkeypatch):
    """Fused Conv -> BN -> LIF over T with spike inputs: the bf16-only scheme gives the same spikes as the
    tf32 + correction scheme up to the spike-mismatch bar and the same final membranes to 1e-4."""
    pkg = _pkg()
    torch.manual_seed(5)
    T, B, C, H, W = 6, 2, 32, 40, 24
    x = (torch.rand(T * B, C, H, W) < 0.3).float().to(DEV).contiguous(memory_format=torch.channels_last)
    pkg.ops.mark_exact(x)
    w = torch.randn(C, C, 3, 3) * 0.25
    pc = pkg.ops.pack_conv(w.to(DEV), 1, (torch.rand(C) + 0.5).to(DEV), (torch.randn(C) * 0.1).to(DEV), kind=1)
    z1, v1, i1 = pkg.ops.conv_lif(x, pc, T, pkg.ops.LIF_DEFAULT)
    monkeypatch.setenv("SODA_TC_SPK", "0")
    z0, v0, i0 = pkg.ops.conv_lif(x, pc, T, pkg.ops.LIF_DEFAULT)
    assert 0.01 < float(z0.mean()) < 0.9
    assert float((z1 != z0).float().mean()) <= 1e-4
    same = (z1 == z0).reshape(T, B, C, H, W).all(0)   # neurons with identical spike trains: membranes to 1e-4
    assert float(((v1 - v0).abs() / (v0.abs() + 1.0))[same].max()) < 1e-4
    assert pkg.ops.exact_bound(z1) == 1


def _roi_check(anchors, labels, thr):
    from oracle import roi_restated as rr

    ops = _pkg().ops
    off, msk, cls = ops.roi_assign(torch.from_numpy(anchors).to(DEV), torch.from_numpy(labels).to(DEV), thr)
    o_off, o_msk, o_cls = rr.roi(anchors, labels, thr)
    assert cls.dtype == torch.int64
    assert np.array_equal(cls.cpu().numpy(), o_cls)            # index work: identical
    assert np.array_equal(msk.cpu().numpy(), o_msk)
    np.testing.assert_allclose(off.cpu().numpy(), o_off, rtol=2e-6, atol=2e-6)   # logf vs np.log
    return o_cls, o_msk


@pytest.mark.gpu
def test_roi_assign_matches_reference_golden():
    """soda_roi_assign against the output of the reference's OWN utils/roi.py (tests/golden/roi.npz)."""
    ops = _pkg().ops
    g = golden("roi.npz")
    off, msk, cls = ops.roi_assign(torch.from_numpy(g["anchors"]).to(DEV), torch.from_numpy(g["labels"]).to(DEV),
                                   float(g["iou_threshold"]))
    assert np.array_equal(cls.cpu().numpy(), g["class_labels"])
    assert np.array_equal(msk.cpu().numpy(), g["bbox_mask"])
    np.testing.assert_allclose(off.cpu().numpy(), g["bbox_offset"], rtol=2e-6, atol=2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("A,G,B,thr", [(13545, 12, 5, 0.4), (700, 9, 4, 0.5), (33, 40, 2, 0.3), (4097, 1, 3, 0.5),
                                       (1000, 0, 2, 0.5)])
def test_roi_assign_random_against_oracle(A, G, B, thr):
    """GEN1-sized anchor set; ragged label counts padded with -1; more boxes than free anchors; a single box; no boxes."""
    rng = np.random.default_rng(A + G)
    ctr = rng.random((A, 2), dtype=np.float32)
    wh = rng.random((A, 2), dtype=np.float32) * 0.3 + 0.03
    anchors = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=1).astype(np.float32)
    anchors[A // 2] = anchors[0]                       # duplicate anchors: exact IoU ties -> smallest index wins
    labels = np.full((B, G, 5), -1.0, np.float32)
    for b in range(B):
        for k in range(min(G, (3 * b + 1) if G < 33 else G)):
            labels[b, k, 0] = k % 3
            src = anchors[0] if k == 0 else anchors[rng.integers(A)]
            labels[b, k, 1:] = src + (0 if k < 2 else (rng.random(4, dtype=np.float32) - 0.5) * 0.08)
    if G > A:
        pytest.skip("more label rows than anchors: the reference starts re-picking discarded entries")
    o_cls, o_msk = _roi_check(anchors, labels, thr)
    if G > 0:
        assert int(o_msk[:, :, 0].sum()) >= B * min(G, A)   # every label row, padded ones included, claims an anchor


@pytest.mark.gpu
def test_detection_loss_matches_reference_expression():
    """SODa._loss (roi kernel + cross entropy / L1 as models/soda.py:259-281) against the same expression evaluated
    on the oracle's labelling; gradients flow to the predictions."""
    from oracle import roi_restated as rr
    from snn_for_object_detection_b200 import TinyYolo

    torch.manual_seed(2)
    g = golden("roi.npz")
    anchors = torch.from_numpy(g["anchors"]).to(DEV)
    labels = torch.from_numpy(g["labels"]).to(DEV)
    A = anchors.shape[0]
    model = TinyYolo(num_classes=2, iou_threshold=float(g["iou_threshold"]))
    cls_preds = torch.randn(3, A, 3, device=DEV, requires_grad=True)
    bbox_preds = torch.randn(3, A, 4, device=DEV, requires_grad=True)
    loss = model._loss((anchors, cls_preds, bbox_preds), labels)
    loss.backward()
    o_off, o_msk, o_cls = (torch.from_numpy(x) for x in rr.roi(g["anchors"], g["labels"], float(g["iou_threshold"])))
    ce = F.cross_entropy(cls_preds.detach().cpu().reshape(-1, 3), o_cls.reshape(-1), reduction="none")
    m = o_cls.reshape(-1) > 0
    want = ce[m].mean() * 0.04 + ce[~m].mean() * 0.96 + (bbox_preds.detach().cpu() * o_msk - o_off * o_msk).abs().mean()
    assert abs(float(loss.detach()) - float(want)) < 1e-5 * max(1.0, abs(float(want)))
    assert cls_preds.grad is not None and float(cls_preds.grad.abs().sum()) > 0 and float(bbox_preds.grad.abs().sum()) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("T,B,C,H,W,affine", [(4, 3, 32, 17, 23, True), (8, 5, 64, 30, 38, True), (1, 2, 8, 24, 30, True),
                                              (3, 2, 256, 8, 10, False), (2, 1, 128, 5, 7, True)])
def test_bn_train_per_time_step_matches_torch(T, B, C, H, W, affine):
    """soda_bn_train_fwd / _bwd against T calls of F.batch_norm(training=True) (what the reference's loop does,
    soda.py:142 + layer_gen.py:211): outputs, running statistics after T sequential updates, and all gradients.
    Floating-point kernel: torch fp32 on the same GPU is the reference, tolerance 2e-5 of the tensor scale."""
    ops = _pkg().ops
    torch.manual_seed(T * 100 + C)
    x = (torch.randn(T * B, C, H, W, device=DEV) * 1.7 + 0.3).contiguous(memory_format=torch.channels_last)
    gamma = (torch.rand(C, device=DEV) + 0.5) if affine else None
    beta = torch.randn(C, device=DEV) if affine else None
    gy = torch.randn(T * B, C, H, W, device=DEV).contiguous(memory_format=torch.channels_last)
    rm0, rv0 = torch.randn(C, device=DEV) * 0.1, torch.rand(C, device=DEV) + 0.5
    # reference: per-time-step torch
    xr = x.clone().requires_grad_(True)
    gr = gamma.clone().requires_grad_(True) if affine else None
    br = beta.clone().requires_grad_(True) if affine else None
    rm, rv = rm0.clone(), rv0.clone()
    yr = torch.cat([F.batch_norm(xr[t * B:(t + 1) * B], rm, rv, gr, br, True, 0.1, 1e-5) for t in range(T)])
    yr.backward(gy)
    # ours
    rm2, rv2 = rm0.clone(), rv0.clone()
    y, mean, invstd, xk = ops.bn_train_fwd(x, T, gamma, beta, 1e-5, 0.1, rm2, rv2)
    gx, gg, gb = ops.bn_train_bwd(gy, xk, T, gamma, mean, invstd, want_affine=affine)

    def close(a, b, tol=2e-5):
        assert float((a - b).abs().max()) <= tol * max(1.0, float(b.abs().max())), float((a - b).abs().max())

    close(y, yr.detach())
    close(rm2, rm)
    close(rv2, rv)
    close(gx, xr.grad, 5e-5)
    if affine:
        close(gg, gr.grad, 5e-5)
        close(gb, br.grad, 5e-5)
    # deterministic
    y2, *_ = ops.bn_train_fwd(x, T, gamma, beta, 1e-5, 0.1, rm0.clone(), rv0.clone())
    assert torch.equal(y, y2)


@pytest.mark.gpu
def test_bn_train_module_route_and_fallback(monkeypatch):
    """SodaBatchNorm2d.forward_seq in train mode takes the kernel for C = 4 * 2^k and ATen otherwise; both give the
    per-time-step result, update num_batches_tracked by T and let gradients reach the input."""
    from snn_for_object_detection_b200.modules.common import SodaBatchNorm2d

    for C in (32, 24):
        torch.manual_seed(C)
        bn = SodaBatchNorm2d(C).to(DEV).train()
        ref = torch.nn.BatchNorm2d(C).to(DEV).train()
        T, B = 3, 2
        x = torch.randn(T * B, C, 9, 11, device=DEV, requires_grad=True)
        y = bn.forward_seq(x, T)
        yr = torch.cat([ref(x.detach()[t * B:(t + 1) * B]) for t in range(T)])
        assert float((y.detach() - yr).abs().max().item()) < 2e-5
        assert int(bn.num_batches_tracked) == T
        assert float((bn.running_var - ref.running_var).abs().max()) < 1e-5
        y.square().mean().backward()
        assert x.grad is not None and float(x.grad.abs().sum()) > 0


@pytest.mark.gpu
def test_sli_scan_matches_reference_golden_and_carries_state():
    """soda_sli_fwd against the reference's own sli_feed_forward_step (tests/golden/sli.npz): floating-point kernel
    (expf vs ATen's exp), 2e-6 of the membrane scale; a scan split in two with carried state equals one scan."""
    ops = _pkg().ops
    from snn_for_object_detection_b200.modules.cells import SLICell

    g = golden("sli.npz")
    x = torch.from_numpy(g["x"])
    T, B = x.shape[:2]
    xd = x.reshape(T * B, *x.shape[2:]).to(DEV)
    cell = SLICell()
    with torch.no_grad():
        out, st = cell.forward_seq(xd, T)
    want = torch.from_numpy(g["out"]).reshape(T * B, *x.shape[2:])
    scale = float(want.abs().max())
    assert float((out.cpu() - want).abs().max()) < 2e-6 * scale
    assert float((st.v.cpu() - torch.from_numpy(g["v"])).abs().max()) < 2e-6 * scale
    assert float((st.i.cpu() - torch.from_numpy(g["i"])).abs().max()) < 2e-6 * float(np.abs(g["i"]).max())
    with torch.no_grad():
        o1, s1 = cell.forward_seq(xd[: 5 * B], 5)
        o2, s2 = cell.forward_seq(xd[5 * B:], T - 5, s1)
    assert torch.equal(torch.cat([o1, o2]), out) and torch.equal(s2.v, st.v) and torch.equal(s2.i, st.i)
    # training route: the reference's expression in torch ops, same values, gradients flow
    xg = xd.clone().requires_grad_(True)
    og, _ = cell.forward_seq(xg, T)
    assert float((og.detach() - out).abs().max()) < 2e-6 * scale
    og.sum().backward()
    assert float(xg.grad.abs().sum()) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("tag,sigma", [("s0", 0.0), ("s07", 0.7)])
def test_synapse_scan_matches_reference_golden_bit_exact(tag, sigma):
    """soda_synapse_fwd against the reference's own synapse_feed_forward_step (tests/golden/synapse.npz): only
    round-to-nearest adds / multiplies and a compare, so bit-exact; carried state; the differentiable torch route."""
    from snn_for_object_detection_b200.modules.cells import SynapseCell, SynapseParameters

    g = golden("synapse.npz")
    x = torch.from_numpy(g["x"])
    T, B = x.shape[:2]
    xd = x.reshape(T * B, *x.shape[2:]).to(DEV)
    cell = SynapseCell(SynapseParameters(sigma_inhibition=sigma))
    with torch.no_grad():
        out, st = cell.forward_seq(xd, T)
        o1, s1 = cell.forward_seq(xd[: 4 * B], 4)
        o2, s2 = cell.forward_seq(xd[4 * B:], T - 4, s1)
    assert torch.equal(out.cpu().reshape(x.shape), torch.from_numpy(g["out_" + tag]))
    assert torch.equal(st.p.cpu(), torch.from_numpy(g["p_" + tag]))
    assert torch.equal(torch.cat([o1, o2]), out) and torch.equal(s2.p, st.p)
    xg = xd.clone().requires_grad_(True)
    og, _ = cell.forward_seq(xg, T)
    assert float((og.detach() - out).abs().max()) < 1e-6
    og.sum().backward()
    assert float(xg.grad.abs().sum()) > 0
    with pytest.raises(ValueError):
        SynapseCell(SynapseParameters(sigma_inhibition=0.3))


@pytest.mark.gpu
def test_st_sample_stream_batches_to_device_frames():
    """events.STSampleStream.next_batch: two recordings, windows located as the reference's parse_data does, raw records
    staged through pinned memory and binned on the GPU: the golden recording's frames equal the reference's own
    features bit for bit, the second sample equals the oracle; exhausted recordings end the stream."""
    from snn_for_object_detection_b200.events import STSampleStream

    g = golden("binning_st.npz")
    T, H, W, step_us = int(g["T"]), int(g["H"]), int(g["W"]), int(g["step_us"])
    rng = np.random.default_rng(9)
    rec2 = ob.make_records(rng, 60000, 0, 12 * step_us, H, W)
    st = STSampleStream([ob.dat_bytes(g["records"], H, W), ob.dat_bytes(rec2, H, W)], [[6], [5, 10]], num_steps=T,
                        time_shift=1, time_step_ms=step_us // 1000, height=H, width=W, events_threshold=1000)
    X, steps, files = st.next_batch(4)
    assert X.shape == (T, 3, 2, H, W) and X.is_cuda and steps == [6, 5, 10] and files == [0, 1, 1]
    want = np.unpackbits(g["features_bits"])[: T * 2 * H * W].reshape(T, 2, H, W)
    assert np.array_equal(X[:, 0].cpu().numpy(), want)
    for b, s in ((1, 5), (2, 10)):
        t0 = s * step_us - step_us * (T - 1)
        keep = (rec2[:, 0] >= t0) & (rec2[:, 0] < s * step_us + step_us)
        assert np.array_equal(X[:, b].cpu().numpy(), ob.bin_st(rec2[keep], t0, step_us, T, H, W))
    assert st.next_batch(4) is None and not any(st.alive)


@pytest.mark.gpu
def test_mt_sample_stream_consecutive_windows_match_reference_golden():
    """events.MTSampleStream: two consecutive windows of one recording == the reference's own
    MTPropheseeDataset.parse_data called twice (binning_mt.npz), bit for bit; then the recording restarts."""
    from snn_for_object_detection_b200.events import MTSampleStream

    g = golden("binning_mt.npz")
    T, H, W, step_us = int(g["T"]), int(g["H"]), int(g["W"]), int(g["step_us"])
    ref = _unpack(g["features_bits"], (2, T, 2, H, W))
    st = MTSampleStream([ob.dat_bytes(g["records"], H, W)], num_steps=T, time_step_ms=step_us // 1000, height=H, width=W)
    X, starts, files = st.next_batch(2)
    assert X.shape == (T, 2, 2, H, W) and starts == [0, T] and files == [0, 0]
    assert np.array_equal(X[:, 0].cpu().numpy(), ref[0]) and np.array_equal(X[:, 1].cpu().numpy(), ref[1])
    X2, starts2, _ = st.next_batch(1)
    if st.readers[0].current_time == T * step_us:          # the recording ended after two windows: restarted
        assert starts2 == [0] and np.array_equal(X2[:, 0].cpu().numpy(), ref[0])


@pytest.mark.gpu
def test_mt_sample_stream_visits_every_recording_and_reuses_staging_buffers_safely():
    """More recordings than the batch size: the round-robin cursor survives the call (the reference's MT
    samples_generator visits every loaded file, datasets.py:305-309); and many batches in flight without a host
    sync (the pinned staging buffers are only rewritten after the copy that read them has finished) all come out
    equal to the oracle's binning."""
    from snn_for_object_detection_b200.events import MTSampleStream

    T, H, W, step_us = 3, 16, 24, 16000
    rng = np.random.default_rng(9)
    recs = [ob.make_records(rng, 6000, 0, 4 * T * step_us, H, W) for _ in range(3)]
    st = MTSampleStream([ob.dat_bytes(r, H, W) for r in recs], num_steps=T, time_step_ms=step_us // 1000, height=H, width=W)
    seen, outs = [], []
    for _ in range(6):                      # 12 windows, no synchronisation in between
        X, starts, files = st.next_batch(2)
        seen += files
        outs.append((X, starts, files))
    assert seen == [0, 1, 2, 0, 1, 2, 0, 1, 2, 0, 1, 2]
    for X, starts, files in outs:
        for b, (s0, k) in enumerate(zip(starts, files)):
            rec = recs[k]
            lo, hi = np.searchsorted(rec[:, 0], [s0 * step_us, (s0 + T) * step_us])
            ref = ob.bin_batch([rec[lo:hi]], [s0 * step_us], step_us, T, H, W, clip_x=False)
            assert np.array_equal(X[:, b].cpu().numpy(), ref[:, 0])


@pytest.mark.gpu
def test_folded_batchnorm_cache_follows_the_running_statistics():
    """eval forward, then a train-mode forward with frozen weights (running statistics move through the CUDA kernel's
    raw pointers), then an eval forward: the cached Conv+BN fold must be rebuilt (round-1 advice)."""
    pkg = _pkg()
    from snn_for_object_detection_b200 import BlockGen
    from snn_for_object_detection_b200.modules import LIF, Conv, Norm

    torch.manual_seed(2)
    block = BlockGen(2, [Conv(32, 3), Norm(), LIF()]).to(DEV)
    block.net[0][1].weight.data.fill_(6.0)       # with the reference's raw initialisation the layer is silent
    block.net[0][1].running_mean.fill_(-0.3)
    X = (torch.rand(8, 2, 16, 24, device=DEV) < 0.3).float()
    block.eval()
    with torch.no_grad():
        y0, _ = block.forward_seq(X, 4)
    block.train()
    for p in block.parameters():
        p.requires_grad_(False)
    with torch.no_grad():
        block.forward_seq(X * 3.0, 4)        # moves running_mean / running_var only
    block.eval()
    with torch.no_grad():
        y1, _ = block.forward_seq(X, 4)
    conv, norm = block.net[0][0], block.net[0][1]
    conv._pack_cache.clear()
    with torch.no_grad():
        y2, _ = block.forward_seq(X, 4)      # fresh pack: the truth for the current statistics
    assert torch.equal(y1, y2)
    assert float(norm.running_var.mean()) != 1.0 and 0.01 < float(y0.mean()) < 0.9 and not torch.equal(y0, y1)


@pytest.mark.gpu
def test_sync_batchnorm_sums_exchange_equals_full_batch():
    """SyncBatchNorm route (config.yaml:76) on one GPU: the batch is split between two emulated ranks, their
    [T, 2, C] sums are added (what the all-reduce does) and each rank normalises its own frames with the global sums:
    outputs, running statistics and input gradients equal the unsplit per-time-step BatchNorm; the ranks' dgamma /
    dbeta add up to the full ones."""
    ops = _pkg().ops
    torch.manual_seed(21)
    T, B, C, H, W = 4, 6, 32, 9, 13
    x = (torch.randn(T * B, C, H, W, device=DEV) * 1.3 + 0.2).contiguous(memory_format=torch.channels_last)
    gy = torch.randn_like(x)
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV)
    rm0, rv0 = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    rmF, rvF = rm0.clone(), rv0.clone()
    yF, meanF, invF, xF = ops.bn_train_fwd(x, T, gamma, beta, 1e-5, 0.1, rmF, rvF)
    gxF, ggF, gbF = ops.bn_train_bwd(gy, xF, T, gamma, meanF, invF)
    idx = [torch.cat([torch.arange(t * B + r * 3, t * B + r * 3 + 3) for t in range(T)]).to(DEV) for r in range(2)]
    xs = [x[i].contiguous(memory_format=torch.channels_last) for i in idx]
    gs = [gy[i].contiguous(memory_format=torch.channels_last) for i in idx]
    total = ops.bn_train_stats(xs[0], T) + ops.bn_train_stats(xs[1], T)

    def fake_all_reduce_fwd(s):
        s.copy_(total)
        return 2

    outs = [ops.bn_train_fwd_sync(xs[r], T, gamma, beta, 1e-5, 0.1, rm0.clone(), rv0.clone(), fake_all_reduce_fwd)
            for r in range(2)]
    local_b = [ops.bn_train_stats(xs[r], T, gs[r], outs[r][1], outs[r][2]) for r in range(2)]
    total_b = local_b[0] + local_b[1]

    def fake_all_reduce_bwd(s):
        s.copy_(total_b)
        return 2

    gg_sum, gb_sum = 0, 0
    for r in range(2):
        y, mean, invstd, xk, world = outs[r]
        assert world == 2
        assert float((y - yF[idx[r]]).abs().max()) < 2e-5 * float(yF.abs().max())
        assert float((mean - meanF).abs().max()) < 1e-6 and float((invstd - invF).abs().max()) < 1e-5
        gx, gg, gb = ops.bn_train_bwd_sync(gs[r], xk, T, gamma, mean, invstd, fake_all_reduce_bwd)
        assert float((gx - gxF[idx[r]]).abs().max()) < 5e-5 * max(1.0, float(gxF.abs().max()))
        gg_sum, gb_sum = gg_sum + gg, gb_sum + gb
    assert float((gg_sum - ggF).abs().max()) < 5e-5 * float(ggF.abs().max())
    assert float((gb_sum - gbF).abs().max()) < 5e-5 * float(gbF.abs().max())
    rm1, rv1 = rm0.clone(), rv0.clone()
    ops.bn_train_fwd_sync(xs[0], T, gamma, beta, 1e-5, 0.1, rm1, rv1, fake_all_reduce_fwd)
    assert float((rm1 - rmF).abs().max()) < 1e-6 and float((rv1 - rvF).abs().max()) < 1e-5


@pytest.mark.gpu
def test_stateless_layergens_on_cuda_kernels():
    """LayerGens ReLU / SiLU / Tanh / Up (reference layer_gen.py:176-194, :255-284) on the elementwise kernels against
    torch on the same GPU: ReLU and nearest upsampling bit-exact, SiLU / Tanh to 2e-6 (expf / tanhf vs ATen);
    channels_last and plain layouts; gradients take the torch route."""
    from snn_for_object_detection_b200.modules.layer_gen import ReLU, SiLU, Tanh, Up

    torch.manual_seed(8)
    for x in (torch.randn(3, 20, 9, 11, device=DEV) * 3, (torch.randn(2, 8, 6, 5, device=DEV) * 3).contiguous(memory_format=torch.channels_last)):
        with torch.no_grad():
            assert torch.equal(ReLU().get(20)[0](x), torch.relu(x))
            assert float((SiLU().get(20)[0](x) - F.silu(x)).abs().max()) < 2e-6 * 9
            assert float((Tanh().get(20)[0](x) - torch.tanh(x)).abs().max()) < 2e-6
            for s in (2, 3):
                up = Up(scale=s).get(20)[0]
                y = up(x)
                assert y.shape == (x.shape[0], x.shape[1], x.shape[2] * s, x.shape[3] * s)
                assert torch.equal(y, F.interpolate(x, scale_factor=s, mode="nearest"))
            yb = Up(scale=2, mode="bilinear").get(20)[0](x)      # other modes: torch
            assert torch.equal(yb, F.interpolate(x, scale_factor=2.0, mode="bilinear"))
    xg = torch.randn(2, 4, 5, 5, device=DEV, requires_grad=True)
    SiLU().get(4)[0](xg).sum().backward()
    assert float(xg.grad.abs().sum()) > 0
