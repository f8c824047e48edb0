"""Kernel-level parity of the packed-activation path (rows N1 / N2): soda_act_convert, soda_pconv_fwd,
soda_pconv_lif_fwd, soda_lif_fmt_fwd against fp64 / torch references on the same inputs.

The convolution is a floating-point kernel: the reference is torch's conv2d in float64 on the SAME (decoded)
operands, tolerance 3e-6 of the output scale (fp16 hi/lo operand split = 22 significant bits, fp32 accumulation;
measured 2e-7 .. 1.6e-6).  Format round trips and the LIF arithmetic are exact.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 3e-6   # tensor-core fp32 accumulation truncates: ~1e-6 at K = 288 .. 1152 (measured), operands carry 22 bits


def _pkg():
    import snn_for_object_detection_b200 as pkg

    return pkg


def _cl(x):
    return x.contiguous(memory_format=torch.channels_last)


def _ref_conv(x, w, scale, shift, stride, residual=None):
    y = F.conv2d(x.double(), (w.double() * scale.double().view(-1, 1, 1, 1)), None, stride, w.shape[2] // 2)
    y = y + shift.double().view(1, -1, 1, 1)
    if residual is not None:
        y = y + residual.double()
    return y


def test_format_round_trips_are_exact_where_they_must_be():
    ops = _pkg().ops
    torch.manual_seed(0)
    x = _cl(torch.randn(3, 64, 9, 11, device=DEV) * 3)
    p = ops.act_convert(x, ops.FMT_PAIR)
    assert p.dtype == torch.int32 and p.shape == x.shape and p.stride() == x.stride()
    back = ops.to_f32(p)
    # hi + lo carries 22 significant bits
    assert float((back - x).abs().max()) <= 2.0 ** -21 * float(x.abs().max())
    # idempotent: what PAIR can represent survives another round trip bit for bit
    assert torch.equal(ops.to_f32(ops.act_convert(back, ops.FMT_PAIR)), back)
    # exact small integers in all three formats
    z = _cl((torch.rand(3, 64, 9, 11, device=DEV) < 0.3).float() * torch.randint(1, 5, (3, 64, 9, 11), device=DEV))
    h = ops.act_convert(z, ops.FMT_HALF)
    assert h.dtype == torch.float16 and torch.equal(h.float(), z)
    assert torch.equal(ops.to_f32(ops.act_convert(h, ops.FMT_PAIR)), z)
    # into a channel slice of a wider PAIR buffer, neighbours untouched
    buf = ops.empty_act(3, 128, 9, 11, DEV, ops.FMT_PAIR)
    buf.zero_()
    ops.act_convert(x, ops.FMT_PAIR, out=buf[:, 32:96])
    full = ops.to_f32(buf)
    assert torch.equal(full[:, 32:96], back) and float(full[:, :32].abs().max()) == 0 and float(full[:, 96:].abs().max()) == 0
    assert ops.overflow_count() == 0


CASES = [
    # N, Cin, H, W, Cout, k, stride, in_fmt, out_fmt, residual fmt, note
    (4, 64, 24, 40, 64, 1, 1, "pair", "pair", None),        # flat 1x1
    (4, 64, 24, 40, 32, 1, 1, "half", "f32", None),         # spikes in, N = 32
    (3, 32, 33, 29, 32, 3, 1, "pair", "pair", "pair"),      # halo, odd map, residual, N = 32 (unit split)
    (3, 64, 30, 38, 64, 3, 1, "pair", "f32", None),         # weights resident at BN = 64
    (2, 128, 30, 38, 128, 3, 1, "pair", "pair", None),      # streamed weights, two tiles per item
    (2, 64, 31, 37, 128, 3, 2, "pair", "half", None),       # stride 2 (tap mode), odd map
    (2, 128, 15, 19, 256, 3, 2, "half", "pair", None),      # stride 2 on spikes, 15x19 -> 8x10
    (2, 320, 15, 19, 128, 1, 1, "pair", "pair", "f32"),     # wide 1x1
    (2, 256, 8, 10, 96, 1, 1, "pair", "f32", None),         # Cout = 96: one absent 32-channel group at BN = 128
    (5, 128, 16, 16, 64, 1, 1, "half", "half", "half"),     # all fp16
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c[1]}x{c[2]}x{c[3]}->{c[4]}k{c[5]}s{c[6]}{c[7]}>{c[8]}")
def test_pconv_against_fp64(case):
    ops = _pkg().ops
    N, Cin, H, W, Cout, k, stride, in_fmt, out_fmt, res_fmt = case
    F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF}
    torch.manual_seed(sum(v for v in case if isinstance(v, int)))
    if in_fmt == "half":
        x = _cl((torch.rand(N, Cin, H, W, device=DEV) < 0.2).float() * torch.randint(1, 4, (N, Cin, H, W), device=DEV))
    else:
        x = _cl(torch.randn(N, Cin, H, W, device=DEV) * 2 + (torch.rand(N, Cin, H, W, device=DEV) < 0.2).float())
    w = torch.randn(Cout, Cin, k, k, device=DEV) * (2.0 / (Cin * k * k)) ** 0.5
    scale = torch.rand(Cout, device=DEV) * 3 + 0.2
    shift = torch.randn(Cout, device=DEV)
    xp = ops.act_convert(x, F_[in_fmt])
    xd = ops.to_f32(xp)                       # the operand values the kernel actually sees
    Ho, Wo = ops.conv_out_hw(H, W, k, stride)
    res = resd = None
    if res_fmt is not None:
        r = _cl(torch.randn(N, Cout, Ho, Wo, device=DEV)) if res_fmt != "half" else \
            _cl((torch.rand(N, Cout, Ho, Wo, device=DEV) < 0.3).float())
        res = ops.act_convert(r, F_[res_fmt])
        resd = ops.to_f32(res)
    pc = ops.pack_pconv(w, stride, scale, shift, F_[in_fmt])
    # destination = a channel slice of a wider buffer, to exercise the strided store
    wide = ops.empty_act(N, Cout + 64, Ho, Wo, DEV, F_[out_fmt])
    wide.zero_()
    out = wide[:, 32:32 + Cout]
    y = ops.pconv(xp, pc, residual=res, out=out)
    ref = _ref_conv(xd, w, scale, shift, stride, resd)
    got = ops.to_f32(y).double()
    scale_ref = float(ref.abs().max())
    err = float((got - ref).abs().max()) / scale_ref
    if out_fmt == "f32":
        assert err < TOL, err
    elif out_fmt == "pair":
        assert err < TOL + 2.0 ** -21, err       # + the re-encoding of the output
    else:
        assert err < 1e-3, err                   # fp16 output of a real-valued result: 11 bits
    lo, hi = ops.to_f32(wide[:, :32]), ops.to_f32(wide[:, 32 + Cout:])
    assert float(lo.abs().max()) == 0 and float(hi.abs().max()) == 0
    # and against the UNSPLIT input: PAIR carries 22 bits of it
    if in_fmt == "pair" and out_fmt == "f32":
        ref0 = _ref_conv(x, w, scale, shift, stride, resd)
        assert float((got - ref0).abs().max()) / scale_ref < 2e-6
    print(f"pconv {case}: max err {err:.2e} of the output scale")
    assert ops.overflow_count() == 0


@pytest.mark.parametrize("shape", [(32, 32, 3, 1, "pair"), (64, 64, 3, 1, "pair"), (64, 128, 3, 2, "pair"),
                                   (128, 128, 1, 1, "half"), (2 * 32, 64, 3, 2, "half"),
                                   (64, 96, 3, 1, "pair")])   # 96 channels: the second 64-channel chunk has one absent group
@pytest.mark.parametrize("out_fmt", ["f32", "pair", "half"])
def test_pconv_lif_equals_pconv_then_scan(shape, out_fmt):
    """Fused Conv -> BN -> LIF (+ residual) over T steps with the state in registers == the same convolution written
    as fp32 followed by the stand-alone bit-exact LIF scan; also with the state carried across two calls."""
    ops = _pkg().ops
    Cin, Cout, k, stride, in_fmt = shape
    F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF}
    T, B, H, W = 5, 3, 21, 27
    torch.manual_seed(Cin + Cout + k)
    if in_fmt == "half":
        x = _cl((torch.rand(T * B, Cin, H, W, device=DEV) < 0.25).float())
    else:
        x = _cl(torch.randn(T * B, Cin, H, W, device=DEV))
    w = torch.randn(Cout, Cin, k, k, device=DEV) * (6.0 / (Cin * k * k)) ** 0.5
    scale, shift = torch.rand(Cout, device=DEV) + 0.5, torch.randn(Cout, device=DEV) * 0.3
    xp = ops.act_convert(x, F_[in_fmt])
    pc = ops.pack_pconv(w, stride, scale, shift, F_[in_fmt])
    Ho, Wo = ops.conv_out_hw(H, W, k, stride)
    use_res = stride == 1 and out_fmt != "half"
    res = ops.act_convert(_cl(torch.randn(T * B, Cout, Ho, Wo, device=DEV)), ops.FMT_PAIR) if use_res else None
    cur = ops.pconv(xp, pc)                                   # fp32 currents
    z_ref, v_ref, i_ref, _, _, _ = ops.lif_scan(cur, T)
    assert 0.01 < float(z_ref.mean()) < 0.9
    want = z_ref + (ops.to_f32(res) if use_res else 0)
    y, v, i = ops.pconv_lif(xp, pc, T, ops.LIF_DEFAULT, residual=res, out_fmt=F_[out_fmt])
    got = ops.to_f32(y)
    if out_fmt == "pair" and use_res:
        assert float((got - want).abs().max()) <= 2.0 ** -21 * float(want.abs().max())
    else:
        assert torch.equal(got, want)
    assert torch.equal(v, v_ref) and torch.equal(i, i_ref)
    # streaming: 2 steps, then the rest with the carried state
    n1 = 2 * B
    y1, v1, i1 = ops.pconv_lif(xp[:n1], pc, 2, ops.LIF_DEFAULT, residual=None if res is None else res[:n1],
                               out_fmt=F_[out_fmt])
    y2, v2, i2 = ops.pconv_lif(xp[n1:], pc, T - 2, ops.LIF_DEFAULT, v0=v1, i0=i1,
                               residual=None if res is None else res[n1:], out_fmt=F_[out_fmt])
    assert torch.equal(torch.cat([ops.to_f32(y1), ops.to_f32(y2)]), got)
    assert torch.equal(v2, v) and torch.equal(i2, i)


@pytest.mark.parametrize("hw", [(13, 19), (16, 8), (40, 23), (5, 3)])
def test_pconv_lif_two_tiles_per_item_with_an_odd_tile_count(hw):
    """The fused kernel at 32 output channels takes TWO 16 x 8 tiles per work item (one per epilogue warpgroup, step-
    interleaved).  Maps with an odd number of tiles (13 x 19: 3; 40 x 23: 9), exactly one tile and a map smaller than a
    tile: the unpaired last tile is computed, the absent partner neither stored nor carried into the state."""
    ops = _pkg().ops
    H, W = hw
    T, B, C = 4, 3, 32
    torch.manual_seed(H * 100 + W)
    x = _cl(torch.randn(T * B, C, H, W, device=DEV))
    w = torch.randn(C, C, 3, 3, device=DEV) * (6.0 / (C * 9)) ** 0.5
    scale, shift = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV) * 0.3
    xp = ops.act_convert(x, ops.FMT_PAIR)
    pc = ops.pack_pconv(w, 1, scale, shift, ops.FMT_PAIR)
    res = ops.act_convert(_cl(torch.randn(T * B, C, H, W, device=DEV)), ops.FMT_PAIR)
    cur = ops.pconv(xp, pc)
    z_ref, v_ref, i_ref, _, _, _ = ops.lif_scan(cur, T)
    assert 0.01 < float(z_ref.mean()) < 0.9
    wide = ops.empty_act(T * B, C + 64, H, W, DEV, ops.FMT_PAIR)
    wide.zero_()
    y, v, i = ops.pconv_lif(xp, pc, T, ops.LIF_DEFAULT, residual=res, out=wide[:, 32:64])
    want = z_ref + ops.to_f32(res)
    assert float((ops.to_f32(y) - want).abs().max()) <= 2.0 ** -21 * float(want.abs().max())
    assert torch.equal(v, v_ref) and torch.equal(i, i_ref)
    assert float(ops.to_f32(wide[:, :32]).abs().max()) == 0 and float(ops.to_f32(wide[:, 64:]).abs().max()) == 0


@pytest.mark.parametrize("res_fmt", [None, "f32", "pair", "half"])
@pytest.mark.parametrize("out_fmt", ["f32", "pair", "half"])
def test_lif_fmt_scan_equals_plain_scan(res_fmt, out_fmt):
    ops = _pkg().ops
    F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF}
    T, B, Cc, H, W = 6, 2, 64, 13, 17
    torch.manual_seed(3)
    x = _cl(torch.randn(T * B, Cc, H, W, device=DEV) * 3)
    z_ref, v_ref, i_ref, _, _, _ = ops.lif_scan(x, T)
    res = None
    if res_fmt is not None:
        r = _cl((torch.rand(T * B, Cc, H, W, device=DEV) < 0.4).float()) if "half" in (res_fmt, out_fmt) else \
            _cl(torch.randn(T * B, Cc, H, W, device=DEV))
        res = ops.act_convert(r, F_[res_fmt])
    wide = ops.empty_act(T * B, Cc + 32, H, W, DEV, F_[out_fmt])
    wide.zero_()
    y, v, i = ops.lif_scan_fmt(x, T, residual=res, out=wide[:, 32:])
    want = z_ref + (ops.to_f32(res) if res is not None else 0)
    got = ops.to_f32(y)
    if out_fmt == "pair":
        assert float((got - want).abs().max()) <= 2.0 ** -21 * max(1.0, float(want.abs().max()))
    else:
        assert torch.equal(got, want)
    assert torch.equal(v, v_ref) and torch.equal(i, i_ref)
    assert float(ops.to_f32(wide[:, :32]).abs().max()) == 0
    # carried state
    h = 2 * B
    _, v1, i1 = ops.lif_scan_fmt(x[:h], 2, out_fmt=F_[out_fmt])
    _, v2, i2 = ops.lif_scan_fmt(x[h:], T - 2, v0=v1, i0=i1, out_fmt=F_[out_fmt])
    assert torch.equal(v2, v_ref) and torch.equal(i2, i_ref)


@pytest.mark.parametrize("cin,cout,out_fmt", [(2, 8, "half"), (2, 8, "f32"), (8, 32, "pair"), (8, 32, "half"), (8, 16, "f32")])
def test_conv_lif_pool_equals_conv_then_scan_then_pool(cin, cout, out_fmt):
    """Conv(3x3) -> BN -> LIF -> 2x2 sum pool in one kernel (pooled counts in the consumer's format, full-resolution
    spikes never written) == FP32-pipe convolution, bit-exact LIF scan and pooling kernel run one after the other:
    pooled counts identical up to borderline spikes (different fp32 summation order; <= 1e-4 of neurons, measured 0),
    membrane state within 1e-5; also with the state carried across two calls."""
    pkg = _pkg()
    ops = pkg.ops
    F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF}
    T, B, H, W = 5, 2, 22, 36
    torch.manual_seed(cin * 100 + cout)
    if cin == 2:
        x = (torch.rand(T * B, cin, H, W, device=DEV) < 0.3).float()                      # NCHW event frames
        xin = x
    else:
        x = _cl((torch.rand(T * B, cin, H, W, device=DEV) < 0.3).float() * torch.randint(1, 5, (T * B, cin, H, W), device=DEV))
        xin = ops.act_convert(x, ops.FMT_HALF)
    w = torch.randn(cout, cin, 3, 3, device=DEV) * (30.0 / (cin * 9)) ** 0.5
    scale, shift = torch.rand(cout, device=DEV) + 0.5, torch.randn(cout, device=DEV) * 0.2 + 0.3
    pc = ops.pack_conv(w, 1, scale, shift, pkg._lib.SODA_CONV_FP32)
    assert ops.conv_lif_pool_ok(xin, pc)
    cur = ops.conv2d(x, pc)
    z, v_ref, i_ref, _, _, _ = ops.lif_scan(cur, T)
    want = ops.pool2d(z, 2, 2, "S")
    assert 0.02 < float(z.mean()) < 0.9
    y, v, i = ops.conv_lif_pool(xin, pc, T, ops.LIF_DEFAULT, out_fmt=F_[out_fmt])
    got = ops.to_f32(y)
    assert got.shape == want.shape
    assert float((got != want).float().mean()) <= 1e-4
    assert float((v - v_ref).abs().max()) <= 1e-5 * max(1.0, float(v_ref.abs().max())) or \
        float(((v - v_ref).abs() > 1e-5).float().mean()) <= 1e-4
    n1 = 2 * B
    y1, v1, i1 = ops.conv_lif_pool(xin[:n1], pc, 2, ops.LIF_DEFAULT, out_fmt=F_[out_fmt])
    y2, v2, i2 = ops.conv_lif_pool(xin[n1:], pc, T - 2, ops.LIF_DEFAULT, v0=v1, i0=i1, out_fmt=F_[out_fmt])
    assert torch.equal(torch.cat([ops.to_f32(y1), ops.to_f32(y2)]), got)
    assert torch.equal(v2, v) and torch.equal(i2, i)


def test_event_frame_layer_as_im2col_plus_1x1_matches_the_direct_kernel():
    """The detector's first layer (Conv(64, 3, stride 2) + BN + LIF on the binned frames, K = 18) as im2col of the
    frames + a 1x1 layer on the tensor-core kernel == the direct FP32-pipe kernel: spikes equal up to borderline
    neurons (<= 1e-4), membranes within 1e-5."""
    pkg = _pkg()
    ops = pkg.ops
    T, B, H, W = 6, 3, 48, 64
    torch.manual_seed(5)
    x = (torch.rand(T * B, 2, H, W, device=DEV) < 0.1).float()
    conv = pkg.modules.common.SodaConv2d(2, 64, 3, stride=2, padding=1, bias=False).to(DEV)
    torch.nn.init.normal_(conv.weight, std=1.2)
    norm = pkg.modules.common.SodaBatchNorm2d(64).to(DEV).eval()
    norm.bias = None
    norm.weight.data.uniform_(0.5, 2.0)
    norm.running_mean.uniform_(-0.2, 0.2)
    col = ops.im2col_frames(x, 2)
    ref_col = torch.nn.functional.unfold(x, 3, padding=1, stride=2)             # [N, C*9 (ci, ky, kx), L]
    Ho, Wo = ops.conv_out_hw(H, W, 3, 2)
    ref_col = ref_col.reshape(T * B, 2, 9, Ho, Wo).permute(0, 2, 1, 3, 4).reshape(T * B, 18, Ho, Wo)
    got_col = ops.to_f32(col)
    assert torch.equal(got_col[:, :18], ref_col) and float(got_col[:, 18:].abs().max()) == 0
    y, v, i = ops.pconv_lif(col, conv.packed_im2col(norm), T, ops.LIF_DEFAULT, out_fmt=ops.FMT_HALF)
    pc = conv.packed(norm, pkg._lib.SODA_CONV_FP32)
    y0, v0, i0 = ops.conv_lif(x, pc, T, ops.LIF_DEFAULT)
    assert 0.02 < float(y0.mean()) < 0.9
    assert float((ops.to_f32(y) != y0).float().mean()) <= 1e-4
    assert float(((v - v0).abs() > 1e-5 * max(1.0, float(v0.abs().max()))).float().mean()) <= 1e-4
