"""GPU parity at network level: generator nets and TinyYolo, layer-major CUDA path vs the
time-major CPU oracle (which is pinned to the reference's own Python by tests/golden/).

Tolerances (BASELINE.json north_star): fp32 membrane potentials within 1e-4 relative, spike-map
mismatch <= 1e-4 of neurons, detections identical.
"""
import pytest
import torch
import torch.nn.functional as F

from helpers import golden, calibrated_tiny_yolo, tiny_yolo_from_golden, vgg_block_from_golden
from oracle import net_restated as onr
from snn_for_object_detection_b200 import generator

pytestmark = pytest.mark.gpu
DEV = "cuda"
SPIKE_TOL = 1e-4
V_RTOL = 1e-4


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def test_vgg_net_matches_golden_reference_output():
    block, g = vgg_block_from_golden()
    X = torch.from_numpy(g["X"]).float()
    T, B = X.shape[:2]
    Yref = torch.from_numpy(g["Y"])
    block = block.to(DEV)
    with torch.no_grad():
        Y, state = block.forward_seq(X.to(DEV).reshape(T * B, *X.shape[2:]), T)
    Y = Y.reshape(T, B, *Y.shape[1:]).cpu()
    assert float(Yref.mean()) > 0.05
    assert float((Y != Yref).float().mean()) <= SPIKE_TOL
    last = [s for s in state[0] if s is not None][-1]
    vref = torch.from_numpy(g["v_final"])
    ok = (last.v.cpu() - vref).abs() <= V_RTOL * vref.abs().max()
    assert float((~ok).float().mean()) <= SPIKE_TOL  # a flipped spike resets v: same budget as spikes


def test_vgg_time_major_equals_layer_major():
    """The per-step protocol forward(X, state) and forward_seq give identical results."""
    block, g = vgg_block_from_golden()
    X = torch.from_numpy(g["X"]).float().to(DEV)
    T, B = X.shape[:2]
    block = block.to(DEV)
    with torch.no_grad():
        Yseq, sseq = block.forward_seq(X.reshape(T * B, *X.shape[2:]), T)
        ys, state = [], None
        for ts in X:
            y, state = block(ts, state)
            ys.append(y)
    assert torch.equal(torch.cat(ys), Yseq)
    a = [s for s in state[0] if s is not None][-1]
    b = [s for s in sseq[0] if s is not None][-1]
    assert torch.equal(a.v, b.v) and torch.equal(a.i, b.i)
    # chunked streaming: first 2 steps, then the rest with the carried state
    with torch.no_grad():
        y1, s1 = block.forward_seq(X[:2].reshape(2 * B, *X.shape[2:]), 2)
        y2, s2 = block.forward_seq(X[2:].reshape((T - 2) * B, *X.shape[2:]), T - 2, s1)
    assert torch.equal(torch.cat([y1, y2]), Yseq)


def test_tiny_yolo_matches_golden_reference_output(monkeypatch):
    """End-to-end golden outputs (oracle/make_golden.py, produced by the reference's own model code).
    Run on the FP32-pipe convolution: the calibrated random-weight net is chaotic (one borderline spike
    flip in an early layer grows to percent-level differences 15 LIF layers later, DESIGN.md "Parity"),
    so an end-to-end equality only holds for an accumulation that happens not to flip any neuron on this
    seed.  The tensor-core path is held to the per-block gate and the first-divergence test below."""
    monkeypatch.setenv("SODA_CONV_KIND", "fp32")
    model, X, g, digest_ok = tiny_yolo_from_golden()
    assert digest_ok
    cls_ref, bbox_ref = torch.from_numpy(g["cls"]), torch.from_numpy(g["bbox"])
    model = model.to(DEV)
    with torch.no_grad():
        anchors, cls, bbox = model(X.to(DEV))
    assert torch.equal(anchors.cpu(), torch.from_numpy(g["anchors"]))
    # head outputs are conv(tanh(LI membrane)): membrane tolerance carried through a 1x1 conv
    assert _rel(cls.cpu(), cls_ref) < 5e-4, _rel(cls.cpu(), cls_ref)
    assert _rel(bbox.cpu(), bbox_ref) < 5e-4, _rel(bbox.cpu(), bbox_ref)


def _conv_norm_cell_triples(model):
    from snn_for_object_detection_b200 import BlockGen
    from snn_for_object_detection_b200.modules.common import SodaBatchNorm2d, SodaConv2d, StateStorage

    out = []
    for name, blk in model.named_modules():
        if not isinstance(blk, BlockGen):
            continue
        for bi, branch in enumerate(blk.net):
            for i in range(len(branch) - 2):
                if (isinstance(branch[i], SodaConv2d) and isinstance(branch[i + 1], SodaBatchNorm2d)
                        and isinstance(branch[i + 2], StateStorage)):
                    out.append((f"{name}.net.{bi}.{i}", branch[i], branch[i + 1], branch[i + 2]))
    return out


def _storage_model(seed=5, T=6, B=2, H=64, W=96):
    from snn_for_object_detection_b200 import TinyYolo

    torch.manual_seed(seed)
    model = TinyYolo(num_classes=2, state_storage=True)
    X = (torch.rand(T, B, 2, H, W) < 0.08).float()
    model.eval()
    onr.calibrate(model, X, gain=5.0)
    model.eval()
    return model, X


def test_tiny_yolo_every_conv_norm_neuron_block_teacher_forced():
    """THE parity gate.  Every Conv -> Norm -> LIF/LI block of TinyYolo (19 LIF + 3 LI) is run on the
    GPU over all T steps on the ORACLE's input to that block and compared with the oracle's output:
    spike mismatch <= 1e-4 of neurons, membranes of never-flipped neurons within 1e-4 relative.
    (Chained end to end the net amplifies a single borderline flip chaotically - see the next test
    and DESIGN.md - so per-block comparison on identical inputs is the meaningful statement.)"""
    model, X = _storage_model()
    T, B = X.shape[:2]
    taps = {}
    onr.soda_forward(model, X, taps)
    triples = _conv_norm_cell_triples(model)
    assert len(triples) == 22
    ref_in = {n: torch.cat(taps[("in", id(conv))]) for n, conv, _, _ in triples}
    ref_out = {n: taps[id(cell)] for n, _, _, cell in triples}
    model = model.to(DEV)
    worst_spike = worst_v = 0.0
    for n, conv, norm, cell in triples:
        outs = torch.stack([o for o, _ in ref_out[n]])
        v_ref = torch.stack([s.v for _, s in ref_out[n]])
        x = ref_in[n].to(DEV)
        if x.shape[1] != 2:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.no_grad():
            y = generator.conv_forward(conv, norm, x)     # the planner's route: packed-activation tcgen05 kernel
            cell.forward_seq(y, T, None)
        got, v = cell.get_spikes().cpu(), cell.get_state().v.cpu()
        assert got.shape == outs.shape
        if type(cell.module).__name__ == "LIFCell":
            assert float(outs.mean()) > 0.02, f"{n}: dead layer makes the parity check vacuous"
            flips = got != outs
            mism = float(flips.float().mean())
            worst_spike = max(worst_spike, mism)
            assert mism <= SPIKE_TOL, (n, mism)
            never = ~flips.any(dim=0, keepdim=True).expand_as(flips)
            err = float((v - v_ref)[never].abs().max() / v_ref.abs().max())
        else:
            err = _rel(v, v_ref)
        worst_v = max(worst_v, err)
        assert err < V_RTOL, (n, err)
    print(f"teacher-forced: worst spike mismatch {worst_spike:.2e}; worst membrane rel err {worst_v:.2e}")


def test_tiny_yolo_end_to_end_fp32_pipe(monkeypatch):
    """End to end through all 19 LIF layers with the FP32-pipe convolution (SODA_CONV_KIND=fp32):
    on this seed no neuron sits within rounding distance of the threshold, so spikes and
    membranes agree everywhere."""
    from snn_for_object_detection_b200.modules.common import StateStorage

    monkeypatch.setenv("SODA_CONV_KIND", "fp32")
    model, X = _storage_model()
    taps = {}
    preds_ref, _ = onr.soda_forward(model, X, taps)
    cells = [(n, m) for n, m in model.named_modules() if isinstance(m, StateStorage)]
    assert len(cells) == 22
    ref = {n: taps[id(m)] for n, m in cells}
    model = model.to(DEV)
    with torch.no_grad():
        preds = model(X.to(DEV))
    for n, m in cells:
        outs = torch.stack([o for o, _ in ref[n]])
        v_ref = torch.stack([s.v for _, s in ref[n]])
        got, v = m.get_spikes().cpu(), m.get_state().v.cpu()
        if type(m.module).__name__ == "LIFCell":
            assert float((got != outs).float().mean()) <= SPIKE_TOL, n
            bad = (v - v_ref).abs() > V_RTOL * v_ref.abs().max()
            assert float(bad.float().mean()) <= SPIKE_TOL, n
        else:
            assert _rel(v, v_ref) < V_RTOL, n
    for a, b in zip(preds, preds_ref):
        assert _rel(a.cpu(), b) < 5e-4


def test_tiny_yolo_end_to_end_tensor_core_first_divergence():
    """End to end on the tcgen05 path (the default).  Walk the 19 LIF layers in execution order: every
    layer before the first one with a flipped spike must agree with the oracle exactly (spikes) and
    within 1e-4 relative (membranes); in the first layer that has flips, every flipped neuron must have
    been within 1e-4 * max|v| of the threshold on the non-spiking side - i.e. the divergence is a
    threshold decision inside the membrane tolerance, not an arithmetic error.  Later layers see
    different inputs and are not comparable (chaotic amplification)."""
    from snn_for_object_detection_b200.modules.common import StateStorage

    model, X = _storage_model()
    taps = {}
    onr.soda_forward(model, X, taps)
    cells = [(n, m) for n, m in model.named_modules()
             if isinstance(m, StateStorage) and type(m.module).__name__ == "LIFCell"]
    assert len(cells) == 19
    ref = {n: taps[id(m)] for n, m in cells}
    model = model.to(DEV)
    with torch.no_grad():
        model(X.to(DEV))
    clean_layers = 0
    for n, m in cells:
        outs = torch.stack([o for o, _ in ref[n]])
        v_ref = torch.stack([s.v for _, s in ref[n]])
        got, v = m.get_spikes().cpu(), m.get_state().v.cpu()
        flips = got != outs
        vmax = float(v_ref.abs().max())
        if not flips.any():
            assert float((v - v_ref).abs().max()) <= V_RTOL * vmax, n
            clean_layers += 1
            continue
        assert float(flips.float().mean()) <= SPIKE_TOL, n
        # earliest flipped step of each neuron: its input history was identical up to there
        first_t = flips.float().argmax(dim=0, keepdim=True)
        first = torch.zeros_like(flips).scatter_(0, first_t, True) & flips
        side = torch.where(outs == 0, v_ref, v)[first]   # post-step membrane of the side that did not spike
        assert float((side - 1.0).abs().max()) <= V_RTOL * vmax, (n, float((side - 1.0).abs().max()))
        break
    assert clean_layers >= 1
    print(f"first divergence after {clean_layers} bit-identical LIF layers")


def test_tiny_yolo_detections_identical(monkeypatch):
    from snn_for_object_detection_b200.utils import box

    monkeypatch.setenv("SODA_CONV_KIND", "fp32")  # see test_tiny_yolo_matches_golden_reference_output
    model, X, g, _ = tiny_yolo_from_golden()
    anchors, cls_ref, bbox_ref = (torch.from_numpy(g[k]) for k in ("anchors", "cls", "bbox"))
    model = model.to(DEV)
    with torch.no_grad():
        _, cls, bbox = model(X.to(DEV))
    det_ref = box.multibox_detection(F.softmax(cls_ref, dim=2), bbox_ref, anchors)
    det = box.multibox_detection(F.softmax(cls.cpu(), dim=2), bbox.cpu(), anchors)
    for b in range(det.shape[0]):
        kr, kg = det_ref[b][det_ref[b, :, 0] >= 0], det[b][det[b, :, 0] >= 0]
        assert kr.shape == kg.shape
        assert torch.equal(kr[:, 0], kg[:, 0])
        assert torch.allclose(kr[:, 1:], kg[:, 1:], atol=1e-3)


def test_streaming_predict_protocol():
    """SODa.predict(X[2,H,W], state): time-major, state in / state out (reference soda.py:202-233)."""
    model, X, g, _ = tiny_yolo_from_golden()
    model = model.to(DEV)
    state = None
    with torch.no_grad():
        for ts in X[:, 0].to(DEV):
            det, state = model.predict(ts, state)
        (anc, cls_seq, bbox_seq), _ = model.forward_seq(X[:, :1].to(DEV))
        # time-major protocol reproduces the layer-major result exactly
        st, preds = None, None
        for ts in X[:, :1].to(DEV):
            preds, st = model._forward_impl(ts, st)
    assert det.dim() == 2 and det.shape[1] == 6
    assert len(state) == 3
    assert torch.equal(preds[1], cls_seq) and torch.equal(preds[2], bbox_seq)


def test_full_size_gen1_forward_properties():
    """BASELINE config 2 size (GEN1 304x240, B=5, T=32): batch sharding leaves results unchanged
    (bit-exact: a sample's accumulation order does not depend on the batch), outputs finite, net alive."""
    model = calibrated_tiny_yolo(seed=0).to(DEV)
    T, B, H, W = 32, 5, 240, 304
    gen = torch.Generator().manual_seed(1)
    X = (torch.rand(T, B, 2, H, W, generator=gen) < 0.07).float().to(DEV)
    with torch.no_grad():
        anchors, cls, bbox = model(X)
        _, cls_a, bbox_a = model(X[:, :2])
        _, cls_b, bbox_b = model(X[:, 2:])
    assert anchors.shape == (13545, 4) and cls.shape == (B, 13545, 3) and bbox.shape == (B, 13545, 4)
    assert torch.isfinite(cls).all() and torch.isfinite(bbox).all()
    assert torch.equal(torch.cat([cls_a, cls_b]), cls) and torch.equal(torch.cat([bbox_a, bbox_b]), bbox)
    assert float(cls.std()) > 0
    # ---- the planner's routes at this size (the benchmarked configuration), by the launches it makes ----------
    from snn_for_object_detection_b200 import _lib, ops

    prof = {}
    ops.PROFILE = prof
    try:
        n0 = _lib.launch_count()
        with torch.no_grad():
            model(X)
        n_launch = _lib.launch_count() - n0
    finally:
        ops.PROFILE = None
    conv_tags, lif_tags = prof.get("conv2d_tags", []), prof.get("conv_lif_tags", [])
    # 48 convolutions of the net -> 38 launches: each C2f block's entry 1x1 is folded into the two stacked stem
    # convolutions that open its Dense block (5 blocks x 2 launches saved)
    assert len(conv_tags) + len(lif_tags) == 38, (len(conv_tags), len(lif_tags))
    # fused Conv -> BN -> LIF where a layer has >= 600 work items: the event-frame layer (as an im2col'd 1x1 layer:
    # 32 "channels") and the two 32 -> 32 bottlenecks of the first C2f block
    assert sorted(t.split(" exact")[0] for t in lif_tags) == sorted([
        "(160, 32, 120, 152)->64 k1 s1 kindP", "(160, 32, 120, 152)->32 k3 s1 kindP", "(160, 32, 120, 152)->32 k3 s1 kindP"]), lif_tags
    # every other convolution of backbone and neck runs on the packed-activation tensor-core kernel; only the six
    # box / class convolutions of the head (fp32 LI + tanh input, last time step only) take the fp32-input kernel
    assert sum("kindP" in t for t in conv_tags) == 29 and sum("kind1" in t and t.startswith("(5, 256") for t in conv_tags) == 6, conv_tags
    assert len(prof.get("lif_scan", [])) == 16 and len(prof.get("li_scan", [])) == 3
    assert n_launch <= 62, n_launch   # one step = 59 kernels of this library (+ 3 torch cat copies in the head)
    assert ops.overflow_count() == 0


def _flat_states(state, acc=None):
    acc = [] if acc is None else acc
    if state is None:
        return acc
    if isinstance(state, (list,)):
        for s in state:
            _flat_states(s, acc)
    elif hasattr(state, "v"):
        acc.append(state.v)
        acc.append(state.i)
    return acc


def test_fused_conv_norm_lif_planner_equals_unfused(monkeypatch):
    """The planner's fused route (Conv -> Norm -> LIF (+ Residual identity) in ONE kernel, membrane
    state on-chip over all T steps) gives bit-identical predictions and final states to the unfused
    route (conv over T*B images, stand-alone scan, add_n), and launches fewer kernels."""
    from snn_for_object_detection_b200 import _lib

    model = calibrated_tiny_yolo(seed=3).to(DEV)
    gen = torch.Generator().manual_seed(7)
    X = (torch.rand(5, 2, 2, 128, 160, generator=gen) < 0.07).float().to(DEV)
    with torch.no_grad():
        monkeypatch.setenv("SODA_FUSE_LIF_MIN_ITEMS", "1")
        monkeypatch.setenv("SODA_FUSE_LIF_MAX_COUT", "100000")
        model.forward_seq(X)  # weight packing out of the launch count
        n0 = _lib.launch_count()
        preds_f, state_f = model.forward_seq(X)
        n_fused = _lib.launch_count() - n0
        # streaming: split the sequence and carry the state through the fused kernel's v0 / i0
        _, s1 = model.forward_seq(X[:2])
        preds_s, state_s = model.forward_seq(X[2:], s1)
        monkeypatch.setenv("SODA_FUSE_LIF", "0")
        n0 = _lib.launch_count()
        preds_u, state_u = model.forward_seq(X)
        n_unfused = _lib.launch_count() - n0
    assert n_fused < n_unfused - 10, (n_fused, n_unfused)
    for a, b, c in zip(preds_f, preds_u, preds_s):
        assert torch.equal(a, b) and torch.equal(a, c)
    fa, fb, fc = _flat_states(state_f), _flat_states(state_u), _flat_states(state_s)
    assert len(fa) == len(fb) == len(fc) == 2 * 22
    for a, b, c in zip(fa, fb, fc):
        assert torch.equal(a, b) and torch.equal(a, c)
    assert float(preds_f[1].std()) > 0


def test_bptt_training_step_gradients_match_oracle(monkeypatch):
    """BPTT through a generated net in TRAIN mode (batch-statistics BatchNorm per time step, surrogate-gradient
    LIF, LI readout): our forward kernels + fused reverse scans (+ cuDNN dgrad/wgrad) against autograd through the
    CPU oracle (reference executor order, restated norse SuperSpike).  Reference: models/soda.py:146-158."""
    import copy

    from snn_for_object_detection_b200 import BlockGen
    from snn_for_object_detection_b200.modules import LI, LIF, Conv, Norm, Pass, Residual

    monkeypatch.setenv("SODA_CONV_KIND", "fp32")  # fp32-exact products: no borderline spike flips in this check
    torch.manual_seed(11)
    cfg = [Conv(32, 3, stride=2), Norm(), LIF(),
           Residual([[Conv(kernel_size=3), Norm(), LIF()], [Pass()]]),
           Conv(8, 1), Norm(), LI()]
    block = BlockGen(2, cfg)
    for m in block.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.fill_(3.0)
    T, B, H, W = 4, 3, 20, 24
    X = (torch.rand(T, B, 2, H, W) < 0.15).float()
    ref = copy.deepcopy(block).train()
    state, ys = None, []
    for ts in X:   # the reference's time loop, with autograd recording
        y, state = onr.block_forward(ref, ts, state)
        ys.append(y)
    Yr = torch.stack(ys)
    Wt = torch.randn_like(Yr)
    (Yr * Wt).sum().backward()

    dev = copy.deepcopy(block).to(DEV).train()
    Y, _ = dev.forward_seq(X.to(DEV).reshape(T * B, 2, H, W), T)
    Y = Y.reshape(T, B, *Y.shape[1:])
    assert float((Y.detach().cpu() - Yr.detach()).abs().max()) <= 1e-4 * float(Yr.detach().abs().max())
    (Y * Wt.to(DEV)).sum().backward()
    n_checked = 0
    for (n, pr), (_, pd) in zip(ref.named_parameters(), dev.named_parameters()):
        assert pd.grad is not None, n
        scale = float(pr.grad.abs().max())
        assert scale > 0, n
        err = float((pd.grad.cpu() - pr.grad).abs().max()) / scale
        assert err < 2e-3, (n, err)
        n_checked += 1
    assert n_checked == 6
    # running statistics were updated T times, per time step
    for (n, br), (_, bd) in zip(ref.named_buffers(), dev.named_buffers()):
        if br.dtype.is_floating_point:
            assert torch.allclose(bd.cpu(), br, rtol=1e-4, atol=1e-6), n
        else:
            assert int(bd) == int(br) == T, n


def test_vgg_planner_fuses_lif_and_pool():
    """Config 1 net (Conv/Norm/LIF/Pool blocks): the planner's LIF+Pool fused scan gives the golden output and
    launches fewer kernels than scan + pool."""
    from snn_for_object_detection_b200 import _lib

    block, g = vgg_block_from_golden()
    X = torch.from_numpy(g["X"]).float().to(DEV)
    T, B = X.shape[:2]
    block = block.to(DEV)
    with torch.no_grad():
        block.forward_seq(X.reshape(T * B, *X.shape[2:]), T)
        n0 = _lib.launch_count()
        Y, _ = block.forward_seq(X.reshape(T * B, *X.shape[2:]), T)
        n = _lib.launch_count() - n0
    Y = Y.reshape(T, B, *Y.shape[1:]).cpu()
    assert float((Y != torch.from_numpy(g["Y"])).float().mean()) <= SPIKE_TOL
    assert n <= 7, n   # conv+lif, pool, conv, lif+pool, [pooled counts -> tensor-core operand format], conv+lif


@pytest.mark.gpu
def test_planner_tags_spike_inputs_through_pools_residuals_and_concats():
    """Every Conv of a Conv/Norm/LIF/Pool chain, a Residual with [Pass] and a Dense concat sees an input tagged with
    its spike-count bound (1, 4 after a 2x2 sum pool, 2 after spike + spike), and the tagged run equals an
    untagged run to the block bars."""
    from snn_for_object_detection_b200 import BlockGen, ops
    from snn_for_object_detection_b200.modules.layer_gen import LIF, Conv, Dense, Norm, Pass, Pool, Residual

    torch.manual_seed(11)
    cb = lambda c, k=3: (Conv(c, kernel_size=k), Norm(), LIF())
    cfg = [*cb(32), Pool("S"), *cb(32), Residual([[*cb(32)], [Pass()]]), Dense([[*cb(32)], [Pass()]]), *cb(64, 1)]
    block = BlockGen(2, [cfg]).to(DEV).eval()
    for m in block.modules():
        if isinstance(m, torch.nn.Conv2d):
            torch.nn.init.normal_(m.weight, std=6.0 / (m.in_channels * m.kernel_size[0] ** 2) ** 0.5)
    T, B = 8, 2
    X = (torch.rand(T * B, 2, 32, 48, device=DEV) < 0.3).float()
    prof = {}
    ops.PROFILE = prof
    try:
        with torch.no_grad():
            Y, _ = block.forward_seq(X, T)
    finally:
        ops.PROFILE = None
    tags = prof.get("conv2d_tags", []) + prof.get("conv_lif_tags", [])
    bounds = sorted(t.split("exact")[1] for t in tags)
    # first conv: event frame (untagged); then pooled spikes (4), spikes (1), spike + spike (2), concat (2)
    assert bounds == ["1", "2", "2", "4", "None"], tags
    assert ops.exact_bound(Y) == 1 and 0.0 < float(Y.mean()) < 1.0
    real_mark = ops.mark_exact
    ops.mark_exact = lambda t, bound=1: t
    try:
        with torch.no_grad():
            Y0, _ = block.forward_seq(X, T)
    finally:
        ops.mark_exact = real_mark
    assert float((Y != Y0).float().mean()) <= 2e-3   # end to end through 6 spiking layers (flips propagate)


@pytest.mark.gpu
def test_generated_net_with_sli_layers_matches_time_major_oracle():
    """A ListGen with the reference's SLI LayerGen (Conv, Norm, SLI, Conv, Norm, SLI with state_storage) is a drop-in:
    layer-major forward == the per-step protocol == the oracle's SLI restatement on the same conv outputs."""
    from oracle import sli_restated as sr
    from snn_for_object_detection_b200 import BlockGen
    from snn_for_object_detection_b200.modules.layer_gen import SLI, Conv, Norm

    torch.manual_seed(4)
    block = BlockGen(2, [Conv(32, kernel_size=3), Norm(), SLI(), Conv(32, kernel_size=3), Norm(), SLI(state_storage=True)])
    block = block.to(DEV).eval()
    T, B = 6, 2
    X = (torch.rand(T * B, 2, 20, 24, device=DEV) < 0.3).float()
    with torch.no_grad():
        Y, state = block.forward_seq(X, T)
        ys, st = [], None
        for t in range(T):                       # the reference's per-step protocol
            y, st = block(X[t * B:(t + 1) * B], st)
            ys.append(y)
    assert torch.equal(torch.cat(ys), Y)
    storage = block.net[0][5]
    assert storage.get_spikes().shape[0] == T and storage.get_state().v.shape[0] == T
    # oracle: SLI restatement applied to this net's own second-conv currents
    with torch.no_grad():
        c1 = block.net[0][0](X, norm=block.net[0][1])
        v1, _ = block.net[0][2].forward_seq(c1, T)
        c2 = block.net[0][3](v1, norm=block.net[0][4])
    want, _, _ = sr.sli_scan(c2.cpu().reshape(T, B, *c2.shape[1:]))
    assert float((Y.cpu().reshape(want.shape) - want).abs().max()) < 5e-6 * max(1.0, float(want.abs().max()))


@pytest.mark.gpu
def test_conv_lstm_matches_reference_golden_and_protocols():
    """LayerGen LSTM: our ConvLSTM (input part of the gate conv for all T at once, recurrent part per step with the
    input part as fused residual, gate kernel) against the reference's OWN ConvLSTM (tests/golden/conv_lstm.npz),
    same state_dict key; layer-major == per-step protocol; inside a generated block; training route in torch ops."""
    from snn_for_object_detection_b200 import BlockGen
    from snn_for_object_detection_b200.modules.conv_lstm import ConvLSTM
    from snn_for_object_detection_b200.modules.layer_gen import LSTM, Conv, Norm

    g = golden("conv_lstm.npz")
    x = torch.from_numpy(g["x"])
    T, B = x.shape[:2]
    m = ConvLSTM(8, 16)
    m.load_state_dict({"conv.weight": torch.from_numpy(g["weight"])}, strict=True)
    m = m.to(DEV).eval()
    xd = x.reshape(T * B, *x.shape[2:]).to(DEV)
    with torch.no_grad():
        out, (h, c) = m.forward_seq(xd, T)
        ys, st = [], None
        for t in range(T):
            y, st = m(xd[t * B:(t + 1) * B], st)
            ys.append(y)
    want = torch.from_numpy(g["out"]).reshape(T * B, 16, *x.shape[3:])
    assert float((out.cpu() - want).abs().max()) < 5e-6          # fp32 conv order + expf / tanhf vs ATen
    assert float((h.cpu() - torch.from_numpy(g["h"])).abs().max()) < 5e-6
    assert float((c.cpu() - torch.from_numpy(g["c"])).abs().max()) < 5e-6
    assert torch.equal(torch.cat(ys), out)
    og, _ = m.forward_seq(xd.clone().requires_grad_(True), T)      # training route
    assert float((og.detach() - out).abs().max()) < 5e-3           # torch's conv runs cuDNN TF32 here
    block = BlockGen(2, [Conv(32, kernel_size=3), Norm(), LSTM(32), Conv(16, kernel_size=1)]).to(DEV).eval()
    X = (torch.rand(4 * 2, 2, 12, 16, device=DEV) < 0.3).float()
    with torch.no_grad():
        Y, state = block.forward_seq(X, 4)
        ys, st = [], None
        for t in range(4):
            y, st = block(X[t * 2:(t + 1) * 2], st)
            ys.append(y)
    assert Y.shape == (8, 16, 12, 16) and torch.equal(torch.cat(ys), Y)


@pytest.mark.gpu
def test_generated_net_with_up_and_activations_is_a_drop_in():
    """ListGen mixing Conv/Norm/LIF with Up, ReLU and SiLU: layer-major == per-step protocol; the upsampled spike map
    keeps its spike-count tag, so the following Conv takes the bf16-only operand scheme."""
    from snn_for_object_detection_b200 import BlockGen, ops
    from snn_for_object_detection_b200.modules.layer_gen import LIF, Conv, Norm, ReLU, SiLU, Up

    torch.manual_seed(12)
    block = BlockGen(2, [Conv(32, kernel_size=3, stride=2), Norm(), LIF(), Up(2), Conv(32, kernel_size=3), ReLU(),
                         Conv(16, kernel_size=1), SiLU()]).to(DEV).eval()
    for m in block.modules():
        if isinstance(m, torch.nn.Conv2d):
            torch.nn.init.normal_(m.weight, std=6.0 / (m.in_channels * m.kernel_size[0] ** 2) ** 0.5)
    T, B = 6, 2
    X = (torch.rand(T * B, 2, 24, 32, device=DEV) < 0.3).float()
    prof = {}
    ops.PROFILE = prof
    try:
        with torch.no_grad():
            Y, _ = block.forward_seq(X, T)
    finally:
        ops.PROFILE = None
    assert Y.shape == (T * B, 16, 24, 32) and float(Y.abs().sum()) > 0
    tags = prof.get("conv2d_tags", []) + prof.get("conv_lif_tags", [])
    assert any("(12, 32, 24, 32)->32" in t and t.endswith("exact1") for t in tags), tags
    with torch.no_grad():
        ys, st = [], None
        for t in range(T):
            y, st = block(X[t * B:(t + 1) * B], st)
            ys.append(y)
    assert torch.equal(torch.cat(ys), Y)


@pytest.mark.gpu
def test_package_calibration_reaches_the_fused_layers_and_the_network_is_alive():
    """bench.py's synthetic-weight calibration (utils/calibrate.py) must also cover the layers the planner runs as
    fused Conv -> Norm -> LIF kernels (the event-frame layer is one): with raw initialisation they are silent
     and every output downstream is exactly zero."""
    import snn_for_object_detection_b200 as pkg
    from snn_for_object_detection_b200.utils.calibrate import calibrate

    torch.manual_seed(0)
    model = pkg.TinyYolo(num_classes=2).to(DEV).eval()
    X = (torch.rand(8, 2, 2, 96, 128, device=DEV) < 0.07).float()
    with torch.no_grad():
        _, cls0, _ = model(X)
        assert float((cls0 != 0).float().mean()) == 0.0      # the premise: reference init is silent
        calibrate(model, X, gain=5.0)
        stem_bn = next(m for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d))
        assert float(stem_bn.weight.mean()) == 5.0 and float(stem_bn.running_var.mean()) != 1.0
        _, cls, bbox = model(X)
    assert float((cls != 0).float().mean()) > 0.9 and float((bbox != 0).float().mean()) > 0.9
    assert torch.isfinite(cls).all() and torch.isfinite(bbox).all()
