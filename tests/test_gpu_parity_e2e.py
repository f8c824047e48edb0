"""End-to-end parity of the path that is BENCHMARKED (the tensor-core convolution, default planner routes),
adjudicated against the comparison's own noise floor, and parity at BASELINE.json configuration sizes.

A threshold is a discontinuity and the calibrated random-weight TinyYolo amplifies one borderline spike flip
chaotically (DESIGN.md "Parity").  "Any two correct fp32 implementations diverge the same way" is therefore
MEASURED here: the oracle is run twice - stock fp32 convolutions, and convolutions accumulated in fp64 and rounded
once (oracle.net_restated.CONV_DTYPE) - and the distance between those two CPU runs is the
floor every GPU-vs-oracle distance is held against, per LIF layer, together with detection agreement (AP@0.5,
north_star "detection boxes/mAP identical to 0.1"; reference utils/box.py:102-153, models/soda.py:294-321).
"""
import os

import pytest
import torch
import torch.nn.functional as F

from helpers import ap50, calibrated_tiny_yolo, vgg_cfg
from oracle import net_restated as onr
from snn_for_object_detection_b200 import generator

pytestmark = pytest.mark.gpu
DEV = "cuda"
SPIKE_TOL = 1e-4
V_RTOL = 1e-4
FLOOR_K = 4.0          # a GPU layer must be inside the north_star tolerance outright, or mismatch the fp32 oracle
                       # at most K x what the fp64-conv oracle itself does


def _report(name, lines):
    path = os.environ.get("SODA_PARITY_REPORT_DIR")
    text = "\n".join(lines)
    print(text)
    if path:
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, name), "w") as f:
            f.write(text + "\n")


def _storage_model(seed=5, T=6, B=2, H=64, W=96, density=0.08):
    from snn_for_object_detection_b200 import TinyYolo

    torch.manual_seed(seed)
    model = TinyYolo(num_classes=2, state_storage=True)
    X = (torch.rand(T, B, 2, H, W) < density).float()
    model.eval()
    onr.calibrate(model, X, gain=5.0)
    model.eval()
    return model, X


def _oracle_run(model, X, conv_dtype):
    onr.CONV_DTYPE = conv_dtype
    try:
        taps = {}
        preds, state = onr.soda_forward(model, X, taps)
    finally:
        onr.CONV_DTYPE = "fp32"
    return preds, state, taps


def _cells(model):
    from snn_for_object_detection_b200.modules.common import StateStorage

    return [(n, m) for n, m in model.named_modules() if isinstance(m, StateStorage)]


def _detections(anchors, cls, bbox):
    from snn_for_object_detection_b200.utils import box

    return box.multibox_detection(F.softmax(cls, dim=2), bbox.clone(), anchors)


def test_tensor_core_path_end_to_end_small_net_report_and_first_divergence():
    """TinyYolo three ways on the same input (64x96, B = 2, T = 6): oracle fp32, oracle with fp64-accumulated
    convolutions, and the GPU's default (tcgen05, packed activations) path, per LIF layer in execution order.

    At this size the fp64-conv oracle reproduces the fp32 oracle EXACTLY (measured: no neuron of 19 layers sits within
    fp32 rounding distance of its threshold), so "k x the floor" is vacuous here; the floor comparison proper is the
    GEN1-size test below, where the two CPU runs do diverge.  What is asserted here: every LIF layer up to and
    including the first one with a flipped spike is inside the north_star tolerance (<= 1e-4 of neurons), the
    divergence downstream stays bounded, and the detections agree (AP@0.5, reference utils/box.py:102-153).  The
    per-layer curve goes to profiles/ through SODA_PARITY_REPORT_DIR."""
    model, X = _storage_model()
    p32, _, t32 = _oracle_run(model, X, "fp32")
    p64, _, t64 = _oracle_run(model, X, "fp64")
    cells = _cells(model)
    assert len(cells) == 22
    ref32 = {n: torch.stack([o for o, _ in t32[id(m)]]) for n, m in cells}
    ref64 = {n: torch.stack([o for o, _ in t64[id(m)]]) for n, m in cells}
    model = model.to(DEV)
    with torch.no_grad():
        preds = model(X.to(DEV))
    lines = ["# layer | firing rate | mismatch(fp64-conv oracle, fp32 oracle) | mismatch(GPU tcgen05, fp32 oracle)"]
    checks = []
    for n, m in cells:
        if type(m.module).__name__ != "LIFCell":
            continue
        got = m.get_spikes().cpu()
        f64 = float((ref64[n] != ref32[n]).float().mean())
        gpu = float((got != ref32[n]).float().mean())
        lines.append(f"{n:38s} {float(ref32[n].mean()):.4f} {f64:.3e} {gpu:.3e}")
        checks.append((n, gpu, f64))
    anchors = p32[0]
    det32 = _detections(anchors, p32[1], p32[2])
    det64 = _detections(anchors, p64[1], p64[2])
    detg = _detections(anchors, preds[1].cpu(), preds[2].cpu())
    ap_floor, ap_gpu = ap50(det32, det64), ap50(det32, detg)
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())  # noqa: E731
    lines.append(f"cls  rel err: fp64-oracle {rel(p64[1], p32[1]):.3e}  GPU {rel(preds[1].cpu(), p32[1]):.3e}")
    lines.append(f"bbox rel err: fp64-oracle {rel(p64[2], p32[2]):.3e}  GPU {rel(preds[2].cpu(), p32[2]):.3e}")
    lines.append(f"AP@0.5 vs fp32-oracle detections: fp64-oracle {ap_floor:.4f}  GPU {ap_gpu:.4f}")
    _report("noise_floor_tiny_yolo_64x96.txt", lines)
    assert torch.allclose(preds[0].cpu(), anchors, rtol=0, atol=1e-6)   # (bit-equality vs the reference: test_gpu_nets golden)
    seen_flip = False
    for n, gpu, f64 in checks:
        if not seen_flip:
            assert gpu <= SPIKE_TOL, (n, gpu)           # up to and including the first divergent layer
        seen_flip = seen_flip or gpu > 0
        assert gpu <= max(5e-2, FLOOR_K * f64), (n, gpu, f64)   # chaotic growth stays bounded over 19 layers
    assert ap_gpu >= 0.75, (ap_gpu, ap_floor)


def _flat_cells(state, acc=None):
    acc = [] if acc is None else acc
    if state is None:
        return acc
    if isinstance(state, list):
        for s in state:
            _flat_cells(s, acc)
    elif hasattr(state, "v"):
        acc.append(state)
    return acc


def test_fused_routes_at_config2_slice_against_the_oracle():
    """BASELINE config 2 geometry (GEN1 304x240, odd maps 15x19 -> 8x10), B = 2, T = 8, state_storage OFF: the planner
    takes its production routes (fused Conv->Norm->LIF kernels, event-frame kernel, LIF + residual scans) and the
    final (v, i) of all 22 cells plus the last-step predictions are compared with the oracle DIRECTLY, each cell
    held against the fp64-conv noise floor of the same comparison."""
    T, B, H, W = 8, 2, 240, 304
    model = calibrated_tiny_yolo(seed=0)
    gen = torch.Generator().manual_seed(1)
    X = (torch.rand(T, B, 2, H, W, generator=gen) < 0.07).float()
    p32, s32, _ = _oracle_run(model, X, "fp32")
    p64, s64, _ = _oracle_run(model, X, "fp64")
    c32, c64 = _flat_cells(s32), _flat_cells(s64)
    model = model.to(DEV)
    with torch.no_grad():
        preds, state = model.forward_seq(X.to(DEV))
    cg = _flat_cells(state)
    assert len(c32) == len(c64) == len(cg) == 22
    lines = ["# cell | neurons | frac(|dv| > 1e-4 max|v|): fp64-conv oracle vs fp32 oracle | GPU vs fp32 oracle"]
    checks = []
    for k, (a, b, c) in enumerate(zip(c32, c64, cg)):
        vmax = float(a.v.abs().max())
        f64 = float(((b.v - a.v).abs() > V_RTOL * vmax).float().mean())
        gpu = float(((c.v.cpu() - a.v).abs() > V_RTOL * vmax).float().mean())
        lines.append(f"cell {k:2d} {a.v.numel():9d} {f64:.3e} {gpu:.3e}")
        assert c.v.shape == a.v.shape
        checks.append((k, gpu, f64))
    anchors = p32[0]
    assert anchors.shape[0] == 13545
    ap_floor = ap50(_detections(anchors, p32[1], p32[2]), _detections(anchors, p64[1], p64[2]))
    ap_gpu = ap50(_detections(anchors, p32[1], p32[2]), _detections(anchors, preds[1].cpu(), preds[2].cpu()))
    lines.append(f"AP@0.5 vs fp32-oracle detections: fp64-oracle {ap_floor:.4f}  GPU {ap_gpu:.4f}")
    _report("noise_floor_tiny_yolo_gen1_b2_t8_fused.txt", lines)
    for k, gpu, f64 in checks:
        assert gpu <= max(SPIKE_TOL, FLOOR_K * f64), (k, gpu, f64)
    assert ap_gpu >= min(0.9, ap_floor - 0.1), (ap_gpu, ap_floor)


def _triples(model):
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


def test_every_block_teacher_forced_at_gen1_size():
    """The per-block parity gate at BASELINE config 2 geometry (304x240; maps 120x152 ... 8x10), B = 1, T = 4: every
    Conv -> Norm -> LIF/LI block on the ORACLE's input to that block; spike mismatch <= 1e-4 of neurons, membranes
    of never-flipped neurons within 1e-4 relative."""
    model, X = _storage_model(seed=0, T=4, B=1, H=240, W=304, density=0.07)
    T = X.shape[0]
    taps = {}
    onr.soda_forward(model, X, taps)
    triples = _triples(model)
    assert len(triples) == 22
    model = model.to(DEV)
    worst_s = worst_v = 0.0
    for n, conv, norm, cell in triples:
        outs = torch.stack([o for o, _ in taps[id(cell)]])
        v_ref = torch.stack([s.v for _, s in taps[id(cell)]])
        x = torch.cat(taps[("in", id(conv))]).to(DEV)
        if x.shape[1] != 2:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.no_grad():
            y = generator.conv_forward(conv, norm, x)     # the planner's route: packed-activation tcgen05 kernel
            cell.forward_seq(y, T, None)
        got, v = cell.get_spikes().cpu(), cell.get_state().v.cpu()
        if type(cell.module).__name__ == "LIFCell":
            assert float(outs.mean()) > 0.01, f"{n}: dead layer"
            flips = got != outs
            mism = float(flips.float().mean())
            worst_s = max(worst_s, mism)
            assert mism <= SPIKE_TOL, (n, mism)
            never = ~flips.any(dim=0, keepdim=True).expand_as(flips)
            err = float((v - v_ref)[never].abs().max() / v_ref.abs().max())
        else:
            err = float((v - v_ref).abs().max() / v_ref.abs().max())
        worst_v = max(worst_v, err)
        assert err < V_RTOL, (n, err)
    _report("teacher_forced_gen1.txt", [f"teacher-forced @304x240: worst spike mismatch {worst_s:.2e}; "
                                        f"worst membrane rel err {worst_v:.2e}"])


def test_config1_exactly_against_the_oracle():
    """BASELINE config 1 as written: vgg_block(8)/Pool/vgg_block(32)/Pool/vgg_block(64) (reference README.md:57-64),
    GEN1 304x240, batch 1, T = 8 - the planner's production routes against the time-major CPU oracle."""
    from snn_for_object_detection_b200 import BlockGen

    torch.manual_seed(7)
    T, B, H, W = 8, 1, 240, 304
    X = (torch.rand(T, B, 2, H, W) < 0.07).float()
    block = BlockGen(2, vgg_cfg())
    for m in block.modules():
        if isinstance(m, torch.nn.Conv2d):
            torch.nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
    block.eval()
    onr.calibrate(block, X, gain=5.0, block=block)
    block.eval()
    Yo, so = onr.seq_block_forward(block, X)
    assert Yo.shape == (T, B, 64, 60, 76) and float(Yo.mean()) > 0.02
    block = block.to(DEV)
    with torch.no_grad():
        Y, sg = block.forward_seq(X.to(DEV).reshape(T * B, 2, H, W), T)
    Y = Y.float().reshape(T, B, *Y.shape[1:]).cpu()
    mism = float((Y != Yo).float().mean())
    assert mism <= SPIKE_TOL, mism
    for a, b in zip(_flat_cells(so), _flat_cells(sg)):
        bad = (b.v.cpu() - a.v).abs() > V_RTOL * float(a.v.abs().max())
        assert float(bad.float().mean()) <= SPIKE_TOL
    _report("config1_exact.txt", [f"config 1 (vgg, 304x240, B=1, T=8): output spike mismatch {mism:.2e}, "
                                  f"firing rate {float(Yo.mean()):.3f}"])
