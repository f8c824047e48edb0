"""Shared builders for the tests (configs of BASELINE.json, golden loading)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def vgg_cfg():
    """vgg_block(8)/Pool/vgg_block(32)/Pool/vgg_block(64) (reference README.md:57-64, config 1)."""
    from snn_for_object_detection_b200.modules import LIF, Conv, Norm, Pool

    def vgg_block(out_channels, kernel=3):
        return Conv(out_channels, kernel), Norm(), LIF()

    return [*vgg_block(8), Pool("S"), *vgg_block(32), Pool("S"), *vgg_block(64)]


def vgg_block_from_golden():
    from snn_for_object_detection_b200 import BlockGen

    g = golden("vgg_net.npz")
    block = BlockGen(2, vgg_cfg())
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd::")}
    block.load_state_dict(sd, strict=True)
    block.eval()
    return block, g


def tiny_yolo_from_golden():
    """Rebuilds the golden TinyYolo from its seed (weights are not stored) and checks the digest."""
    from oracle import net_restated as onr
    from oracle.make_golden import weights_digest
    from snn_for_object_detection_b200 import TinyYolo

    g = golden("tiny_yolo.npz")
    torch.manual_seed(int(g["seed"]))
    model = TinyYolo(num_classes=2)
    T, B, H, W = int(g["T"]), int(g["B"]), int(g["H"]), int(g["W"])
    X = (torch.rand(T, B, 2, H, W) < float(g["density"])).float()
    model.eval()
    onr.calibrate(model, X, gain=float(g["gain"]))
    model.eval()
    ok = abs(weights_digest(model.state_dict()) - float(g["digest"])) <= 1e-9 * abs(float(g["digest"]))
    return model, X, g, ok


def calibrated_tiny_yolo(seed=0, gain=5.0, calib_hw=(64, 64), T=6, B=2, density=0.07):
    from oracle import net_restated as onr
    from snn_for_object_detection_b200 import TinyYolo

    torch.manual_seed(seed)
    model = TinyYolo(num_classes=2)
    X = (torch.rand(T, B, 2, *calib_hw) < density).float()
    model.eval()
    onr.calibrate(model, X, gain=gain)
    model.eval()
    return model


def ap50(det_ref: torch.Tensor, det: torch.Tensor) -> float:
    """In-repo AP@0.5 (both COCO back ends of the reference's ``MeanAveragePrecision``,
    models/soda.py:89,294-321, are absent): rows ``(class, conf, x1, y1, x2, y2)`` as ``multibox_detection``
    returns them (class -1 = suppressed).  The kept rows of ``det_ref`` are the ground truth; ``det`` is scored
    against them per class over the whole batch (greedy matching by descending confidence at IoU >= 0.5,
    all-point interpolated precision envelope), mean over the classes present."""
    aps = []
    classes = sorted(set(det_ref[..., 0].flatten().tolist()) - {-1.0})
    for c in classes:
        gts = [det_ref[b][det_ref[b, :, 0] == c][:, 2:6] for b in range(det_ref.shape[0])]
        n_gt = sum(g.shape[0] for g in gts)
        rows = [(float(r[1]), b, r[2:6]) for b in range(det.shape[0]) for r in det[b][det[b, :, 0] == c]]
        rows.sort(key=lambda t: -t[0])
        used = [torch.zeros(g.shape[0], dtype=torch.bool) for g in gts]
        tp = []
        for _, b, bx in rows:
            g = gts[b]
            hit = False
            if g.shape[0]:
                lt, rb = torch.maximum(g[:, :2], bx[:2]), torch.minimum(g[:, 2:], bx[2:])
                inter = (rb - lt).clamp(min=0).prod(dim=1)
                area = (g[:, 2:] - g[:, :2]).prod(dim=1) + (bx[2:] - bx[:2]).prod() - inter
                iou = torch.where(used[b], torch.full_like(inter, -1.0), inter / area)
                j = int(iou.argmax())
                if float(iou[j]) >= 0.5:
                    used[b][j] = True
                    hit = True
            tp.append(hit)
        if n_gt == 0:
            continue
        tp = np.asarray(tp, dtype=np.float64)
        ctp, cfp = np.cumsum(tp), np.cumsum(1.0 - tp)
        rec = np.concatenate([[0.0], ctp / n_gt, [1.0]])
        prec = np.concatenate([[1.0], ctp / np.maximum(ctp + cfp, 1e-12), [0.0]])
        prec = np.maximum.accumulate(prec[::-1])[::-1]
        aps.append(float(np.sum((rec[1:] - rec[:-1]) * prec[1:])))
    return float(np.mean(aps)) if aps else 1.0
