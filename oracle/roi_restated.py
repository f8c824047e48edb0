"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's anchor labelling
(SURVEY.md 8f-2), pinned against the reference's own ``utils/roi.py`` by ``tests/golden/roi.npz``
(``oracle/make_golden.py:golden_roi``).

Follows ``utils/roi.py:18-63`` (``RoI.__call__``), ``:65-109`` (``_assign_anchor_to_box``), ``utils/box.py:24-57``
(``box_iou``), ``:8-17`` (``box_corner_to_center``) and ``:60-67`` (``offset_boxes``).  numpy fp32, explicit loops
over the ground-truth boxes only.
"""
import numpy as np

F = np.float32


def box_iou(anchors: np.ndarray, gt: np.ndarray) -> np.ndarray:
    """[A,4] x [G,4] corner boxes -> IoU [A,G], same operation order as utils/box.py:38-57 (fp32)."""
    a, g = anchors.astype(F), gt.astype(F)
    area_a = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
    area_g = (g[:, 2] - g[:, 0]) * (g[:, 3] - g[:, 1])
    ul = np.maximum(a[:, None, :2], g[None, :, :2])
    lr = np.minimum(a[:, None, 2:], g[None, :, 2:])
    wh = np.maximum(lr - ul, F(0))
    inter = wh[..., 0] * wh[..., 1]
    union = (area_a[:, None] + area_g[None, :]) - inter
    with np.errstate(invalid="ignore", divide="ignore"):
        return (inter / union).astype(F)


def assign_anchor_to_box(gt: np.ndarray, anchors: np.ndarray, iou_threshold: float) -> np.ndarray:
    """utils/roi.py:65-109: per-anchor best box above the threshold, then every box claims its best free anchor
    (global argmax with row / column discards; ties -> smallest flattened index, torch.argmax's CPU rule)."""
    A, G = anchors.shape[0], gt.shape[0]
    jac = box_iou(anchors, gt)
    amap = np.full((A,), -1, dtype=np.int64)
    if G == 0:
        return amap
    best = jac.argmax(axis=1)                       # first maximal index, as torch.max(dim=1)
    mx = jac[np.arange(A), best]
    keep = mx >= F(iou_threshold)
    amap[keep] = best[keep]
    for _ in range(G):
        flat = int(jac.argmax())                    # first maximal index of the row-major flattening
        j = flat % G
        a = int(F(flat) / F(G))                     # (max_idx / num_gt_boxes).long(): fp32 division, truncation
        amap[a] = j
        jac[:, j] = F(-1)
        jac[a, :] = F(-1)
    return amap


def offset_boxes(anchors: np.ndarray, assigned: np.ndarray, eps: float = 1e-6) -> np.ndarray:
    """utils/box.py:60-67 in fp32."""
    a, b = anchors.astype(F), assigned.astype(F)
    acx, acy = (a[:, 0] + a[:, 2]) / F(2), (a[:, 1] + a[:, 3]) / F(2)
    aw, ah = a[:, 2] - a[:, 0], a[:, 3] - a[:, 1]
    bcx, bcy = (b[:, 0] + b[:, 2]) / F(2), (b[:, 1] + b[:, 3]) / F(2)
    bw, bh = b[:, 2] - b[:, 0], b[:, 3] - b[:, 1]
    with np.errstate(divide="ignore", invalid="ignore"):
        ox = F(10) * (bcx - acx) / aw
        oy = F(10) * (bcy - acy) / ah
        ow = F(5) * np.log(F(eps) + bw / aw, dtype=F)
        oh = F(5) * np.log(F(eps) + bh / ah, dtype=F)
    return np.stack([ox, oy, ow, oh], axis=1).astype(F)


def roi(anchors: np.ndarray, labels: np.ndarray, iou_threshold: float = 0.5):
    """utils/roi.py:18-63.  anchors [A,4], labels [B,G,5] = (class, x1, y1, x2, y2), padded rows are all -1.
    Returns (bbox_offset [B,A,4] f32, bbox_mask [B,A,4] f32, class_labels [B,A] i64)."""
    B, A = labels.shape[0], anchors.shape[0]
    off = np.zeros((B, A, 4), F)
    msk = np.zeros((B, A, 4), F)
    cls = np.zeros((B, A), np.int64)
    for b in range(B):
        lab = labels[b].astype(F)
        amap = assign_anchor_to_box(lab[:, 1:], anchors, iou_threshold)
        m = amap >= 0
        msk[b] = np.repeat(m.astype(F)[:, None], 4, axis=1)
        assigned = np.zeros((A, 4), F)
        cls[b, m] = lab[amap[m], 0].astype(np.int64) + 1
        assigned[m] = lab[amap[m], 1:]
        off[b] = offset_boxes(anchors, assigned) * msk[b]
    return off, msk, cls
