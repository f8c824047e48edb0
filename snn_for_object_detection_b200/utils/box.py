"""Box decode + greedy per-class NMS (role of the reference's ``utils/box.py:72-153``).

Torch restatement of the same algorithm.  ``SODa.predict`` runs the CUDA kernel
(``ops.multibox_detection`` -> ``soda_multibox_detection``, csrc/detect.cu); this module is
its cross-check in the tests and the route for anchor sets beyond the kernel's limit (262 143).
"""

from __future__ import annotations

import torch


def corner_to_center(b: torch.Tensor) -> torch.Tensor:
    x1, y1, x2, y2 = b.unbind(-1)
    return torch.stack(((x1 + x2) / 2, (y1 + y2) / 2, x2 - x1, y2 - y1), dim=-1)


def center_to_corner(b: torch.Tensor) -> torch.Tensor:
    cx, cy, w, h = b.unbind(-1)
    return torch.stack((cx - 0.5 * w, cy - 0.5 * h, cx + 0.5 * w, cy + 0.5 * h), dim=-1)


def box_iou(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Pairwise IoU ``[len(a), len(b)]`` of corner-format boxes."""
    assert a.shape == (a.shape[0], 4) and b.shape == (b.shape[0], 4), "Wrong box shape"
    area_a = torch.prod(a[:, 2:] - a[:, :2], dim=1)
    area_b = torch.prod(b[:, 2:] - b[:, :2], dim=1)
    lo = torch.max(a[:, None, :2], b[:, :2])
    hi = torch.min(a[:, None, 2:], b[:, 2:])
    inter = torch.prod(torch.clamp(hi - lo, min=0), dim=2)
    return inter / (area_a[:, None] + area_b - inter)


def offset_inverse(anchors: torch.Tensor, offsets: torch.Tensor) -> torch.Tensor:
    anc = corner_to_center(anchors)
    xy = (offsets[:, :2] * anc[:, 2:] / 10) + anc[:, :2]
    wh = torch.exp(offsets[:, 2:] / 5) * anc[:, 2:]
    return center_to_corner(torch.cat((xy, wh), dim=1))


def nms(boxes, scores, class_id, num_classes: int, iou_threshold: float) -> torch.Tensor:
    """Greedy NMS per foreground class; returns kept indices, class by class, best first."""
    keep = []
    for c in range(num_classes - 1):
        idx = torch.nonzero(class_id == c).squeeze(1)
        if idx.numel() == 0:
            continue
        order = torch.argsort(scores[idx], descending=True)
        cand = boxes[idx]
        while order.numel() > 0:
            top = order[0]
            keep.append(idx[top])
            if order.numel() == 1:
                break
            iou = box_iou(cand[top].reshape(1, 4), cand[order[1:]]).reshape(-1)
            order = order[1:][iou <= iou_threshold]
    if not keep:
        return torch.empty(0, dtype=torch.long, device=boxes.device)
    return torch.stack(keep).to(torch.long)


def multibox_detection(cls_probs, offset_preds, anchors, nms_threshold: float = 0.1,
                       pos_threshold: float = 0.009999999) -> torch.Tensor:
    """``[batch, anchor, 6]`` rows ``(class, conf, xlu, ylu, xrd, yrd)``; kept boxes first, class -1 =
    background / suppressed (same row order and values as the reference)."""
    batch, num_anchors, num_classes = cls_probs.shape
    out = []
    for b in range(batch):
        conf, class_id = torch.max(cls_probs[b], 1)
        boxes = offset_inverse(anchors, offset_preds[b])
        class_id = class_id - 1
        keep = nms(boxes, conf, class_id, num_classes, nms_threshold)
        mask = torch.ones(num_anchors, dtype=torch.bool, device=boxes.device)
        mask[keep] = False
        non_keep = torch.nonzero(mask).squeeze(1)
        order = torch.cat((keep, non_keep))
        class_id[non_keep] = -1
        class_id, conf, boxes = class_id[order], conf[order], boxes[order]
        low = conf < pos_threshold
        class_id[low] = -1
        conf = torch.where(low, 1 - conf, conf)
        out.append(torch.cat((class_id.unsqueeze(1), conf.unsqueeze(1), boxes), dim=1))
    return torch.stack(out)
