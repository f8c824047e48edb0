"""Anchor labelling for the detection loss: the reference's ``utils/roi.py`` interface on the CUDA kernel."""

from __future__ import annotations

from typing import Tuple

import torch

from .. import ops


class RoI:
    """Label anchor boxes using ground-truth bounding boxes (reference ``utils/roi.py:8-63``).

    ``roi(anchors [A,4], labels [B,G,5]) -> (bbox_offset [B,A,4], bbox_mask [B,A,4], class_labels [B,A])`` with
    ``labels`` rows ``(class, x1, y1, x2, y2)`` padded with ``-1``; class 0 = background."""

    def __init__(self, iou_threshold: float = 0.5) -> None:
        self.iou_threshold = iou_threshold

    def __call__(self, anchors: torch.Tensor, labels: torch.Tensor
                 ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        return ops.roi_assign(anchors, labels, self.iou_threshold)
