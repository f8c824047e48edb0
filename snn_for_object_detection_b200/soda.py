"""Detector base class: builds backbone / neck / head from the ``*_cfgs`` hooks and runs the
time loop (role of the reference's ``models/soda.py:18-321``; same constructor arguments,
``forward`` / ``predict`` / ``_forward_impl`` signatures and sub-module names).

The reference is a ``LightningModule``; Lightning (trainer, logging, checkpoint callbacks) is
outside the hot path and not installed here, so this is a plain ``nn.Module`` with the same
``hparams`` attribute access.  ``forward`` executes layer-major (all T steps per layer) — legal
because no state crosses layers — and returns the same last-step predictions.
"""

from __future__ import annotations

from types import SimpleNamespace
from typing import Optional, Tuple

import torch
from torch import nn
from torch.nn import functional as F

from .generator import BackboneGen, Head, ListGen, ListState, NeckGen
from . import ops
from .utils import box
from .utils.roi import RoI


class SODa(nn.Module):
    """Basic object detector class.  Sub-classes define ``backbone_cfgs`` / ``neck_cfgs`` / ``head_cfgs``."""

    def __init__(
        self,
        num_classes: int,
        loss_ratio: float = 0.04,
        time_window: int = 16,
        iou_threshold: float = 0.4,
        learning_rate: float = 0.001,
        state_storage: bool = False,
        init_weights: bool = True,
        plotter=None,
    ):
        super().__init__()
        self.hparams = SimpleNamespace(
            num_classes=num_classes, loss_ratio=loss_ratio, time_window=time_window,
            iou_threshold=iou_threshold, learning_rate=learning_rate, state_storage=state_storage,
            init_weights=init_weights,
        )
        self.plotter = plotter
        self.base_net = BackboneGen(self.backbone_cfgs, in_channels=2, init_weights=init_weights)
        self.neck_net = NeckGen(self.neck_cfgs, self.base_net.out_channels, init_weights=init_weights)
        self.head_net = Head(self.head_cfgs, num_classes, self.neck_net.out_shape, init_weights=init_weights)
        self.roi_blk = RoI(iou_threshold)
        self.cls_loss = nn.CrossEntropyLoss(reduction="none")
        self.box_loss = nn.L1Loss(reduction="none")

    # ---- configuration hooks -------------------------------------------------------------
    def backbone_cfgs(self) -> ListGen:
        raise NotImplementedError

    def neck_cfgs(self) -> ListGen:
        raise NotImplementedError

    def head_cfgs(self, box_out: int, cls_out: int) -> ListGen:
        raise NotImplementedError

    def configure_optimizers(self) -> torch.optim.Optimizer:
        return torch.optim.Adamax(self.parameters(), lr=self.hparams.learning_rate)

    # ---- forward ----------------------------------------------------------------------------
    def forward(self, X: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """``X[T, B, 2, H, W]`` -> last-step ``(anchors[A,4], cls[B,A,C+1], bbox[B,A,4])``."""
        preds, _ = self.forward_seq(X, None)
        return preds

    def forward_seq(self, X: torch.Tensor, state: Optional[ListState] = None):
        """Layer-major pass over all time steps of ``X[T,B,2,H,W]``; also returns the detector
        state after the last step so that a stream can be processed in chunks."""
        T, B = X.shape[0], X.shape[1]
        state = [None] * 3 if state is None else list(state)
        Xf = X.reshape(T * B, *X.shape[2:])
        # between the three generators the feature maps stay in the planner's own activation formats
        base_out, state[0] = self.base_net.forward_seq(Xf, T, state[0], packed_out=True)
        neck_out, state[1] = self.neck_net.forward_seq(base_out, T, state[1], packed_out=True)
        anchors, cls_preds, bbox_preds, state[2] = self.head_net.forward_seq(neck_out, T, state[2])
        return (anchors, cls_preds, bbox_preds), state

    def _forward_impl(self, X: torch.Tensor, state: Optional[ListState]):
        """One time step, state in / state out (reference soda.py:235-244)."""
        state = [None] * 3 if state is None else list(state)
        # = the generators' forward(X, state) with T = 1; the feature maps between them stay in the planner's
        # activation formats exactly as in forward_seq, so both protocols give bit-identical results
        base_out, state[0] = self.base_net.forward_seq(X, 1, state[0], packed_out=True)
        neck_out, state[1] = self.neck_net.forward_seq(base_out, 1, state[1], packed_out=True)
        anchors, cls_preds, bbox_preds, state[2] = self.head_net.forward_seq(neck_out, 1, state[2])
        return (anchors, cls_preds, bbox_preds), state

    def predict(self, X: torch.Tensor, state: Optional[ListState]):
        """Streaming inference on one frame ``X[2,H,W]``: rows ``(class, conf, xlu, ylu, xrd, yrd)``."""
        preds, state = self._forward_impl(X.unsqueeze(0), state)
        anchors, cls, bbox = preds
        probs = F.softmax(cls, dim=2)
        if anchors.shape[0] <= ops.MULTIBOX_MAX_ANCHORS:
            prep_pred = ops.multibox_detection(probs, bbox, anchors).squeeze(0)   # one kernel, no host sync
        else:  # > 262143 anchors: torch restatement of the same algorithm, still on the GPU
            prep_pred = box.multibox_detection(probs, bbox, anchors).squeeze(0)
        prep_pred = prep_pred[prep_pred[:, 0] >= 0]
        prep_pred[:, 2:] = torch.clamp(prep_pred[:, 2:], min=0.0, max=1.0)
        return prep_pred, state

    def _loss(self, preds: Tuple[torch.Tensor, torch.Tensor, torch.Tensor], labels: torch.Tensor) -> torch.Tensor:
        """Detection loss of the reference (``models/soda.py:259-281``): anchors labelled by ``roi_blk`` (one CUDA
        kernel, no host sync), per-anchor cross entropy split into object / background means weighted by
        ``loss_ratio``, plus the masked L1 box loss."""
        anchors, cls_preds, bbox_preds = preds
        bbox_offset, bbox_mask, class_labels = self.roi_blk(anchors, labels)
        num_classes = cls_preds.shape[2]
        cls = self.cls_loss(cls_preds.reshape(-1, num_classes), class_labels.reshape(-1))
        bbox = self.box_loss(bbox_preds * bbox_mask, bbox_offset * bbox_mask)
        mask = class_labels.reshape(-1) > 0
        gt_loss = cls[mask].mean()
        background_loss = cls[~mask].mean()
        return gt_loss * self.hparams.loss_ratio + background_loss * (1 - self.hparams.loss_ratio) + bbox.mean()

    def training_step(self, batch: Tuple[torch.Tensor, torch.Tensor], batch_idx: int = 0) -> torch.Tensor:
        """``models/soda.py:146-158``: forward over a random-length window, loss against the labels."""
        data, labels = batch
        preds = self.forward(data[self._rand_start_time():])
        return self._loss(preds, labels)

    def _rand_start_time(self) -> int:
        if not self.hparams.time_window:
            return 0
        return int(torch.randint(0, self.hparams.time_window, (1,)))
