"""Convolutional LSTM of the reference (``models/modules/conv_lstm.py:9-78``) on the sm_100a kernels.

Same module tree and parameter (``conv.weight [4*hid, in+hid, k, k]``, no bias) as the reference.  The gate
convolution over ``cat([X, h])`` is split along its input channels: ``W_x * X`` does not depend on the recurrence and is
computed for ALL T steps in one launch; per step only ``W_h * h_{t-1}`` runs, with the input part as the convolution
kernel's fused residual, followed by one gate kernel (``soda_lstm_gates``).  No concatenation is materialised."""

from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import nn

from .. import ops


class ConvLSTM(nn.Module):
    def __init__(self, in_channels: int, hidden_channels: int, kernel_size: int = 1, bias: bool = False):
        super().__init__()
        if kernel_size != 1 or bias:
            # the reference's conv has no padding, so any k > 1 shrinks h against X and its own cat() fails on step 2
            raise NotImplementedError("ConvLSTM: kernel_size 1 without bias (the reference's LSTM LayerGen, layer_gen.py:300-302)")
        self.in_channels, self.hidden_channels = in_channels, hidden_channels
        self.conv = nn.Conv2d(in_channels + hidden_channels, 4 * hidden_channels, kernel_size, bias=False)
        self._packs = {}

    def _packed(self, part: str, x: torch.Tensor) -> ops.PackedConv:
        w = self.conv.weight
        sl = w[:, : self.in_channels] if part == "x" else w[:, self.in_channels :]
        kind = ops.conv_kind_for(x, sl.shape[1], 1, 1)
        key = (w.data_ptr(), w._version, w.device, kind)
        hit = self._packs.get(part)
        if hit is None or hit[0] != key:
            hit = (key, ops.pack_conv(sl, 1, None, None, kind))
            self._packs[part] = hit
        return hit[1]

    def forward(self, X: torch.Tensor, state: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
        return self.forward_seq(X, 1, state)

    def forward_seq(self, X: torch.Tensor, T: int, state: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
        """``X [T*B, C, H, W]`` time-major; returns ``(h per step [T*B, hid, H, W], (h_T, c_T))``."""
        if torch.is_grad_enabled() and (X.requires_grad or self.conv.weight.requires_grad):
            return self._forward_torch(X, T, state)
        B = X.shape[0] // T
        h, c = (None, None) if state is None else state
        pre_x = ops.conv2d(X, self._packed("x", X))                      # input part, all T steps
        outs = []
        for t in range(T):
            px = pre_x[t * B : (t + 1) * B]
            pre = px if h is None else ops.conv2d(h, self._packed("h", h), residual=px)   # h_0 = 0: no recurrent part
            h, c = ops.lstm_gates(pre, c)
            outs.append(h)
        return (outs[0] if T == 1 else torch.cat(outs)), (h, c)

    def _forward_torch(self, X, T, state):
        """Training route: the reference's own expression (differentiable torch ops)."""
        B = X.shape[0] // T
        hid = self.hidden_channels
        if state is None:
            h = torch.zeros((B, hid) + tuple(X.shape[2:]), device=X.device)
            c = torch.zeros_like(h)
        else:
            h, c = state
        outs = []
        for t in range(T):
            comb = self.conv(torch.cat([X[t * B : (t + 1) * B], h], dim=1))
            gi, gf, go, gg = torch.split(comb, hid, dim=1)
            c = torch.sigmoid(gf) * c + torch.sigmoid(gi) * torch.tanh(gg)
            h = torch.sigmoid(go) * torch.tanh(c)
            outs.append(h)
        return (outs[0] if T == 1 else torch.cat(outs)), (h, c)
