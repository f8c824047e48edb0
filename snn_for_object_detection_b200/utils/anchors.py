"""Anchor boxes for one pyramid level (role of the reference's ``utils/anchors.py:7-85``).

Runs once per feature-map shape and is cached, so it is host-side set-up, not a kernel.  The
arithmetic keeps the reference's fp32 operation order so the anchors are bit-identical.
"""

from __future__ import annotations

import torch
from torch import nn


class AnchorGenerator(nn.Module):
    """``sizes x ratios`` boxes centred on every pixel, as ``(xlu, ylu, xrd, yrd)`` in [0,1] units."""

    def __init__(self, sizes: torch.Tensor, ratios: torch.Tensor, step: int = 1) -> None:
        super().__init__()
        self.step = step
        # registered (frozen) so that state_dict keys match the reference's (utils/anchors.py:26-27)
        self.sizes = nn.Parameter(sizes, requires_grad=False)
        self.ratios = nn.Parameter(ratios, requires_grad=False)
        self._cache = {}

    def forward(self, X: torch.Tensor) -> torch.Tensor:
        key = (int(X.shape[-2]), int(X.shape[-1]), X.device)
        if key not in self._cache:
            self._cache[key] = self._make(*key)
        return self._cache[key]

    def _make(self, fh: int, fw: int, device) -> torch.Tensor:
        sizes, ratios = self.sizes.to(device), self.ratios.to(device)
        cy = (torch.arange(fh, device=device) + 0.5) * (1.0 / fh)
        cx = (torch.arange(fw, device=device) + 0.5) * (1.0 / fw)
        gy, gx = torch.meshgrid(cy, cx, indexing="ij")
        centres = torch.stack((gx, gy, gx, gy), dim=-1).reshape(fh * fw, 1, 4)
        bw = torch.cat([sizes * r for r in ratios]) * fh / fw
        bh = torch.cat([sizes / r for r in ratios]) * fw / fh
        half = (torch.stack((-bw, -bh, bw, bh)).T / 2).reshape(1, -1, 4)
        return (centres + half).reshape(-1, 4)
