"""Synthetic-weight calibration on the CUDA path.

There are no trained weights for the reference (its checkpoint is a git-LFS pointer) and with its
raw initialisation (kaiming conv, BatchNorm = identity) every LIF layer is silent.
``calibrate`` runs one layer-major pass in which every ``Conv -> Norm`` pair first measures the
per-channel statistics of its own convolution output and stores them as the BatchNorm running
statistics, with gamma = ``gain``: LIF inputs become ~N(0, gain^2) and every layer fires.
Only the statistics reduction uses a stock torch op (``var_mean``); the convolutions run on the
package's kernels.
"""

from __future__ import annotations

import torch

from .. import generator, ops
from ..modules.common import SodaConv2d


@torch.no_grad()
def calibrate(model: torch.nn.Module, X: torch.Tensor, gain: float = 5.0) -> None:
    """``X[T,B,2,H,W]`` on the model's CUDA device.  Mutates BatchNorm buffers / gamma in place."""
    orig = SodaConv2d.forward

    def measuring_forward(self, x, norm=None, residual=None, out=None):
        if norm is not None:
            raw = ops.conv2d(x, ops.pack_conv(self.weight, self.stride[0]))
            var, mean = torch.var_mean(raw, dim=(0, 2, 3), unbiased=False)
            norm.running_mean.copy_(mean)
            norm.running_var.copy_(var.clamp_min(1e-12))
            norm.weight.fill_(gain)
        return orig(self, x, norm=norm, residual=residual, out=out)

    was_training = model.training
    model.eval()
    SodaConv2d.forward = measuring_forward
    # the planner's fused Conv -> Norm -> LIF kernels never call SodaConv2d.forward: plan them out while measuring,
    # or the layers they cover keep identity statistics and stay silent (and with them everything downstream)
    fused_plan = generator._fused_lif_plan
    generator._fused_lif_plan = lambda *a, **k: None
    # ... and neither does the packed-activation route (ops.pconv): measure on the fp32 route
    import os

    saved = {k: os.environ.get(k) for k in ("SODA_PACKED", "SODA_FUSE_LIF")}
    os.environ["SODA_PACKED"] = "0"
    os.environ["SODA_FUSE_LIF"] = "0"
    try:
        model(X)
    finally:
        SodaConv2d.forward = orig
        generator._fused_lif_plan = fused_plan
        for k, v in saved.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v
        model.train(was_training)
