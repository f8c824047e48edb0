"""Training-mode routes (BPTT).

The fused surrogate-gradient LIF / LI reverse scans are in ``modules/cells.py``
(``soda_lif_bwd`` / ``soda_li_bwd``).  Convolution dgrad / wgrad and train-mode BatchNorm
statistics kernels are not built in this revision: asking for them raises instead of silently
running somewhere else.
"""

from __future__ import annotations


def conv2d_autograd(conv, x, norm, residual, out):
    raise NotImplementedError(
        "SodaConv2d: training-mode convolution (dgrad/wgrad sm_100a kernels) is not built in this "
        "revision; call model.eval() for the inference forward.  No fallback is taken."
    )


def batchnorm_train(norm, x):
    raise NotImplementedError(
        "SodaBatchNorm2d: train-mode batch statistics kernel is not built in this revision; "
        "call model.eval().  No fallback is taken."
    )
