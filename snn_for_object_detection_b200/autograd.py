"""Training-mode routes (BPTT; reference models/soda.py:146-158).

What is hand-written for training this round: every convolution FORWARD, the LIF / LI forward scans and their fused
surrogate-gradient REVERSE scans (``soda_lif_bwd`` / ``soda_li_bwd``, wired in ``modules/cells.py``), and train-mode
BatchNorm forward and backward over all T steps (``soda_bn_train_fwd`` / ``soda_bn_train_bwd``).
The data gradient of stride-1 convolutions runs on our forward kernel (flipped, transposed filter).
What is a library call, stated plainly: stride-2 dgrad and every wgrad go through ``aten::convolution_backward``
(cuDNN) - CUDA, not a CPU or silent fallback, and a next-round kernel.  Residual sums, concat stores
and pooling use differentiable torch ops when a graph is being recorded.
"""

from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import ops


# SODA_DGRAD=0 sends every dgrad to cuDNN (A/B runs)
_DGRAD_NATIVE = os.environ.get("SODA_DGRAD", "1") != "0"


class _ConvFn(torch.autograd.Function):
    """Forward: our sm_100a convolution.  Backward: dgrad of stride-1 layers on the same kernel (flipped, transposed
    filter); stride-2 dgrad and every wgrad: cuDNN (library)."""

    @staticmethod
    def forward(ctx, x, weight, conv):
        k, s = conv.kernel_size[0], conv.stride[0]
        kind = ops.conv_kind_for(x, conv.in_channels, k, s)
        y = ops.conv2d(x, conv.packed(None, kind))
        ctx.save_for_backward(x, weight)
        ctx.stride, ctx.pad = s, k // 2
        ctx.conv = conv
        return y

    @staticmethod
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gx = None
        if need_x and ctx.stride == 1 and _DGRAD_NATIVE and weight.shape[0] % 32 == 0:
            # dgrad of a stride-1 'same' convolution IS a convolution of gy with the spatially flipped, (Cout, Cin)-
            # transposed filter: it runs on our own forward kernel (tcgen05, fp32-accurate operand scheme)
            # (a channel slice of a concat gradient is pixel-major addressable: the kernel reads it through its pixel
            # stride, no copy - ops.conv2d only makes a tensor contiguous when it is not addressable that way)
            gx = ops.conv2d(gy, ctx.conv.packed_dgrad())
            need_x = False
        if need_x or need_w:
            gx2, gw, _ = torch.ops.aten.convolution_backward(
                gy, x, weight, None, [ctx.stride] * 2, [ctx.pad] * 2, [1, 1], False, [0, 0], 1,
                [need_x, need_w, False])
            gx = gx if gx is not None else gx2
        else:
            gw = None
        return gx, gw, None


def conv2d_autograd(conv, x, norm, residual, out):
    if norm is not None:
        raise RuntimeError("a folded (eval-mode) BatchNorm cannot be trained through; call model.train()")
    y = _ConvFn.apply(x, conv.weight, conv)
    if residual is not None:
        y = y + residual
    if out is not None:
        out.copy_(y)   # differentiable slice store (autograd CopySlices)
        return out
    return y


def _all_reduce_sums(sums: torch.Tensor) -> int:
    """The SyncBatchNorm exchange: one all-reduce of the [T, 2, C] fp64 sums; returns the world size."""
    dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    return dist.get_world_size()


class _BNTrainFn(torch.autograd.Function):
    """Per-time-step train-mode BatchNorm on our kernels (soda_bn_train_fwd / soda_bn_train_bwd)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, norm, T):
        ctx.sync = bool(getattr(norm, "sync", False)) and dist.is_available() and dist.is_initialized() \
            and dist.get_world_size() > 1
        if ctx.sync:
            y, mean, invstd, xk, _ = ops.bn_train_fwd_sync(x, T, gamma, beta, norm.eps, norm.momentum,
                                                           norm.running_mean, norm.running_var, _all_reduce_sums)
        else:
            y, mean, invstd, xk = ops.bn_train_fwd(x, T, gamma, beta, norm.eps, norm.momentum, norm.running_mean,
                                                   norm.running_var)
        ctx.save_for_backward(xk, gamma, mean, invstd)
        ctx.T = T
        ctx.affine = gamma is not None
        return y

    @staticmethod
    def backward(ctx, gy):
        xk, gamma, mean, invstd = ctx.saved_tensors
        if ctx.sync:
            gx, gg, gb = ops.bn_train_bwd_sync(gy, xk, ctx.T, gamma, mean, invstd, _all_reduce_sums,
                                               want_affine=ctx.affine)
        else:
            gx, gg, gb = ops.bn_train_bwd(gy, xk, ctx.T, gamma, mean, invstd, want_affine=ctx.affine)
        return gx, (gg if ctx.needs_input_grad[1] else None), (gb if ctx.needs_input_grad[2] else None), None, None


def batchnorm_train(norm, x, T: int = 1):
    """Train-mode BatchNorm exactly as the reference runs it: batch statistics PER CALL, i.e. per time step
    (models/modules/layer_gen.py:211 inside the loop of models/soda.py:142), running statistics updated T times
    in sequence with the module's momentum."""
    if not x.is_cuda:
        raise RuntimeError("snn_for_object_detection_b200 runs on CUDA only; there is no CPU fallback")
    if (ops.bn_train_supported(x) and norm.momentum is not None and os.environ.get("SODA_BN_TRAIN", "1") != "0"
            and (norm.running_mean is None) == (norm.running_var is None)):
        if norm.running_mean is not None and norm.num_batches_tracked is not None:
            norm.num_batches_tracked.add_(T)
        return _BNTrainFn.apply(x, norm.weight, norm.bias, norm, T)
    # channel counts the kernel does not take (C not 4 * 2^k) and cumulative-average momentum: ATen, per time step
    B = x.shape[0] // T
    outs = []
    for t in range(T):
        # nn.BatchNorm2d.forward: the counter always advances; momentum=None means the cumulative average 1 / n
        mom = norm.momentum
        if norm.running_mean is not None and norm.num_batches_tracked is not None:
            norm.num_batches_tracked.add_(1)
            if mom is None:
                mom = 1.0 / float(norm.num_batches_tracked)
        outs.append(F.batch_norm(x[t * B : (t + 1) * B], norm.running_mean, norm.running_var, norm.weight, norm.bias,
                                 True, 0.0 if mom is None else mom, norm.eps))
    return outs[0] if T == 1 else torch.cat(outs)
