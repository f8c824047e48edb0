"""Small custom modules of the generator API, CUDA-backed.

Mirrors the public surface of the reference's ``models/modules/common.py`` (``SumPool2d`` :18-49,
``Storage`` :52-83, ``StateStorage`` :86-123) and adds the CUDA-backed stand-ins for the torch
layers the reference's LayerGens instantiate (Conv2d, BatchNorm2d, Avg/Max pool).
"""

from __future__ import annotations

from typing import Any, List, Optional

import torch
import torch.nn.functional as F
from torch import nn

from .. import _lib, ops
from .cells import LICell, LIFCell, SLICell, SNNCell

common_list = ("SumPool2d", "Storage", "StateStorage")

__all__ = list(common_list) + ["SodaConv2d", "SodaBatchNorm2d", "SodaPool2d", "merge_states"]


def merge_states(states: List[Any]):
    """Stack every field of a list of state tuples (role of norse ``_merge_states``)."""
    cls = states[0].__class__
    fields = {}
    for key in states[0]._fields:
        vals = [getattr(s, key) for s in states]
        fields[key] = merge_states(vals) if isinstance(vals[0], tuple) else torch.stack(vals)
    return cls(**fields)


class SodaConv2d(nn.Conv2d):
    """``nn.Conv2d(bias=False, padding=k//2)`` whose forward is the sm_100a kernel.

    Subclassing keeps the parameter name/shape (``weight [Cout,Cin,k,k]``), ``isinstance`` checks of
    the generator's initialiser (models/generator.py:245-252) and ``state_dict`` compatibility.
    The packed / BatchNorm-folded weight is a cache keyed on the parameter versions.
    """

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._pack_cache = {}

    def packed(self, norm: Optional["SodaBatchNorm2d"] = None, kind: int = _lib.SODA_CONV_FP32) -> ops.PackedConv:
        key = (self.weight.data_ptr(), self.weight._version, self.weight.device)
        scale = shift = None
        if norm is not None:
            key = key + norm.fold_key()
        hit = self._pack_cache.get(kind)
        if hit is None or hit[0] != key:
            if norm is not None:
                scale, shift = norm.fold()
            hit = (key, ops.pack_conv(self.weight, self.stride[0], scale, shift, kind))
            self._pack_cache[kind] = hit
        return hit[1]

    def packed_p(self, norm: Optional["SodaBatchNorm2d"], in_fmt: int) -> "ops.PackedPConv":
        """fp16 hi | lo weight planes (BatchNorm folded) for the packed-activation convolution, cached like
        :meth:`packed`."""
        key = (self.weight.data_ptr(), self.weight._version, self.weight.device)
        scale = shift = None
        if norm is not None:
            key = key + norm.fold_key()
        slot = ("p", in_fmt)
        hit = self._pack_cache.get(slot)
        if hit is None or hit[0] != key:
            if norm is not None:
                scale, shift = norm.fold()
            hit = (key, ops.pack_pconv(self.weight, self.stride[0], scale, shift, in_fmt))
            self._pack_cache[slot] = hit
        return hit[1]

    def packed_im2col(self, norm: Optional["SodaBatchNorm2d"]) -> "ops.PackedPConv":
        """The 3x3 filter of a layer on few input channels (Cin * 9 <= 32: the event-frame layer) as the 1x1 filter
        over ops.im2col_frames' tap-major 32-channel rows, BatchNorm folded, packed for the tensor-core kernel."""
        key = (self.weight.data_ptr(), self.weight._version, self.weight.device)
        scale = shift = None
        if norm is not None:
            key = key + norm.fold_key()
        hit = self._pack_cache.get("im2col")
        if hit is None or hit[0] != key:
            if norm is not None:
                scale, shift = norm.fold()
            co, ci = self.out_channels, self.in_channels
            w = torch.zeros(co, 32, 1, 1, device=self.weight.device)
            w[:, : ci * 9, 0, 0] = self.weight.detach().permute(0, 2, 3, 1).reshape(co, ci * 9)   # (ky, kx, ci) order
            hit = (key, ops.pack_pconv(w, 1, scale, shift, ops.FMT_PAIR))
            self._pack_cache["im2col"] = hit
        return hit[1]

    def packed_dgrad(self) -> ops.PackedConv:
        """Packed filter of the data-gradient convolution of a stride-1 layer: ``W'[ci, co, ky, kx] =
        W[co, ci, k-1-ky, k-1-kx]`` (tensor-core kind), cached on the weight version."""
        key = (self.weight.data_ptr(), self.weight._version, self.weight.device)
        hit = self._pack_cache.get("dgrad")
        if hit is None or hit[0] != key:
            wt = self.weight.detach().flip(2, 3).transpose(0, 1).contiguous()
            hit = (key, ops.pack_conv(wt, 1, None, None, _lib.SODA_CONV_TC))
            self._pack_cache["dgrad"] = hit
        return hit[1]

    def forward(self, x: torch.Tensor, norm: Optional["SodaBatchNorm2d"] = None, residual=None, out=None):
        if self.bias is not None or self.groups != 1 or self.dilation != (1, 1):
            raise NotImplementedError("SodaConv2d: bias / groups / dilation are not part of the reference path")
        if self.training and torch.is_grad_enabled() and (x.requires_grad or self.weight.requires_grad):
            from ..autograd import conv2d_autograd  # training route (dgrad / wgrad kernels)

            return conv2d_autograd(self, x, norm, residual, out)
        kind = ops.conv_kind_for(x, self.in_channels, self.kernel_size[0], self.stride[0])
        return ops.conv2d(x, self.packed(norm, kind), residual=residual, out=out)


class SodaBatchNorm2d(nn.BatchNorm2d):
    """Inference BatchNorm as a per-channel affine (folded into the producing conv when it follows one)."""

    def fold_key(self):
        b = self.bias
        # the train-mode kernels update the running statistics through raw pointers (no version bump): they bump
        # `_stats_gen` instead (autograd.batchnorm_train), and num_batches_tracked moves with every train step
        nbt = self.num_batches_tracked
        return (
            self.weight.data_ptr(), self.weight._version, self.running_mean._version, self.running_var._version,
            None if b is None else (b.data_ptr(), b._version),
            getattr(self, "_stats_gen", 0), None if nbt is None else nbt._version,
        )

    def fold(self):
        """(scale, shift) with y = x * scale + shift  ==  (x - mean) / sqrt(var + eps) * gamma + beta."""
        inv = torch.rsqrt(self.running_var.double() + self.eps)
        g = self.weight.detach().double() if self.weight is not None else torch.ones_like(inv)
        scale = g * inv
        shift = -self.running_mean.double() * scale
        if self.bias is not None:
            shift = shift + self.bias.detach().double()
        return scale.float(), shift.float()

    def forward(self, x: torch.Tensor, out=None) -> torch.Tensor:
        return self.forward_seq(x, 1, out=out)

    def forward_seq(self, x: torch.Tensor, T: int, out=None) -> torch.Tensor:
        """``x [T*B, C, H, W]``: train mode normalises every time step with its own batch statistics."""
        if self.training:
            from ..autograd import batchnorm_train

            return batchnorm_train(self, x, T)
        scale, shift = self.fold()
        return ops.channel_affine(x, scale, shift, out=out)


class SodaPool2d(nn.Module):
    """Avg / Max pooling (floor mode, no padding) on the CUDA pooling kernel."""

    def __init__(self, kernel_size: int, stride: Optional[int] = None, mode: str = "A"):
        super().__init__()
        self.kernel_size = kernel_size
        self.stride = stride if stride is not None else kernel_size
        self.mode = mode

    def forward(self, X: torch.Tensor) -> torch.Tensor:
        if torch.is_grad_enabled() and X.requires_grad:   # training: differentiable torch pooling
            fn = F.avg_pool2d if self.mode == "A" else F.max_pool2d
            return fn(X, self.kernel_size, self.stride)
        return ops.pool2d(X, self.kernel_size, self.stride, self.mode)


class SumPool2d(nn.Module):
    """Sum over each pooling window (= ``avg_pool2d * k * k`` in the reference, common.py:45-49)."""

    def __init__(self, kernel_size: int, stride: int = 1, padding: int = 0):
        super().__init__()
        if padding:
            raise NotImplementedError("SumPool2d: the generator never passes padding (layer_gen.py:173)")
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding

    def forward(self, X: torch.Tensor) -> torch.Tensor:
        if torch.is_grad_enabled() and X.requires_grad:   # training: differentiable torch pooling
            return F.avg_pool2d(X, self.kernel_size, self.stride) * (self.kernel_size * self.kernel_size)
        return ops.pool2d(X, self.kernel_size, self.stride, "S")


class Storage(nn.Module):
    """Stores the forward-pass value (feature-pyramid tap), returns it unchanged."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.storage = None

    def forward(self, X: torch.Tensor) -> torch.Tensor:
        self.storage = X
        return X

    def get_storage(self) -> torch.Tensor:
        temp = self.storage
        self.storage = None
        return temp


class StateStorage(nn.Module):
    """Wraps a stateful neuron cell and, in eval mode, records per-step state and output.

    In layer-major execution the cell's scan kernel writes the per-step ``(v, i)`` traces directly
    (``v_trace`` / ``i_trace`` of soda_lif_fwd); they are split per time step here so that
    ``get_state()`` / ``get_spikes()`` return what the reference's per-step appends would.
    """

    def __init__(self, m: SNNCell):
        super().__init__()
        self.module = m
        self.state_list: List[Any] = []
        self.spike_list: List[torch.Tensor] = []

    def get_state(self):
        return merge_states(self.state_list)

    def get_spikes(self) -> torch.Tensor:
        return torch.stack(self.spike_list)

    def forward(self, input_tensor: torch.Tensor, state: Optional[Any] = None):
        return self.forward_seq(input_tensor, 1, state)

    def forward_seq(self, x: torch.Tensor, T: int, state: Optional[Any] = None, **kw):
        if state is None:
            self.state_list.clear()
            self.spike_list.clear()
        if self.training or not isinstance(self.module, (LIFCell, LICell, SLICell)):
            return self.module.forward_seq(x, T, state, **kw)
        out, new_state, trace = self.module.forward_seq(x, T, state, want_trace=True, **kw)
        B = x.shape[0] // T
        cls = new_state.__class__
        v_steps = trace.v if trace.v is not None else out
        for t in range(T):
            sl = slice(t * B, (t + 1) * B)
            self.state_list.append(cls(v_steps[sl], trace.i[sl]))
            self.spike_list.append(out[sl])
        return out, new_state


class _CudaUnary:
    """Inference on the elementwise kernel; when a graph is being recorded, the torch module's own forward."""

    _mode = "relu"

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if (torch.is_grad_enabled() and x.requires_grad) or not x.is_cuda or x.dtype != torch.float32:
            return super().forward(x)
        return ops.unary(x, self._mode)


class SodaReLU(_CudaUnary, nn.ReLU):
    _mode = "relu"


class SodaSiLU(_CudaUnary, nn.SiLU):
    _mode = "silu"


class SodaTanh(_CudaUnary, nn.Tanh):
    _mode = "tanh"


class SodaUpsample(nn.Upsample):
    """``nn.Upsample``; mode "nearest" with an integer scale runs on the CUDA kernel (inference)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        s = self.scale_factor
        if (self.mode == "nearest" and s is not None and float(s) == int(s) and x.dim() == 4 and x.is_cuda
                and x.dtype == torch.float32 and not (torch.is_grad_enabled() and x.requires_grad)):
            return ops.upsample_nearest(x, int(s))
        return super().forward(x)
