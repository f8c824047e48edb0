"""Torch-facing wrappers of the C ABI (``include/soda_b200.h``).

Each function takes CUDA tensors, derives plain pointers / sizes / strides from them and enqueues
the kernel on torch's current stream.  Nothing here computes on the CPU: a non-CUDA tensor raises.

Layout convention (see the header): activations are fp32 "pixel-major" — logical shape
``[N, C, H, W]`` with ``torch.channels_last`` strides — so that a channel slice of a concat buffer
is addressable as (batch_stride, pix_stride) without a copy.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import NamedTuple, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import ConvDesc, NeuronParams, check, lib

__all__ = [
    "NeuronConst",
    "LIF_DEFAULT",
    "LI_DEFAULT",
    "bin_events",
    "PackedConv",
    "pack_conv",
    "conv2d",
    "conv_lif",
    "lif_scan",
    "lif_scan_out",
    "lif_pool_scan",
    "li_scan",
    "lif_scan_bwd",
    "li_scan_bwd",
    "pool2d",
    "channel_affine",
    "add_n",
    "multibox_detection",
    "roi_assign",
    "to_pixel_major",
    "empty_pixel_major",
]


# When set to a dict, every op records a (start, end) CUDA-event pair on the launching stream under
# its family name: bench.py uses it for the per-kernel-family device time of one eager step.
PROFILE = None
CONV_KERNEL_NAME = ("conv_tc3_kernel (tcgen05/TMEM/TMA implicit GEMM, tf32 main + bf16 correction pass, fused BN/LIF/residual "
                    "epilogue) + conv_ffma_kernel (Cin % 32 != 0)")


def _prof(name):
    def deco(fn):
        def wrapped(*a, **k):
            if PROFILE is None:
                return fn(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            PROFILE.setdefault(name, []).append((e0, e1))
            if name in ("conv2d", "conv_lif"):
                x, pc = a[0], a[1]
                PROFILE.setdefault(name + "_tags", []).append(
                    f"{tuple(x.shape)}->{pc.Cout} k{pc.k} s{pc.stride} kind{getattr(pc, 'kind', 'P')} exact{exact_bound(x)}")
                PROFILE.setdefault(name + "_meta", []).append(
                    dict(N=x.shape[0], Cin=x.shape[1], H=x.shape[2], W=x.shape[3], Cout=pc.Cout, k=pc.k,
                         stride=pc.stride, kind=getattr(pc, 'kind', 2)))   # 2 = packed-activation tensor-core route
            return r

        wrapped.__name__, wrapped.__doc__ = fn.__name__, fn.__doc__
        return wrapped

    return deco


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "snn_for_object_detection_b200 runs on CUDA (sm_100a) only; got a "
                f"{t.device} tensor.  There is no CPU fallback."
            )


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# --------------------------------------------------------------------------------------------
# layout helpers
# --------------------------------------------------------------------------------------------
def empty_pixel_major(n: int, c: int, h: int, w: int, device, dtype=torch.float32) -> torch.Tensor:
    return torch.empty((n, c, h, w), device=device, dtype=dtype, memory_format=torch.channels_last)


def _pm_strides(x: torch.Tensor) -> Optional[Tuple[int, int]]:
    """(batch_stride, pix_stride) if ``x`` ([N,C,H,W]) is pixel-major addressable, else None."""
    n, c, h, w = x.shape
    sn, sc, sh, sw = x.stride()
    if c > 1 and sc != 1:
        return None
    ps = sw if w > 1 else (sh if h > 1 else c)
    if ps < c:
        return None
    if h > 1 and w > 1 and sh != w * ps:
        return None
    if w == 1 and h > 1:
        ps = sh
    bs = sn if n > 1 else h * w * ps
    return bs, ps


def to_pixel_major(x: torch.Tensor) -> torch.Tensor:
    if _pm_strides(x) is not None:
        return x
    return x.contiguous(memory_format=torch.channels_last)


def _is_dense(x: torch.Tensor) -> bool:
    return x.is_contiguous() or x.is_contiguous(memory_format=torch.channels_last)


# --------------------------------------------------------------------------------------------
# event binning (K7)
# --------------------------------------------------------------------------------------------
@_prof("bin_events")
def bin_events(
    events: torch.Tensor,
    sample_offsets: torch.Tensor,
    t0_us: torch.Tensor,
    step_us: int,
    T: int,
    H: int,
    W: int,
    clip_x: bool = True,
    sorted_by_time: bool = True,
    out: Optional[torch.Tensor] = None,
    n_dropped: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Event2D records -> occupancy frames ``[T, B, 2, H, W]`` fp32 {0,1}.

    ``events``: CUDA tensor holding the raw 8-byte little-endian records (uint8 ``[n*8]`` or
    int64/uint64 ``[n]``), samples concatenated; ``sample_offsets`` int64 ``[B+1]``; ``t0_us`` int64 ``[B]``.
    Reference: prophesee_toolbox/src/io/dat_events_tools.py:100-108 + utils/datasets.py:420-433.
    """
    _need_cuda(events, sample_offsets, t0_us, out)
    B = t0_us.numel()
    if sample_offsets.dtype != torch.int64 or t0_us.dtype != torch.int64:
        raise TypeError("sample_offsets and t0_us must be int64")
    if sample_offsets.numel() != B + 1:
        raise ValueError("sample_offsets must have B+1 entries")
    if out is None:
        out = torch.empty((T, B, 2, H, W), device=events.device, dtype=torch.float32)
    else:
        if out.shape != (T, B, 2, H, W) or out.dtype != torch.float32:
            raise ValueError("out must be fp32 [T,B,2,H,W]")
        if out.stride()[2:] != (H * W, W, 1):
            raise ValueError("out frames must be contiguous in (p, y, x)")
    st, sb = out.stride(0), out.stride(1)
    if T == 1:
        st = B * 2 * H * W
    if B == 1:
        sb = 2 * H * W
    check(
        lib.soda_bin_events(
            events.data_ptr(), sample_offsets.data_ptr(), t0_us.data_ptr(), B, int(step_us), T, H, W,
            int(bool(clip_x)), int(bool(sorted_by_time)), out.data_ptr(), st, sb, _ptr(n_dropped), _stream(),
        ),
        "soda_bin_events",
    )
    return out


# --------------------------------------------------------------------------------------------
# convolution (K1, K2, K6)
# --------------------------------------------------------------------------------------------
class PackedConv(NamedTuple):
    packed: torch.Tensor
    shift: Optional[torch.Tensor]
    Cout: int
    Cin: int
    k: int
    stride: int
    kind: int


def pack_conv(
    weight: torch.Tensor,
    stride: int,
    scale: Optional[torch.Tensor] = None,
    shift: Optional[torch.Tensor] = None,
    kind: int = _lib.SODA_CONV_FP32,
) -> PackedConv:
    """Re-layout ``weight [Cout,Cin,k,k]`` (the state_dict tensor) for the kernels; fold ``scale``."""
    _need_cuda(weight, scale, shift)
    Cout, Cin, k, k2 = weight.shape
    if k != k2:
        raise ValueError("square kernels only")
    w = weight.detach().contiguous().float()
    n = lib.soda_conv_packed_elems(Cout, Cin, k, kind)
    if n <= 0:
        raise _lib.SodaError(f"soda_conv_packed_elems: unsupported conv {Cout}x{Cin}x{k} kind {kind}")
    packed = torch.empty(n, device=w.device, dtype=torch.float32)
    sc = None if scale is None else scale.detach().contiguous().float()
    check(
        lib.soda_conv_pack_weights(w.data_ptr(), _ptr(sc), Cout, Cin, k, kind, packed.data_ptr(), _stream()),
        "soda_conv_pack_weights",
    )
    sh = None if shift is None else shift.detach().contiguous().float()
    return PackedConv(packed, sh, Cout, Cin, k, int(stride), kind)


def conv_kind_for(x: torch.Tensor, Cin: int, k: int, stride: int) -> int:
    """SODA_CONV_TC (tcgen05 implicit GEMM) when the shape/layout qualifies, else the FP32-pipe kernel.
    ``SODA_CONV_KIND=fp32`` in the environment forces the FP32-pipe kernel (cross-check runs)."""
    if os.environ.get("SODA_CONV_KIND", "") == "fp32":
        return _lib.SODA_CONV_FP32
    if Cin % 32 != 0 or k not in (1, 3) or stride not in (1, 2):
        return _lib.SODA_CONV_FP32
    st = _pm_strides(x)
    if st is None:
        return _lib.SODA_CONV_TC  # will be made pixel-major by conv2d
    if st[0] % 4 or st[1] % 4 or x.data_ptr() % 16:
        return _lib.SODA_CONV_FP32
    return _lib.SODA_CONV_TC


def mark_exact(t: torch.Tensor, bound: Optional[int] = 1) -> torch.Tensor:
    """Tags a tensor that holds only non-negative integers ``<= bound`` (spikes: 1; sums and sum pools of spike
    maps: the sum of their bounds).  ``bound <= 256`` means exact in bf16, ``<= 2048`` exact in TF32 - the
    tensor-core convolution picks its cheaper operand schemes from that (SODA_CONV_FLAG_INPUT_EXACT_*).
    ``bound=None`` leaves the tensor untagged."""
    if bound is not None:
        t._soda_exact = int(bound)
    return t


def exact_bound(t: Optional[torch.Tensor]) -> Optional[int]:
    b = None if t is None else getattr(t, "_soda_exact", None)
    return None if b is None else int(b)


def sum_bound(*ts) -> Optional[int]:
    """Bound of an elementwise sum (None if any term is untagged)."""
    bs = [exact_bound(t) for t in ts]
    return None if any(b is None for b in bs) else sum(bs)


def _plus(b: Optional[int], k: int) -> Optional[int]:
    return None if b is None else b + k


def is_exact(t: torch.Tensor) -> bool:
    b = exact_bound(t)
    return b is not None and b <= 2048


def _exact_flags(t: torch.Tensor) -> int:
    b = exact_bound(t)
    if b is None or b > 2048:
        return 0
    f = _lib.SODA_CONV_FLAG_INPUT_EXACT_TF32
    if b <= 256:
        f |= _lib.SODA_CONV_FLAG_INPUT_EXACT_BF16
    return f


def conv_out_hw(H: int, W: int, k: int, stride: int) -> Tuple[int, int]:
    pad = k // 2
    return (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1


@_prof("conv2d")
def conv2d(
    x: torch.Tensor,
    pc: PackedConv,
    residual: Optional[torch.Tensor] = None,
    out: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """``conv(x, W*scale) + shift (+ residual)``; x ``[N,Cin,H,W]``; out may be a channel slice.

    Reference: nn.Conv2d(bias=False, padding=k//2) models/modules/layer_gen.py:129-136 with the
    eval BatchNorm affine (:211-214) and the Residual sum (models/generator.py:146) fused.
    """
    _need_cuda(x, residual, out)
    if x.dtype != torch.float32:
        raise TypeError("conv2d: fp32 activations only")
    N, Cin, H, W = x.shape
    if Cin != pc.Cin:
        raise ValueError(f"conv2d: input has {Cin} channels, weights expect {pc.Cin}")
    Ho, Wo = conv_out_hw(H, W, pc.k, pc.stride)
    d = ConvDesc()
    d.N, d.H, d.W, d.Cin, d.Cout, d.k, d.stride, d.kind = N, H, W, Cin, pc.Cout, pc.k, pc.stride, pc.kind
    d.flags = _exact_flags(x)
    if pc.kind == _lib.SODA_CONV_FP32 and x.is_contiguous() and not (Cin == 1 or (H == 1 and W == 1)):
        d.in_nchw = 1  # plain NCHW frames (the binned events)
    else:
        x = to_pixel_major(x)
        d.in_nchw = 0
        d.in_batch_stride, d.in_pix_stride = _pm_strides(x)
    if out is None:
        out = empty_pixel_major(N, pc.Cout, Ho, Wo, x.device)
    elif out.shape != (N, pc.Cout, Ho, Wo):
        raise ValueError(f"conv2d: out has shape {tuple(out.shape)}, expected {(N, pc.Cout, Ho, Wo)}")
    os_ = _pm_strides(out)
    if os_ is None:
        raise ValueError("conv2d: out must be pixel-major (channels_last) addressable")
    d.out_batch_stride, d.out_pix_stride = os_
    if residual is not None:
        if residual.shape != out.shape:
            raise ValueError("conv2d: residual shape mismatch")
        residual = to_pixel_major(residual)
        d.res_batch_stride, d.res_pix_stride = _pm_strides(residual)
    check(
        lib.soda_conv2d_fwd(C.byref(d), x.data_ptr(), pc.packed.data_ptr(), _ptr(pc.shift), _ptr(residual),
                            out.data_ptr(), _stream()),
        "soda_conv2d_fwd",
    )
    return out


def _direct_lif_ok(x: torch.Tensor, pc: "PackedConv") -> bool:
    """Direct FP32-pipe Conv->BN->LIF kernel: plain NCHW frames, tiny K (the event-frame layer)."""
    return (pc.kind == _lib.SODA_CONV_FP32 and x.is_contiguous() and x.shape[1] > 1 and x.shape[2] * x.shape[3] > 1
            and pc.Cout % 8 == 0 and pc.Cin * pc.k * pc.k * pc.Cout <= 11000)


def conv_lif_supported(x: torch.Tensor, pc: "PackedConv") -> bool:
    """The fused Conv->BN->LIF kernels: tensor-core convolutions with Cout % 16 == 0, or the direct
    small-K kernel on NCHW frames."""
    if os.environ.get("SODA_FUSE_LIF", "1") == "0":
        return False
    return (pc.kind == _lib.SODA_CONV_TC and pc.Cout % 16 == 0) or _direct_lif_ok(x, pc)


@_prof("conv_lif")
def conv_lif(
    x: torch.Tensor,
    pc: "PackedConv",
    T: int,
    p: "NeuronConst",
    v0: Optional[torch.Tensor] = None,
    i0: Optional[torch.Tensor] = None,
    residual: Optional[torch.Tensor] = None,
    out: Optional[torch.Tensor] = None,
    out_fmt: Optional[int] = None,
):
    """Fused ``LIF_T(conv(x, W*scale) + shift) (+ residual)`` over ``x [T*B, Cin, H, W]``.

    Returns ``(y, v_T, i_T)``; the membrane state never leaves the SM between time steps.
    Reference: the Conv / Norm / LIF triple of models/tiny_yolo.py:53-58 run T times by
    models/generator.py:184-198 (+ the Residual sum of :146).
    """
    _need_cuda(x, v0, i0, residual, out)
    N, Cin, H, W = x.shape
    if N % T:
        raise ValueError("conv_lif: leading dimension is not a multiple of T")
    B = N // T
    Ho, Wo = conv_out_hw(H, W, pc.k, pc.stride)
    d = ConvDesc()
    d.N, d.H, d.W, d.Cin, d.Cout, d.k, d.stride, d.kind = N, H, W, Cin, pc.Cout, pc.k, pc.stride, pc.kind
    d.flags = _exact_flags(x)
    if _direct_lif_ok(x, pc):
        d.in_nchw = 1
        if residual is not None:
            raise ValueError("conv_lif: the direct (NCHW frames) kernel takes no residual")
    else:
        x = to_pixel_major(x)
        d.in_nchw = 0
        d.in_batch_stride, d.in_pix_stride = _pm_strides(x)
    if out is None:
        out = empty_act(N, pc.Cout, Ho, Wo, x.device, FMT_F32 if out_fmt is None else out_fmt)
    elif out.shape != (N, pc.Cout, Ho, Wo):
        raise ValueError(f"conv_lif: out has shape {tuple(out.shape)}, expected {(N, pc.Cout, Ho, Wo)}")
    if out.dtype != torch.float32:
        if not d.in_nchw:
            raise ValueError("conv_lif: only the direct (NCHW frames) kernel writes packed spikes; use pconv_lif")
        d.flags |= _lib.SODA_CONV_FLAG_OUT_HALF if out.dtype == torch.float16 else _lib.SODA_CONV_FLAG_OUT_PAIR
    os_ = _pm_strides(out)
    if os_ is None:
        raise ValueError("conv_lif: out must be pixel-major (channels_last) addressable")
    d.out_batch_stride, d.out_pix_stride = os_
    if residual is not None:
        if residual.shape != out.shape:
            raise ValueError("conv_lif: residual shape mismatch")
        residual = to_pixel_major(residual)
        d.res_batch_stride, d.res_pix_stride = _pm_strides(residual)
    v = empty_pixel_major(B, pc.Cout, Ho, Wo, x.device)
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    npar = p.c()
    check(
        lib.soda_conv_lif_fwd(C.byref(d), x.data_ptr(), pc.packed.data_ptr(), _ptr(pc.shift), _ptr(residual), T,
                              C.byref(npar), _ptr(v0), _ptr(i0), out.data_ptr(), v.data_ptr(), i.data_ptr(),
                              _stream()),
        "soda_conv_lif_fwd",
    )
    mark_exact(out, 1 if residual is None else _plus(exact_bound(residual), 1))
    return out, v, i


# --------------------------------------------------------------------------------------------
# packed activation formats (rows N1 / N2): PAIR = fp16 hi | fp16 lo per 32-channel group, HALF = fp16
# --------------------------------------------------------------------------------------------
FMT_F32, FMT_PAIR, FMT_HALF = _lib.SODA_ACT_F32, _lib.SODA_ACT_PAIR, _lib.SODA_ACT_HALF
# The storage dtype IS the format tag (it survives views and channel slices): a PAIR tensor has an fp32 tensor's
# shape and strides (4 bytes per value) and is held as int32 so that no float arithmetic can be applied to it by
# accident; a HALF tensor is an ordinary float16 tensor with exact small-integer values.
_FMT_DTYPE = {FMT_F32: torch.float32, FMT_PAIR: torch.int32, FMT_HALF: torch.float16}
_DTYPE_FMT = {v: k for k, v in _FMT_DTYPE.items()}
_STATUS = {}


def fmt_of(t: torch.Tensor) -> int:
    try:
        return _DTYPE_FMT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported activation dtype {t.dtype}") from None


def empty_act(n: int, c: int, h: int, w: int, device, fmt: int) -> torch.Tensor:
    return torch.empty((n, c, h, w), device=device, dtype=_FMT_DTYPE[fmt], memory_format=torch.channels_last)


def _status(device) -> torch.Tensor:
    """Per-device counter of values that left fp16's range when a kernel encoded them as PAIR / HALF."""
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    t = _STATUS.get(key)
    if t is None:
        t = _STATUS[key] = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", key))
    return t


def overflow_count(device=None) -> int:
    """Number of activation values beyond +-65504 met while encoding PAIR / HALF tensors on ``device`` so far
    (synchronises).  Non-zero means the packed route saturated a value: rerun with ``SODA_PACKED=0``."""
    return int(_status("cuda" if device is None else device).item())


def act_convert(x: torch.Tensor, fmt: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Re-encodes a pixel-major activation tensor in another format (or copies it into a slice of the same)."""
    _need_cuda(x, out)
    src_fmt = fmt_of(x)
    if x.dim() != 4:
        raise ValueError("act_convert: [N, C, H, W] tensors only")
    if _pm_strides(x) is None:
        if src_fmt != FMT_F32:
            raise ValueError("act_convert: packed tensors must be pixel-major")
        x = x.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    if out is None:
        out = empty_act(N, Cc, H, W, x.device, fmt)
    elif out.shape != x.shape or fmt_of(out) != fmt:
        raise ValueError("act_convert: out has the wrong shape or format")
    so = _pm_strides(out)
    if so is None:
        raise ValueError("act_convert: out must be pixel-major")
    si = _pm_strides(x)
    check(
        lib.soda_act_convert(x.data_ptr(), src_fmt, si[0], si[1], out.data_ptr(), fmt, so[0], so[1], N, H * W, Cc,
                             _status(x.device).data_ptr(), _stream()),
        "soda_act_convert",
    )
    mark_exact(out, exact_bound(x))
    return out


def im2col_frames(x: torch.Tensor, stride: int) -> torch.Tensor:
    """Plain NCHW fp32 frames ``[N, C, H, W]`` (``C * 9 <= 32``) unfolded for a 3x3 / pad-1 convolution into a PAIR
    tensor ``[N, 32, Ho, Wo]`` (tap-major channels, exact)."""
    _need_cuda(x)
    if x.dim() != 4 or x.dtype != torch.float32 or not x.is_contiguous() or x.shape[1] * 9 > 32:
        raise ValueError("im2col_frames: contiguous NCHW fp32 frames with C * 9 <= 32")
    N, Cc, H, W = x.shape
    Ho, Wo = conv_out_hw(H, W, 3, stride)
    out = empty_act(N, 32, Ho, Wo, x.device, FMT_PAIR)
    so = _pm_strides(out)
    check(lib.soda_im2col3_frames(x.data_ptr(), N, Cc, H, W, stride, out.data_ptr(), so[0], so[1], _stream()),
          "soda_im2col3_frames")
    return mark_exact(out, exact_bound(x) if exact_bound(x) is not None else 1)


def to_f32(x: torch.Tensor) -> torch.Tensor:
    """The fp32 view of an activation tensor of any format (a no-op for fp32)."""
    return x if x.dtype == torch.float32 else act_convert(x, FMT_F32)


class PackedPConv(NamedTuple):
    packed: torch.Tensor
    descale: torch.Tensor
    shift: Optional[torch.Tensor]
    Cout: int
    Cin: int
    k: int
    stride: int
    in_fmt: int


def pconv_shape_ok(Cin: int, Cout: int, k: int, stride: int, in_fmt: int) -> bool:
    return (Cin % (32 if in_fmt == FMT_PAIR else 64) == 0 and Cout % 32 == 0 and k in (1, 3) and stride in (1, 2)
            and os.environ.get("SODA_CONV_KIND", "") != "fp32")


def pack_pconv(weight: torch.Tensor, stride: int, scale: Optional[torch.Tensor], shift: Optional[torch.Tensor],
               in_fmt: int) -> PackedPConv:
    """``weight [Cout,Cin,k,k] * scale`` as fp16 hi | lo planes for soda_pconv_fwd.  The weights are first multiplied
    by a power of two that puts the largest of them near 2^11 (fp16's normal range then covers 15 binades below it);
    the kernel's epilogue multiplies by the inverse.  No host synchronisation: the factor stays on the device."""
    _need_cuda(weight, scale, shift)
    Cout, Cin, k, k2 = weight.shape
    if k != k2:
        raise ValueError("square kernels only")
    w = weight.detach().contiguous().float()
    sc = None if scale is None else scale.detach().contiguous().float()
    m = (w if sc is None else w * sc.view(-1, 1, 1, 1)).abs().max()
    e = torch.floor(torch.log2(2048.0 / m.clamp_min(1e-30))).clamp(-30.0, 30.0)
    mult = torch.exp2(e).reshape(1).float()
    descale = (1.0 / mult).float()
    n = lib.soda_pconv_packed_bytes(Cout, Cin, k)
    if n <= 0:
        raise _lib.SodaError(f"soda_pconv_packed_bytes: unsupported conv {Cout}x{Cin}x{k}")
    packed = torch.empty(n, device=w.device, dtype=torch.uint8)
    check(
        lib.soda_pconv_pack_weights(w.data_ptr(), _ptr(sc), mult.data_ptr(), Cout, Cin, k, in_fmt, packed.data_ptr(),
                                    _stream()),
        "soda_pconv_pack_weights",
    )
    sh = None if shift is None else shift.detach().contiguous().float()
    return PackedPConv(packed, descale, sh, Cout, Cin, k, int(stride), in_fmt)


def _pconv_desc(x, pc: PackedPConv, residual, out, out_fmt):
    N, Cin, H, W = x.shape
    if Cin != pc.Cin:
        raise ValueError(f"pconv: input has {Cin} channels, weights expect {pc.Cin}")
    in_fmt = fmt_of(x)
    if in_fmt != pc.in_fmt:
        raise ValueError("pconv: weights were packed for another input format")
    Ho, Wo = conv_out_hw(H, W, pc.k, pc.stride)
    if out is None:
        out = empty_act(N, pc.Cout, Ho, Wo, x.device, FMT_F32 if out_fmt is None else out_fmt)
    elif out.shape != (N, pc.Cout, Ho, Wo):
        raise ValueError(f"pconv: out has shape {tuple(out.shape)}, expected {(N, pc.Cout, Ho, Wo)}")
    d = _lib.PConvDesc()
    d.N, d.H, d.W, d.Cin, d.Cout, d.k, d.stride = N, H, W, Cin, pc.Cout, pc.k, pc.stride
    d.in_fmt, d.out_fmt = in_fmt, fmt_of(out)
    d.flags = _lib.SODA_PCONV_FLAG_INPUT_EXACT if (in_fmt == FMT_PAIR and exact_bound(x) is not None
                                                   and exact_bound(x) <= 2048) else 0
    si, so = _pm_strides(x), _pm_strides(out)
    if si is None or so is None:
        raise ValueError("pconv: tensors must be pixel-major")
    d.in_batch_stride, d.in_pix_stride = si
    d.out_batch_stride, d.out_pix_stride = so
    if residual is not None:
        if residual.shape != out.shape:
            raise ValueError("pconv: residual shape mismatch")
        sr = _pm_strides(residual)
        if sr is None:
            raise ValueError("pconv: residual must be pixel-major")
        d.res_fmt = fmt_of(residual)
        d.res_batch_stride, d.res_pix_stride = sr
    return d, out


def pconv_ok(x: torch.Tensor, pc_shape, residual=None, out=None) -> bool:
    """Cheap eligibility test of the packed-format tensor-core convolution for this input / destination."""
    Cin, Cout, k, stride = pc_shape
    if x.dtype not in (torch.int32, torch.float16) or not pconv_shape_ok(Cin, Cout, k, stride, fmt_of(x)):
        return False
    for t in (x, residual, out):
        if t is None:
            continue
        st = _pm_strides(t)
        if st is None:
            return False
        b = 2 if t.dtype == torch.float16 else 4
        al = 64 if t.dtype == torch.float16 and t is not x else 128
        if t.data_ptr() % al or (st[0] * b) % al or (st[1] * b) % al:
            return False
    return True


@_prof("conv2d")
def pconv(x: torch.Tensor, pc: PackedPConv, residual: Optional[torch.Tensor] = None,
          out: Optional[torch.Tensor] = None, out_fmt: Optional[int] = None) -> torch.Tensor:
    """``conv(x, W*scale) + shift (+ residual)`` on a PAIR / HALF input; the output format is ``out``'s, else
    ``out_fmt`` (default fp32).  Reference: as :func:`conv2d`."""
    _need_cuda(x, residual, out)
    d, out = _pconv_desc(x, pc, residual, out, out_fmt)
    check(
        lib.soda_pconv_fwd(C.byref(d), x.data_ptr(), pc.packed.data_ptr(), pc.descale.data_ptr(), _ptr(pc.shift),
                           _ptr(residual), out.data_ptr(), _status(x.device).data_ptr(), _stream()),
        "soda_pconv_fwd",
    )
    return out


@_prof("conv_lif")
def pconv_lif(x: torch.Tensor, pc: PackedPConv, T: int, p: "NeuronConst", v0: Optional[torch.Tensor] = None,
              i0: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, out_fmt: Optional[int] = None):
    """Fused ``LIF_T(conv(x, W*scale) + shift) (+ residual)`` on a PAIR / HALF input, see :func:`conv_lif`."""
    _need_cuda(x, v0, i0, residual, out)
    N = x.shape[0]
    if N % T:
        raise ValueError("pconv_lif: leading dimension is not a multiple of T")
    d, out = _pconv_desc(x, pc, residual, out, out_fmt)
    B = N // T
    v = empty_pixel_major(B, pc.Cout, out.shape[2], out.shape[3], x.device)
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    npar = p.c()
    check(
        lib.soda_pconv_lif_fwd(C.byref(d), x.data_ptr(), pc.packed.data_ptr(), pc.descale.data_ptr(), _ptr(pc.shift),
                               _ptr(residual), T, C.byref(npar), _ptr(v0), _ptr(i0), out.data_ptr(), v.data_ptr(),
                               i.data_ptr(), _status(x.device).data_ptr(), _stream()),
        "soda_pconv_lif_fwd",
    )
    mark_exact(out, 1 if residual is None else _plus(exact_bound(residual), 1))
    return out, v, i


def conv_lif_pool_ok(x: torch.Tensor, pc: "PackedConv") -> bool:
    """The fused Conv(3x3) -> BN -> LIF -> 2x2 sum pool kernel takes the NCHW fp32 event frames (Cin = 2) and fp16
    pixel-major pooled counts (Cin = 8)."""
    if pc.kind != _lib.SODA_CONV_FP32 or x.dim() != 4:
        return False
    nchw = x.dtype == torch.float32 and x.is_contiguous() and x.shape[1] > 1
    if not nchw and (x.dtype != torch.float16 or _pm_strides(x) is None):
        return False
    return bool(lib.soda_conv_lif_pool_supported(pc.Cin, pc.Cout, pc.k, pc.stride, x.shape[2], x.shape[3], int(nchw),
                                                 fmt_of(x)))


@_prof("conv_lif")
def conv_lif_pool(x: torch.Tensor, pc: "PackedConv", T: int, p: "NeuronConst", v0=None, i0=None,
                  out: Optional[torch.Tensor] = None, out_fmt: Optional[int] = None):
    """``SumPool2x2(LIF_T(conv3x3(x, W*scale) + shift))`` in one kernel; returns ``(pooled counts, v_T, i_T)`` with the
    state at full resolution.  Reference: the Conv / Norm / LIF / Pool block of README.md:57-64."""
    _need_cuda(x, v0, i0, out)
    N, Cin, H, W = x.shape
    if N % T:
        raise ValueError("conv_lif_pool: leading dimension is not a multiple of T")
    B = N // T
    nchw = x.dtype == torch.float32
    if out is None:
        out = empty_act(N, pc.Cout, H // 2, W // 2, x.device, FMT_F32 if out_fmt is None else out_fmt)
    elif out.shape != (N, pc.Cout, H // 2, W // 2):
        raise ValueError("conv_lif_pool: out has the wrong shape")
    so = _pm_strides(out)
    if so is None:
        raise ValueError("conv_lif_pool: out must be pixel-major")
    si = (0, 0) if nchw else _pm_strides(x)
    v = empty_pixel_major(B, pc.Cout, H, W, x.device)
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    npar = p.c()
    check(
        lib.soda_conv_lif_pool_fwd(x.data_ptr(), fmt_of(x), int(nchw), si[0], si[1], pc.packed.data_ptr(), _ptr(pc.shift),
                                   T, B, H, W, Cin, pc.Cout, C.byref(npar), _ptr(v0), _ptr(i0), out.data_ptr(), fmt_of(out),
                                   so[0], so[1], v.data_ptr(), i.data_ptr(), _stream()),
        "soda_conv_lif_pool_fwd",
    )
    mark_exact(out, 4)
    return out, v, i


def lif_scan_fmt_ok(x: torch.Tensor, residual: Optional[torch.Tensor], out: Optional[torch.Tensor]) -> bool:
    if x.dim() != 4 or x.dtype != torch.float32 or x.shape[1] % 32 or not x.is_contiguous(memory_format=torch.channels_last):
        return False
    for t in (residual, out):
        if t is None:
            continue
        ps = _pix_dense(t)
        if t.shape != x.shape or ps is None:
            return False
        b = 2 if t.dtype == torch.float16 else 4
        al = 128 if t.dtype == torch.int32 else 16
        if t.data_ptr() % al or (ps * b) % al:
            return False
    return True


@_prof("lif_scan")
def lif_scan_fmt(x: torch.Tensor, T: int, v0=None, i0=None, p: "NeuronConst" = None,
                 residual: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                 out_fmt: Optional[int] = None):
    """``LIF_T(x) (+ residual)`` with the residual and the output in any activation format; ``x`` is the dense fp32
    conv output ``[T*B, C, H, W]``.  Returns ``(out, v_T, i_T)``."""
    _need_cuda(x, v0, i0, residual, out)
    p = LIF_DEFAULT if p is None else p
    N, Cc, H, W = x.shape
    if N % T:
        raise ValueError("lif_scan_fmt: leading dimension is not a multiple of T")
    if out is None:
        out = empty_act(N, Cc, H, W, x.device, FMT_F32 if out_fmt is None else out_fmt)
    if not lif_scan_fmt_ok(x, residual, out):
        raise ValueError("lif_scan_fmt: unsupported layout (see lif_scan_fmt_ok)")
    v = torch.empty_like(x[: N // T])
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    pc = p.c()
    n_pix = (N // T) * H * W
    check(
        lib.soda_lif_fmt_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, n_pix, Cc, _ptr(residual),
                             0 if residual is None else fmt_of(residual),
                             0 if residual is None else _pix_dense(residual), out.data_ptr(), fmt_of(out),
                             _pix_dense(out), v.data_ptr(), i.data_ptr(), _status(x.device).data_ptr(), _stream()),
        "soda_lif_fmt_fwd",
    )
    mark_exact(out, 1 if residual is None else _plus(exact_bound(residual), 1))
    return out, v, i


# --------------------------------------------------------------------------------------------
# neuron scans (K3, K4, K8)
# --------------------------------------------------------------------------------------------
class NeuronConst(NamedTuple):
    """fp32 constants exactly as torch forms them for ``python_float * fp32_tensor``."""

    dt_tau_mem: float = 0.1  # fp32(0.001 * 100.0)
    dt_tau_syn: float = 0.2  # fp32(0.001 * 200.0)
    v_leak: float = 0.0
    v_th: float = 1.0
    v_reset: float = 0.0
    alpha: float = 100.0
    jump_first: int = 1
    detach_reset: int = 0

    def c(self) -> NeuronParams:
        return NeuronParams(self.dt_tau_mem, self.dt_tau_syn, self.v_leak, self.v_th, self.v_reset,
                            self.alpha, self.jump_first, self.detach_reset)


LIF_DEFAULT = NeuronConst()
LI_DEFAULT = NeuronConst()


def _flat_T(x: torch.Tensor, T: int) -> Tuple[torch.Tensor, int]:
    if not _is_dense(x):
        x = x.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else x.contiguous()
    total = x.numel()
    if total % T:
        raise ValueError("leading dimension is not a multiple of T")
    return x, total // T


@_prof("lif_scan")
def lif_scan(
    x: torch.Tensor,
    T: int,
    v0: Optional[torch.Tensor] = None,
    i0: Optional[torch.Tensor] = None,
    p: NeuronConst = LIF_DEFAULT,
    want_v_dec: bool = False,
    want_trace: bool = False,
):
    """T steps of norse LIFCell on ``x`` whose leading dim is ``T*B`` (time-major blocks).

    Returns ``(z, v_T, i_T, v_dec_trace|None, v_trace|None, i_trace|None)``; z and the traces have
    x's shape and memory layout; v_T, i_T have the shape of one time block.
    """
    _need_cuda(x, v0, i0)
    x, n = _flat_T(x, T)
    blk_shape = (x.shape[0] // T,) + tuple(x.shape[1:])
    z = mark_exact(torch.empty_like(x))
    v = torch.empty_like(x[: blk_shape[0]])
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    vd = torch.empty_like(x) if want_v_dec else None
    vt = torch.empty_like(x) if want_trace else None
    it = torch.empty_like(x) if want_trace else None
    pc = p.c()
    check(
        lib.soda_lif_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, n, z.data_ptr(), v.data_ptr(),
                         i.data_ptr(), _ptr(vd), _ptr(vt), _ptr(it), _stream()),
        "soda_lif_fwd",
    )
    return z, v, i, vd, vt, it


def _pix_dense(t: torch.Tensor) -> Optional[int]:
    """Pixel stride of a pixel-major tensor whose pixels (over N, H, W) are equally spaced, else None."""
    st = _pm_strides(t)
    if st is None:
        return None
    bs, ps = st
    if t.shape[0] > 1 and bs != t.shape[2] * t.shape[3] * ps:
        return None
    return ps


def lif_scan_out_ok(x: torch.Tensor, residual: Optional[torch.Tensor], out: Optional[torch.Tensor]) -> bool:
    if x.dim() != 4 or x.shape[1] % 4 or not x.is_contiguous(memory_format=torch.channels_last):
        return False
    for t in (residual, out):
        if t is None:
            continue
        ps = _pix_dense(t)
        if t.shape != x.shape or ps is None or ps % 4 or t.data_ptr() % 16:
            return False
    return True


@_prof("lif_scan")
def lif_scan_out(
    x: torch.Tensor,
    T: int,
    v0: Optional[torch.Tensor] = None,
    i0: Optional[torch.Tensor] = None,
    p: NeuronConst = None,
    residual: Optional[torch.Tensor] = None,
    out: Optional[torch.Tensor] = None,
):
    """``LIF_T(x) (+ residual)`` written straight into ``out`` (may be a channel slice of a concat
    buffer).  ``x [T*B, C, H, W]`` dense pixel-major.  Returns ``(out, v_T, i_T)``."""
    _need_cuda(x, v0, i0, residual, out)
    p = LIF_DEFAULT if p is None else p
    N, Cc, H, W = x.shape
    if N % T:
        raise ValueError("lif_scan_out: leading dimension is not a multiple of T")
    if not lif_scan_out_ok(x, residual, out):
        raise ValueError("lif_scan_out: unsupported layout (see lif_scan_out_ok)")
    if out is None:
        out = torch.empty_like(x)
    v = torch.empty_like(x[: N // T])
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    pc = p.c()
    check(
        lib.soda_lif_res_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, (N // T) * H * W, Cc, _ptr(residual),
                             0 if residual is None else _pix_dense(residual), out.data_ptr(), _pix_dense(out),
                             v.data_ptr(), i.data_ptr(), _stream()),
        "soda_lif_res_fwd",
    )
    mark_exact(out, 1 if residual is None else _plus(exact_bound(residual), 1))
    return out, v, i


def lif_pool_ok(x: torch.Tensor, out: Optional[torch.Tensor]) -> bool:
    if x.dim() != 4 or x.shape[1] % 4 or x.shape[2] < 2 or x.shape[3] < 2:
        return False
    if not x.is_contiguous(memory_format=torch.channels_last):
        return False
    if out is not None:
        ps = _pix_dense(out)
        if ps is None or ps % 4 or out.data_ptr() % 16:
            return False
    return True


@_prof("lif_scan")
def lif_pool_scan(x: torch.Tensor, T: int, mode: str, v0: Optional[torch.Tensor] = None,
                  i0: Optional[torch.Tensor] = None, p: NeuronConst = None, out: Optional[torch.Tensor] = None):
    """``Pool2x2(LIF_T(x))`` in one pass (Sum / Avg / Max pooling, stride 2, floor mode); the full-resolution
    spikes never reach HBM.  Returns ``(pooled, v_T, i_T)`` with the state at full resolution."""
    _need_cuda(x, v0, i0, out)
    p = LIF_DEFAULT if p is None else p
    N, Cc, H, W = x.shape
    if N % T:
        raise ValueError("lif_pool_scan: leading dimension is not a multiple of T")
    if not lif_pool_ok(x, out):
        raise ValueError("lif_pool_scan: unsupported layout (see lif_pool_ok)")
    B = N // T
    if out is None:
        out = empty_pixel_major(N, Cc, H // 2, W // 2, x.device)
    elif out.shape != (N, Cc, H // 2, W // 2):
        raise ValueError("lif_pool_scan: out shape mismatch")
    v = torch.empty_like(x[:B])
    i = torch.empty_like(v)
    if v0 is not None:
        v0 = _like_block(v0, v)
        i0 = _like_block(i0, v)
    pc = p.c()
    check(
        lib.soda_lif_pool_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, B, H, W, Cc, _POOL_MODES[mode],
                              out.data_ptr(), _pix_dense(out), v.data_ptr(), i.data_ptr(), _stream()),
        "soda_lif_pool_fwd",
    )
    if mode in ("S", "M"):
        mark_exact(out, 4 if mode == "S" else 1)
    return out, v, i


def _like_block(s: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """State tensor with the same shape and strides as one time block of the activation."""
    if s.shape != ref.shape:
        if s.numel() == 1:  # norse LICell initial v is a scalar tensor
            return torch.full_like(ref, float(s))
        raise ValueError(f"state shape {tuple(s.shape)} does not match activation block {tuple(ref.shape)}")
    if s.stride() != ref.stride() or s.dtype != torch.float32:
        t = torch.empty_like(ref)
        t.copy_(s)
        return t
    return s


@_prof("li_scan")
def li_scan(
    x: torch.Tensor,
    T: int,
    v0: Optional[torch.Tensor] = None,
    i0: Optional[torch.Tensor] = None,
    p: NeuronConst = LI_DEFAULT,
    apply_tanh: bool = False,
    want_trace: bool = False,
):
    """T steps of norse LICell (+ optional Tanh).  Returns ``(out, v_T, i_T, i_trace|None)``."""
    _need_cuda(x, v0, i0)
    x, n = _flat_T(x, T)
    out = torch.empty_like(x)
    v = torch.empty_like(x[: x.shape[0] // T])
    i = torch.empty_like(v)
    if i0 is not None:
        i0 = _like_block(i0, v)
        v0 = _like_block(v0, v)
    it = torch.empty_like(x) if want_trace else None
    pc = p.c()
    check(
        lib.soda_li_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, n, int(apply_tanh), out.data_ptr(),
                        v.data_ptr(), i.data_ptr(), _ptr(it), _stream()),
        "soda_li_fwd",
    )
    return out, v, i, it


@_prof("li_scan")
def sli_scan(x: torch.Tensor, T: int, v0: Optional[torch.Tensor] = None, i0: Optional[torch.Tensor] = None,
             p: NeuronConst = None, want_trace: bool = False):
    """T steps of the reference's saturable leaky integrator (``models/modules/sli.py:110-126``); ``p.v_th`` carries
    the saturation potential.  Returns ``(v per step, v_T, i_T, i_trace|None)``."""
    _need_cuda(x, v0, i0)
    x, n = _flat_T(x, T)
    out = torch.empty_like(x)
    v = torch.empty_like(x[: x.shape[0] // T])
    i = torch.empty_like(v)
    if i0 is not None:
        i0 = _like_block(i0, v)
        v0 = _like_block(v0, v) if v0 is not None and v0.dim() > 0 else None   # a scalar v is the initial v_leak
    it = torch.empty_like(x) if want_trace else None
    pc = p.c()
    check(
        lib.soda_sli_fwd(C.byref(pc), x.data_ptr(), _ptr(v0), _ptr(i0), T, n, out.data_ptr(), v.data_ptr(),
                         i.data_ptr(), _ptr(it), _stream()),
        "soda_sli_fwd",
    )
    return out, v, i, it


@_prof("li_scan")
def synapse_scan(x: torch.Tensor, T: int, p0: Optional[torch.Tensor], tau_secretion: float, tau_dissociation: float,
                 sigma_inhibition: float, dt: float):
    """T steps of the reference's synaptic transmission model (``models/modules/synapse.py:73-103``).
    Returns ``(g per step, p_T)``."""
    _need_cuda(x, p0)
    x, n = _flat_T(x, T)
    out = torch.empty_like(x)
    p = torch.empty_like(x[: x.shape[0] // T])
    if p0 is not None:
        p0 = _like_block(p0, p)
    check(
        lib.soda_synapse_fwd(float(tau_secretion), float(tau_dissociation), float(sigma_inhibition), float(dt),
                             x.data_ptr(), _ptr(p0), T, n, out.data_ptr(), p.data_ptr(), _stream()),
        "soda_synapse_fwd",
    )
    return out, p


@_prof("lstm_gates")
def lstm_gates(pre: torch.Tensor, c_prev: Optional[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(h', c')`` from the gate pre-activations ``pre [B, 4*hid, H, W]`` (pixel-major) and ``c_prev [B, hid, H, W]``
    (reference ``models/modules/conv_lstm.py:62-75``)."""
    _need_cuda(pre, c_prev)
    if not pre.is_contiguous(memory_format=torch.channels_last):
        pre = pre.contiguous(memory_format=torch.channels_last)
    B, C4, H, W = pre.shape
    hid = C4 // 4
    h = empty_pixel_major(B, hid, H, W, pre.device)
    c = empty_pixel_major(B, hid, H, W, pre.device)
    if c_prev is not None and not c_prev.is_contiguous(memory_format=torch.channels_last):
        c_prev = c_prev.contiguous(memory_format=torch.channels_last)
    check(lib.soda_lstm_gates(pre.data_ptr(), _ptr(c_prev), B * H * W, hid, h.data_ptr(), c.data_ptr(), _stream()),
          "soda_lstm_gates")
    return h, c


def lif_scan_bwd(grad_out, grad_vT, grad_iT, v_dec, T: int, p: NeuronConst = LIF_DEFAULT):
    """Fused reverse scan: returns ``(grad_x, grad_v0, grad_i0)``."""
    _need_cuda(grad_out, grad_vT, grad_iT, v_dec)
    g_ps = 0
    if grad_out.stride() != v_dec.stride():
        # a channel slice of a concat gradient (pixel-major, pixels further apart than its own channels) is read in
        # place by the kernel; anything else is re-laid out like v_dec first
        pm = _pm_strides(grad_out) if (grad_out.dim() == 4 and v_dec.dim() == 4 and grad_out.dtype == torch.float32
                                       and v_dec.is_contiguous(memory_format=torch.channels_last)) else None
        Cg = grad_out.shape[1] if grad_out.dim() == 4 else 0
        if (pm is not None and pm[0] == grad_out.shape[2] * grad_out.shape[3] * pm[1] and Cg % 4 == 0 and pm[1] % 4 == 0
                and grad_out.data_ptr() % 16 == 0):
            g_ps = pm[1]
        else:
            g = torch.empty_like(v_dec)
            g.copy_(grad_out)
            grad_out = g
    n = v_dec.numel() // T
    blk = v_dec[: v_dec.shape[0] // T]
    gx = torch.empty_like(v_dec)
    gv0 = torch.empty_like(blk)
    gi0 = torch.empty_like(blk)
    if grad_vT is not None:
        grad_vT = _like_block(grad_vT, blk)
    if grad_iT is not None:
        grad_iT = _like_block(grad_iT, blk)
    pc = p.c()
    if g_ps:
        check(
            lib.soda_lif_bwd_strided(C.byref(pc), grad_out.data_ptr(), grad_out.shape[1], g_ps, _ptr(grad_vT),
                                     _ptr(grad_iT), v_dec.data_ptr(), T, n, gx.data_ptr(), gv0.data_ptr(),
                                     gi0.data_ptr(), _stream()),
            "soda_lif_bwd_strided",
        )
        return gx, gv0, gi0
    check(
        lib.soda_lif_bwd(C.byref(pc), grad_out.data_ptr(), _ptr(grad_vT), _ptr(grad_iT), v_dec.data_ptr(), T, n,
                         gx.data_ptr(), gv0.data_ptr(), gi0.data_ptr(), _stream()),
        "soda_lif_bwd",
    )
    return gx, gv0, gi0


def li_scan_bwd(grad_out, grad_vT, grad_iT, T: int, p: NeuronConst = LI_DEFAULT):
    _need_cuda(grad_out, grad_vT, grad_iT)
    if not _is_dense(grad_out):
        grad_out = grad_out.contiguous()
    n = grad_out.numel() // T
    blk = grad_out[: grad_out.shape[0] // T]
    gx = torch.empty_like(grad_out)
    gv0 = torch.empty_like(blk)
    gi0 = torch.empty_like(blk)
    if grad_vT is not None:
        grad_vT = _like_block(grad_vT, blk)
    if grad_iT is not None:
        grad_iT = _like_block(grad_iT, blk)
    pc = p.c()
    check(
        lib.soda_li_bwd(C.byref(pc), grad_out.data_ptr(), _ptr(grad_vT), _ptr(grad_iT), T, n, gx.data_ptr(),
                        gv0.data_ptr(), gi0.data_ptr(), _stream()),
        "soda_li_bwd",
    )
    return gx, gv0, gi0


# --------------------------------------------------------------------------------------------
# glue kernels (K5, K6)
# --------------------------------------------------------------------------------------------
_POOL_MODES = {"S": 0, "A": 1, "M": 2}


@_prof("pool2d")
def pool2d(x: torch.Tensor, k: int, stride: int, mode: str, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Sum / average / max pooling, floor mode, no padding (models/modules/common.py:45-49)."""
    _need_cuda(x, out)
    x = to_pixel_major(x)
    N, Cc, H, W = x.shape
    Ho, Wo = (H - k) // stride + 1, (W - k) // stride + 1
    if out is None:
        out = empty_pixel_major(N, Cc, Ho, Wo, x.device)
    ibs, ips = _pm_strides(x)
    obs, ops_ = _pm_strides(out)
    check(
        lib.soda_pool2d_fwd(x.data_ptr(), N, H, W, Cc, k, stride, _POOL_MODES[mode], ibs, ips, out.data_ptr(), obs,
                            ops_, _stream()),
        "soda_pool2d_fwd",
    )
    if mode in ("S", "M") and exact_bound(x) is not None:
        mark_exact(out, exact_bound(x) * (k * k if mode == "S" else 1))
    return out


@_prof("channel_affine")
def channel_affine(x: torch.Tensor, scale: Optional[torch.Tensor], shift: Optional[torch.Tensor],
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``y = x * scale[c] + shift[c]`` on pixel-major data; also the strided copy into a concat slice."""
    _need_cuda(x, scale, shift, out)
    x = to_pixel_major(x)
    N, Cc, H, W = x.shape
    if out is None:
        out = empty_pixel_major(N, Cc, H, W, x.device)
    ibs, ips = _pm_strides(x)
    obs, ops_ = _pm_strides(out)
    if N > 1 and (ibs != H * W * ips or obs != H * W * ops_):
        for b in range(N):  # batch-strided slices: one launch per image keeps the kernel 2-D
            channel_affine(x[b : b + 1], scale, shift, out[b : b + 1])
        return out
    check(
        lib.soda_channel_affine(x.data_ptr(), N * H * W, Cc, ips, _ptr(scale), _ptr(shift), out.data_ptr(), ops_,
                                _stream()),
        "soda_channel_affine",
    )
    return out


@_prof("add_n")
def add_n(xs: Sequence[torch.Tensor]) -> torch.Tensor:
    """Residual merge: elementwise sum in list order (models/generator.py:146)."""
    _need_cuda(*xs)
    ref = xs[0] if _is_dense(xs[0]) else xs[0].contiguous(memory_format=torch.channels_last)
    ys = [ref]
    for t in xs[1:]:
        if t.shape != ref.shape:
            raise RuntimeError("residual branches differ in shape")
        if t.stride() != ref.stride():
            u = torch.empty_like(ref)
            u.copy_(t)
            t = u
        ys.append(t)
    out = torch.empty_like(ref)
    arr = (C.c_void_p * len(ys))(*[t.data_ptr() for t in ys])
    check(lib.soda_add_n(arr, len(ys), ref.numel(), out.data_ptr(), _stream()), "soda_add_n")
    return mark_exact(out, sum_bound(*xs))


# --------------------------------------------------------------------------------------------
# detection decode + NMS
# --------------------------------------------------------------------------------------------
MULTIBOX_MAX_ANCHORS = 262143


@_prof("multibox")
def multibox_detection(cls_probs: torch.Tensor, offset_preds: torch.Tensor, anchors: torch.Tensor,
                       nms_threshold: float = 0.1, pos_threshold: float = 0.009999999) -> torch.Tensor:
    """``[B, A, 6]`` rows ``(class, conf, x1, y1, x2, y2)`` exactly as the reference's
    ``utils/box.py:122-153 multibox_detection`` orders and fills them; one CTA per sample, no host sync."""
    _need_cuda(cls_probs, offset_preds, anchors)
    B, A, Cn = cls_probs.shape
    if offset_preds.shape != (B, A, 4) or anchors.shape != (A, 4):
        raise ValueError("multibox_detection: shape mismatch")
    cp = cls_probs.detach().contiguous().float()
    of = offset_preds.detach().contiguous().float()
    an = anchors.detach().contiguous().float()
    ws = torch.empty(lib.soda_multibox_workspace_bytes(B, A), dtype=torch.uint8, device=cp.device)
    out = torch.empty((B, A, 6), dtype=torch.float32, device=cp.device)
    check(
        lib.soda_multibox_detection(cp.data_ptr(), of.data_ptr(), an.data_ptr(), B, A, Cn, float(nms_threshold),
                                    float(pos_threshold), ws.data_ptr(), out.data_ptr(), _stream()),
        "soda_multibox_detection",
    )
    return out


# --------------------------------------------------------------------------------------------
# anchor labelling for the detection loss
# --------------------------------------------------------------------------------------------
@_prof("roi_assign")
def roi_assign(anchors: torch.Tensor, labels: torch.Tensor, iou_threshold: float = 0.5
               ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``(bbox_offset [B,A,4], bbox_mask [B,A,4], class_labels [B,A] int64)`` exactly as the reference's
    ``utils/roi.py:18-63 RoI.__call__`` (padded ``-1`` label rows included); one CTA per sample, no host sync."""
    _need_cuda(anchors, labels)
    if anchors.dim() != 2 or anchors.shape[1] != 4 or labels.dim() != 3 or labels.shape[2] != 5:
        raise ValueError("roi_assign: anchors [A,4] and labels [B,G,5] expected")
    A, (B, G) = anchors.shape[0], labels.shape[:2]
    an = anchors.detach().contiguous().float()
    lb = labels.detach().contiguous().float()
    off = torch.empty((B, A, 4), dtype=torch.float32, device=an.device)
    msk = torch.empty((B, A, 4), dtype=torch.float32, device=an.device)
    cls = torch.empty((B, A), dtype=torch.int64, device=an.device)
    check(
        lib.soda_roi_assign(an.data_ptr(), lb.data_ptr(), B, A, G, float(iou_threshold), off.data_ptr(),
                            msk.data_ptr(), cls.data_ptr(), _stream()),
        "soda_roi_assign",
    )
    return off, msk, cls


# --------------------------------------------------------------------------------------------
# train-mode BatchNorm over T time steps (BPTT route)
# --------------------------------------------------------------------------------------------
def bn_train_supported(x: torch.Tensor) -> bool:
    C = x.shape[1] if x.dim() == 4 else 0
    return (x.dim() == 4 and x.is_cuda and x.dtype == torch.float32 and C >= 4 and C <= 1024 and C & (C - 1) == 0)


@_prof("bn_train")
def bn_train_fwd(x: torch.Tensor, T: int, gamma, beta, eps: float, momentum: float, running_mean, running_var):
    """Per-time-step batch statistics + normalise for ``x [T*B, C, H, W]``; updates the running statistics T times
    in order.  Returns ``(y, save_mean [T,C], save_invstd [T,C])``; ``y`` is channels_last like the kernel's ``x``."""
    _need_cuda(x, gamma, beta, running_mean, running_var)
    if not _is_dense(x) or not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    if N % T:
        raise ValueError("bn_train: leading dimension is not a multiple of T")
    y = torch.empty_like(x)
    mean = torch.empty((T, Cc), dtype=torch.float32, device=x.device)
    invstd = torch.empty_like(mean)
    ws = torch.empty(lib.soda_bn_train_workspace_bytes(T, Cc), dtype=torch.uint8, device=x.device)
    check(
        lib.soda_bn_train_fwd(x.data_ptr(), T, (N // T) * H * W, Cc, _ptr(gamma), _ptr(beta), float(eps), float(momentum),
                              _ptr(running_mean), _ptr(running_var), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                              ws.data_ptr(), _stream()),
        "soda_bn_train_fwd",
    )
    return y, mean, invstd, x


@_prof("bn_train")
def bn_train_bwd(gy: torch.Tensor, x: torch.Tensor, T: int, gamma, mean: torch.Tensor, invstd: torch.Tensor,
                 want_affine: bool = True):
    """``(grad_x, grad_gamma, grad_beta)`` of :func:`bn_train_fwd`; ``x`` is the channels_last tensor it returned."""
    _need_cuda(gy, x, gamma, mean, invstd)
    if gy.stride() != x.stride():
        gy = gy.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    gx = torch.empty_like(x)
    gg = torch.empty(Cc, dtype=torch.float32, device=x.device) if want_affine else None
    gb = torch.empty(Cc, dtype=torch.float32, device=x.device) if want_affine else None
    ws = torch.empty(lib.soda_bn_train_workspace_bytes(T, Cc), dtype=torch.uint8, device=x.device)
    check(
        lib.soda_bn_train_bwd(gy.data_ptr(), x.data_ptr(), T, (N // T) * H * W, Cc, _ptr(gamma), mean.data_ptr(),
                              invstd.data_ptr(), gx.data_ptr(), _ptr(gg), _ptr(gb), ws.data_ptr(), _stream()),
        "soda_bn_train_bwd",
    )
    return gx, gg, gb


def bn_train_stats(x: torch.Tensor, T: int, gy: Optional[torch.Tensor] = None, mean=None, invstd=None) -> torch.Tensor:
    """This rank's per-(t, c) sums ``[T, 2, C]`` fp64: ``(sum x, sum x^2)``, or ``(sum gy, sum gy * xhat)`` with ``gy``."""
    _need_cuda(x, gy, mean, invstd)
    N, Cc, H, W = x.shape
    sums = torch.empty((T, 2, Cc), dtype=torch.float64, device=x.device)
    ws = torch.empty(lib.soda_bn_train_workspace_bytes(T, Cc), dtype=torch.uint8, device=x.device)
    check(lib.soda_bn_train_stats(x.data_ptr(), _ptr(gy), _ptr(mean), _ptr(invstd), T, (N // T) * H * W, Cc,
                                  sums.data_ptr(), ws.data_ptr(), _stream()), "soda_bn_train_stats")
    return sums


@_prof("bn_train")
def bn_train_fwd_sync(x: torch.Tensor, T: int, gamma, beta, eps: float, momentum: float, running_mean, running_var,
                      all_reduce) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, int]:
    """SyncBatchNorm forward: local sums -> ``all_reduce(sums) -> world size`` (in place) -> normalise with the
    global statistics.  Returns ``(y, save_mean, save_invstd, x as used, world)``."""
    _need_cuda(x, gamma, beta, running_mean, running_var)
    if not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    if N % T:
        raise ValueError("bn_train: leading dimension is not a multiple of T")
    n_pix = (N // T) * H * W
    sums = bn_train_stats(x, T)
    world = int(all_reduce(sums))
    y = torch.empty_like(x)
    mean = torch.empty((T, Cc), dtype=torch.float32, device=x.device)
    invstd = torch.empty_like(mean)
    ws = torch.empty(lib.soda_bn_train_workspace_bytes(T, Cc), dtype=torch.uint8, device=x.device)
    check(
        lib.soda_bn_train_fwd_from_sums(x.data_ptr(), sums.data_ptr(), n_pix * world, T, n_pix, Cc, _ptr(gamma), _ptr(beta),
                                        float(eps), float(momentum), _ptr(running_mean), _ptr(running_var), y.data_ptr(),
                                        mean.data_ptr(), invstd.data_ptr(), ws.data_ptr(), _stream()),
        "soda_bn_train_fwd_from_sums",
    )
    return y, mean, invstd, x, world


@_prof("bn_train")
def bn_train_bwd_sync(gy: torch.Tensor, x: torch.Tensor, T: int, gamma, mean, invstd, all_reduce, want_affine=True):
    """SyncBatchNorm backward: dx from the all-reduced ``(sum gy, sum gy * xhat)``, dgamma / dbeta from the local sums."""
    _need_cuda(gy, x, gamma, mean, invstd)
    if gy.stride() != x.stride():
        gy = gy.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    n_pix = (N // T) * H * W
    local = bn_train_stats(x, T, gy, mean, invstd)
    glob = local.clone()
    world = int(all_reduce(glob))
    gx = torch.empty_like(x)
    gg = torch.empty(Cc, dtype=torch.float32, device=x.device) if want_affine else None
    gb = torch.empty(Cc, dtype=torch.float32, device=x.device) if want_affine else None
    ws = torch.empty(lib.soda_bn_train_workspace_bytes(T, Cc), dtype=torch.uint8, device=x.device)
    check(
        lib.soda_bn_train_bwd_from_sums(gy.data_ptr(), x.data_ptr(), glob.data_ptr(), n_pix * world, local.data_ptr(), T,
                                        n_pix, Cc, _ptr(gamma), mean.data_ptr(), invstd.data_ptr(), gx.data_ptr(),
                                        _ptr(gg), _ptr(gb), ws.data_ptr(), _stream()),
        "soda_bn_train_bwd_from_sums",
    )
    return gx, gg, gb


_UNARY_MODES = {"relu": 0, "silu": 1, "tanh": 2}


@_prof("elementwise")
def unary(x: torch.Tensor, mode: str) -> torch.Tensor:
    """ReLU / SiLU / Tanh (reference LayerGens, ``models/modules/layer_gen.py:255-284``); keeps x's memory layout."""
    _need_cuda(x)
    if not _is_dense(x):
        x = x.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else x.contiguous()
    y = torch.empty_like(x)
    check(lib.soda_unary_fwd(x.data_ptr(), x.numel(), _UNARY_MODES[mode], y.data_ptr(), _stream()), "soda_unary_fwd")
    if mode == "relu":
        mark_exact(y, exact_bound(x))
    return y


@_prof("elementwise")
def upsample_nearest(x: torch.Tensor, scale: int) -> torch.Tensor:
    """``nn.Upsample(scale_factor=scale, mode="nearest")`` (reference LayerGen ``Up``, ``layer_gen.py:176-194``) on
    pixel-major frames."""
    _need_cuda(x)
    if not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    N, Cc, H, W = x.shape
    y = empty_pixel_major(N, Cc, H * scale, W * scale, x.device)
    check(lib.soda_upsample_nearest(x.data_ptr(), N, H, W, Cc, int(scale), y.data_ptr(), _stream()),
          "soda_upsample_nearest")
    return mark_exact(y, exact_bound(x))
