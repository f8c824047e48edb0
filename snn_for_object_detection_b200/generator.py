"""Model generation tools: recursive net builder + layer-major executor.

Public surface of the reference's ``models/generator.py``: ``ListGen`` / ``ListState`` types,
``BlockGen`` (:35-198), ``ModelGen`` (:206-275), ``BackboneGen`` (:283-295), ``NeckGen`` (:303-351),
``Head`` (:359-457), ``HeadGen`` (:465-538) with the same constructor arguments, ``forward``
signatures, module tree (hence ``state_dict`` keys) and state nesting.

What is different underneath: the network has no inter-layer recurrence (every state is private to
one cell), so besides the per-step protocol ``forward(X, state)`` every generator also has
``forward_seq(X, T, state)`` which takes all T time steps at once, folded time-major into the batch
dimension (``X[T*B, C, H, W]``).  A layer then runs ONCE over all T steps: convolutions see a batch
of T*B images, neuron cells scan T steps inside one kernel with the membrane state in registers.
``forward(X, state)`` is ``forward_seq(X, 1, state)``.
"""

from __future__ import annotations

import inspect
import os
from typing import List, Optional, Tuple, Union

import torch
from torch import nn

from . import _lib, ops
from .modules import *  # noqa: F401,F403
from .modules.cells import LICell, LIFCell, LIFFeedForwardState, SNNCell
from .modules.common import SodaBatchNorm2d, SodaConv2d, SodaPool2d, StateStorage, Storage, SumPool2d
from .modules.conv_lstm import ConvLSTM
from .modules.layer_gen import Dense, LayerGen, Residual, Return
from .utils.anchors import AnchorGenerator

ListGen = List[Union[LayerGen, "ListGen"]]
ListState = List[Union[torch.Tensor, None, "ListState"]]

__all__ = ["ListGen", "ListState", "BlockGen", "ModelGen", "BackboneGen", "NeckGen", "Head", "HeadGen"]


def _is_module_stateful(module: nn.Module) -> bool:
    """A module is stateful iff its ``forward`` takes a ``state`` argument (norse utils/state.py)."""
    return "state" in inspect.signature(module.forward).parameters or isinstance(module, nn.RNNBase)


def _has_state(block: "BlockGen") -> bool:
    for flags, branch in zip(block.branch_state, block.net):
        for f, layer in zip(flags, branch):
            if f and (not isinstance(layer, BlockGen) or _has_state(layer)):
                return True
    return False


class _CatParts(list):
    """Training route: the pieces of a Dense merge that is the LAST layer of a branch of an enclosing Dense block,
    handed up unconcatenated.  The reference nests its C2f bottlenecks as Dense-in-Dense (models/tiny_yolo.py:67-75),
    so ``torch.cat`` at every level copies the innermost tensors once per level (n + 1 levels for n bottlenecks);
    only the enclosing merge consumes the result, and it concatenates all pieces once."""


def _packed() -> bool:
    """Inference keeps inter-layer activations in the tensor core's own operand formats (ops.FMT_PAIR /
    FMT_HALF, rows N1 / N2 of the scope table) unless ``SODA_PACKED=0``; a recorded (training) graph is fp32."""
    if os.environ.get("SODA_CONV_KIND", "") == "fp32":
        return False   # the bit-faithful FP32-pipe cross-check route keeps fp32 activations end to end
    return os.environ.get("SODA_PACKED", "1") != "0" and not torch.is_grad_enabled()


def _real_fmt(C: int) -> int:
    return ops.FMT_PAIR if (_packed() and C % 32 == 0) else ops.FMT_F32


def _spike_fmt(C: int) -> int:
    if not _packed() or C % 32:
        return ops.FMT_F32
    return ops.FMT_HALF if C % 64 == 0 else ops.FMT_PAIR


def _wants_f32(layer) -> bool:
    """Layers that read fp32: neuron cells (their scans take the conv output at full precision), stand-alone
    BatchNorm, pooling, activations, user modules.  Convs, nested blocks, taps and Pass take any format."""
    return not isinstance(layer, (SodaConv2d, BlockGen, Storage, nn.Identity))


def _copy_into(src: torch.Tensor, dst: torch.Tensor) -> None:
    """Strided pixel-major copy on our own kernel (concat-slice store, generator.py:158)."""
    if torch.is_grad_enabled() and src.requires_grad:
        dst.copy_(src)   # training: differentiable slice store
        return
    if src.dtype != torch.float32 or dst.dtype != torch.float32:
        ops.act_convert(src, ops.fmt_of(dst), out=dst)
        return
    ops.channel_affine(src, None, None, out=dst)
    ops.mark_exact(dst, ops.exact_bound(src))


class _ColConv:
    """The event-frame layer seen as a 1x1 convolution over ops.im2col_frames' 32-channel rows (same interface as
    the SodaConv2d attributes the planner reads)."""

    in_channels, kernel_size, stride, bias, groups, dilation = 32, (1, 1), (1, 1), None, 1, (1, 1)

    def __init__(self, conv: SodaConv2d):
        self.conv, self.out_channels = conv, conv.out_channels

    def packed_p(self, norm, in_fmt: int):
        return self.conv.packed_im2col(norm)


def _im2col_route(conv: SodaConv2d, Y: torch.Tensor) -> bool:
    """K = Cin * 9 <= 32 on plain NCHW fp32 frames (the detector's first layer, models/tiny_yolo.py:16): unfold the
    frames once and run a 1x1 layer on the tensor-core kernel.  Decided by shapes alone - never by batch size or T - so
    that sharded, chunked and per-step execution all take the same arithmetic."""
    return (_packed() and Y.dtype == torch.float32 and Y.dim() == 4 and Y.is_contiguous() and _conv_eligible(conv)
            and conv.kernel_size == (3, 3) and conv.padding == (1, 1) and Y.shape[1] == conv.in_channels
            and conv.in_channels * 9 <= 32 and conv.out_channels % 32 == 0 and Y.shape[1] > 1
            and os.environ.get("SODA_IM2COL_FIRST", "1") != "0")


def _conv_eligible(conv) -> bool:
    return conv.bias is None and conv.groups == 1 and conv.dilation == (1, 1)


def _conv_shape(conv: SodaConv2d):
    return conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.stride[0]


def _packed_input(Y: torch.Tensor, conv) -> Optional[torch.Tensor]:
    """``Y`` in the format the packed-activation convolution reads (converted if it arrives as fp32), or None
    when this convolution (a module, or a ``(Cin, Cout, k, stride)`` tuple) cannot take that route."""
    if isinstance(conv, tuple):
        Cin, Cout, k, s = conv
    else:
        Cin, Cout, k, s = _conv_shape(conv)
        if not _conv_eligible(conv):
            return None
    if not _packed() or Y.dim() != 4:
        return None
    if Y.dtype == torch.int32:
        return Y if ops.pconv_shape_ok(Cin, Cout, k, s, ops.FMT_PAIR) else None
    if Y.dtype == torch.float16:
        if ops.pconv_shape_ok(Cin, Cout, k, s, ops.FMT_HALF):
            return Y
        return ops.act_convert(Y, ops.FMT_PAIR) if ops.pconv_shape_ok(Cin, Cout, k, s, ops.FMT_PAIR) else None
    if Y.dtype != torch.float32 or not ops.pconv_shape_ok(Cin, Cout, k, s, ops.FMT_PAIR):
        return None
    b = ops.exact_bound(Y)
    if b is not None and b <= 2048 and ops.pconv_shape_ok(Cin, Cout, k, s, ops.FMT_HALF):
        return ops.act_convert(Y, ops.FMT_HALF)
    return ops.act_convert(Y, ops.FMT_PAIR)


def _run_conv(conv: SodaConv2d, norm, Y: torch.Tensor, out: Optional[torch.Tensor], out_fmt: int,
              residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Conv (+ folded eval BatchNorm) on an input of any format; result in ``out`` (its format) or a new tensor of
    ``out_fmt``."""
    Yp = _packed_input(Y, conv)
    if Yp is not None:
        shape = (conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.stride[0])
        if ops.pconv_ok(Yp, shape, residual, out):
            return ops.pconv(Yp, conv.packed_p(norm, ops.fmt_of(Yp)), residual=residual, out=out, out_fmt=out_fmt)
        if isinstance(conv, _ColConv):
            # destination not addressable by the tensor-core store (misaligned slice): via a temporary
            tmp = ops.pconv(Yp, conv.packed_p(norm, ops.fmt_of(Yp)), out_fmt=ops.FMT_F32)
            if residual is not None:
                tmp = ops.add_n([tmp, ops.to_f32(residual)])
            if out is None:
                return tmp if out_fmt == ops.FMT_F32 else ops.act_convert(tmp, out_fmt)
            _copy_into(tmp, out)
            return out
    Yf = ops.to_f32(Y)
    rf = None if residual is None else ops.to_f32(residual)
    if out is not None and out.dtype != torch.float32:
        tmp = conv(Yf, norm=norm, residual=rf)
        ops.act_convert(tmp, ops.fmt_of(out), out=out)
        return out
    y = conv(Yf, norm=norm, residual=rf, out=out)
    if out is None and out_fmt != ops.FMT_F32 and not torch.is_grad_enabled():
        y = ops.act_convert(y, out_fmt)
    return y


def conv_forward(conv: SodaConv2d, norm, X: torch.Tensor, out_fmt: int = ops.FMT_F32) -> torch.Tensor:
    """Conv (+ folded eval BatchNorm) exactly as the inference planner runs it for an input of any format - the
    packed-activation tensor-core kernel where the layer qualifies, the event-frame im2col route, else the fp32
    kernels - without any fusion with what follows.  (Used by the per-block parity tests.)"""
    if _im2col_route(conv, X):
        X = ops.im2col_frames(X, conv.stride[0])
        conv = _ColConv(conv)
    return _run_conv(conv, norm, X, None, out_fmt)


def _fuse_items_ok(Cout: int, Ho: int, Wo: int, B: int, T: int) -> bool:
    """The fused kernel walks the T steps inside one work item, so it needs (channel chunks x spatial tiles x
    samples) >= about one wave of CTAs; small maps at small batch stay unfused."""
    items = -(-Cout // (32 if Cout <= 32 else 64)) * (-(-Ho // 16)) * (-(-Wo // 8)) * B
    return T == 1 or items >= int(os.environ.get("SODA_FUSE_LIF_MIN_ITEMS", "600"))


def _fused_lif_plan(conv: SodaConv2d, norm: SodaBatchNorm2d, x: torch.Tensor, T: int):
    """Packed weights for the fused Conv->Norm->LIF kernel, or None when the unfused pair
    (conv over T*B images + stand-alone scan) is the better launch.  The fused kernel walks the T
    steps inside one work item, so it needs (channel chunks x spatial tiles x samples) >= about one
    wave of CTAs; small maps at small batch stay unfused."""
    if conv.bias is not None or conv.groups != 1 or conv.dilation != (1, 1):
        return None
    k, s = conv.kernel_size[0], conv.stride[0]
    kind = ops.conv_kind_for(x, conv.in_channels, k, s)
    pc = conv.packed(norm, kind)
    if not ops.conv_lif_supported(x, pc):
        return None
    if pc.kind != _lib.SODA_CONV_TC:
        return pc  # direct small-K kernel (event-frame layer): thread-per-pixel, always enough work
    Ho, Wo = ops.conv_out_hw(x.shape[2], x.shape[3], k, s)
    B = x.shape[0] // T
    items = -(-pc.Cout // (32 if pc.Cout <= 32 else 64)) * (-(-Ho // 16)) * (-(-Wo // 8)) * B
    if T > 1 and items < int(os.environ.get("SODA_FUSE_LIF_MIN_ITEMS", "600")):
        return None
    # The fused kernel keeps 2 x 32 state registers per epilogue thread, i.e. 64-channel chunks with
    # N = 64 MMAs; measured on B200 it beats conv + scan (+ add) only for the 32-channel layers.
    if pc.Cout > int(os.environ.get("SODA_FUSE_LIF_MAX_COUT", "32")):
        return None
    return pc


class BlockGen(nn.Module):
    """Block generator: inner lists are sequential, outer lists are parallel branches merged by the
    list type (plain = first branch, ``Residual`` = sum, ``Dense`` = channel concat)."""

    def __init__(self, in_channels: int, cfgs: ListGen):
        super().__init__()
        branch_list: List[nn.ModuleList] = []
        self.branch_state: List[List[bool]] = []
        self.out_channels = 0
        if isinstance(cfgs, Residual):
            self.merge = "residual"
        elif isinstance(cfgs, Dense):
            self.merge = "dense"
        else:
            self.merge = "forward"
            cfgs = [cfgs]
        self.branch_channels: List[int] = []
        for branch_cfg in cfgs:
            layer_list, state_layers, out_channels = self._make_branch(in_channels, branch_cfg)
            branch_list.append(layer_list)
            self.branch_state.append(state_layers)
            self.branch_channels.append(out_channels)
            self._channels_calc(out_channels)
        self.net = nn.ModuleList(branch_list)

    def _channels_calc(self, channels: int) -> None:
        if self.merge == "dense":
            self.out_channels += channels
        elif self.merge == "residual":
            if not self.out_channels:
                self.out_channels = channels
            elif self.out_channels != channels:
                raise RuntimeError(
                    "[ERROR]: The number of channels in the residual "
                    "network does not match! Check the configuration settings."
                )
        else:
            self.out_channels = channels

    def _make_branch(self, in_channels: int, cfg: ListGen) -> Tuple[nn.ModuleList, List[bool], int]:
        state_layers: List[bool] = []
        layer_list: List[nn.Module] = []
        channels = in_channels
        for layer_gen in cfg:
            if isinstance(layer_gen, list):
                layer = BlockGen(channels, layer_gen)
                channels = layer.out_channels
            else:
                layer, channels = layer_gen.get(channels)
            layer_list.append(layer)
            state_layers.append(_is_module_stateful(layer))
        return nn.ModuleList(layer_list), state_layers, channels

    # ------------------------------------------------------------------------------------------
    def forward(self, X: torch.Tensor, state: Optional[ListState] = None) -> Tuple[torch.Tensor, ListState]:
        """Per-step protocol of the reference (generator.py:169-198)."""
        return self.forward_seq(X, 1, state)

    def forward_seq(
        self, X: torch.Tensor, T: int, state: Optional[ListState] = None, out: Optional[torch.Tensor] = None,
        packed_out: bool = False, fold: Optional[SodaConv2d] = None, lazy_cat: bool = False,
    ) -> Tuple[torch.Tensor, ListState]:
        """All T steps of the block; ``X`` is ``[T*B, C, H, W]`` time-major.  ``out``: optional
        pre-allocated (possibly channel-sliced) destination for the block's result.  ``packed_out``: the caller
        is another planner and takes the result in whatever activation format it was produced in; otherwise
        (the public call) the result is fp32."""
        Y, out_state = self._forward_seq(X, T, state, out, f32_out=not packed_out and out is None, fold=fold,
                                         lazy_cat=lazy_cat)
        if isinstance(Y, _CatParts):
            return Y, out_state      # (only ever asked for by an enclosing Dense merge of the training route)
        if not packed_out and Y.dtype != torch.float32:
            Y = ops.to_f32(Y)
        return Y, out_state

    def _ch32(self) -> bool:
        """Every channel count inside this block (hence every concat-slice offset) is a multiple of 32."""
        hit = getattr(self, "_ch32_cache", None)
        if hit is None:
            hit = all(c % 32 == 0 for c in self.branch_channels) and all(
                m._ch32() for br in self.net for m in br if isinstance(m, BlockGen))
            self._ch32_cache = hit
        return hit

    # ---- C2f stems: 1x1 convolutions that open the branches of a Dense block --------------------------------
    def _stem_plan(self):
        """A Dense block whose branches all START with a plain 1x1 / stride-1 convolution on the block's input (no
        BatchNorm or neuron directly behind it) - the two ``Conv(half, 1)`` of a C2f block, models/tiny_yolo.py:77-89.
        Returns ``[(branch index, conv, lo, hi)]``, ``[lo, hi)`` = where that convolution's output lives in the
        block's concat buffer, when those ranges are adjacent: ONE convolution with the stacked filters then writes
        all of them and reads the input once (a 1x1 convolution is a matrix product per pixel; stacking output
        channels changes no arithmetic)."""
        hit = getattr(self, "_stem_cache", False)
        if hit is not False:
            return hit
        plan = None
        if self.merge == "dense" and len(self.net) >= 2:
            plan, c0 = [], 0
            for bi, branch in enumerate(self.net):
                s0, s1 = c0, c0 + self.branch_channels[bi]
                c0 = s1
                conv = branch[0] if len(branch) else None
                ok = (isinstance(conv, SodaConv2d) and _conv_eligible(conv) and conv.kernel_size == (1, 1)
                      and conv.stride == (1, 1))
                if ok and len(branch) == 1:
                    lo, hi = s0, s1
                elif ok and len(branch) == 2 and isinstance(branch[1], BlockGen) and branch[1]._identity_slice():
                    a, b = branch[1]._identity_slice()
                    lo, hi = s0 + a, s0 + b
                    ok = branch[1].out_channels == s1 - s0 and branch[1].out_hw(64, 64) == (64, 64)
                else:
                    ok = False
                if not ok or hi - lo != conv.out_channels or conv.in_channels != self.net[0][0].in_channels:
                    plan = None
                    break
                plan.append((bi, conv, lo, hi))
            if plan is not None:
                order = sorted(plan, key=lambda t: t[2])
                if any(a[3] != b[2] for a, b in zip(order, order[1:])):
                    plan = None
                else:
                    plan = order
        self._stem_cache = plan
        return plan

    def _stem_weights(self, plan, fold: Optional[SodaConv2d], in_fmt: int):
        """Stacked (and, with ``fold``, pre-multiplied) filters of the stem convolutions, packed for ops.pconv.
        ``fold`` is the 1x1 convolution that feeds this block and nothing else (C2f's entry ``Conv(out, 1)``,
        models/tiny_yolo.py:79): with no nonlinearity in between, stem(fold(x)) = (W_stem . W_fold) x; the product is
        formed once in float64 and rounded to fp32, the intermediate tensor is never materialised."""
        convs = [c for _, c, _, _ in plan] + ([fold] if fold is not None else [])
        key = tuple((c.weight.data_ptr(), c.weight._version) for c in convs) + (in_fmt, convs[0].weight.device)
        hit = getattr(self, "_stem_pack", None)
        if hit is None or hit[0] != key:
            w = torch.cat([c.weight.detach() for _, c, _, _ in plan], dim=0)
            if fold is not None:
                w = (w[:, :, 0, 0].double() @ fold.weight.detach()[:, :, 0, 0].double()).float()[:, :, None, None]
            hit = (key, ops.pack_pconv(w.contiguous(), 1, None, None, in_fmt))
            self._stem_pack = hit
        return hit[1]

    def _run_stem(self, X, dense_buf, fold: Optional[SodaConv2d]):
        """Runs the fused stem convolution into its slice of ``dense_buf``; returns {branch index: its tensor} or
        None when this block / input / buffer does not qualify."""
        plan = self._stem_plan()
        if plan is None or dense_buf.dtype != torch.int32 or os.environ.get("SODA_STEM_FUSION", "1") == "0":
            return None
        cin = fold.in_channels if fold is not None else plan[0][1].in_channels
        ctot = plan[-1][3] - plan[0][2]
        Xp = _packed_input(X, (cin, ctot, 1, 1))
        if Xp is None:
            return None
        out = dense_buf[:, plan[0][2]:plan[-1][3]]
        if not ops.pconv_ok(Xp, (cin, ctot, 1, 1), None, out):
            return None
        ops.pconv(Xp, self._stem_weights(plan, fold, ops.fmt_of(Xp)), out=out)
        return {bi: dense_buf[:, lo:hi] for bi, _, lo, hi in plan}

    def _forward_seq(self, X, T, state, out, f32_out=False, fold=None, lazy_cat=False):
        """``f32_out``: the result leaves the planner (public call): whatever produces it writes fp32 directly.
        ``fold``: a 1x1 convolution the caller skipped because this block's stems absorb it (see _stem_weights)."""
        outs = []
        out_state = []
        state = [None] * len(self.net) if state is None else state
        dense_buf = None
        # Training route (a graph is being recorded): no in-place slice stores - autograd's CopySlices backward clones
        # the whole concat gradient once per slice - but the reference's own merges, torch.cat / a sum of the branches.
        recording = torch.is_grad_enabled() and (X.requires_grad or any(p.requires_grad for p in self.parameters()))
        if recording:
            out = None
        if self.merge == "dense" and not recording:
            Ho, Wo = self.out_hw(X.shape[2], X.shape[3])
            dense_buf = out if out is not None else ops.empty_act(
                X.shape[0], self.out_channels, Ho, Wo, X.device,
                ops.FMT_PAIR if (_packed() and self._ch32() and not f32_out) else ops.FMT_F32,
            )
        stems = None
        if dense_buf is not None and _packed():
            stems = self._run_stem(X, dense_buf, fold)
        if fold is not None and stems is None:
            X = _run_conv(fold, None, X, None, _real_fmt(fold.out_channels))   # could not be absorbed: run it after all
        c0 = 0
        for bi, (branch, flags, branch_state) in enumerate(zip(self.net, self.branch_state, state)):
            branch_state = [None] * len(branch) if branch_state is None else list(branch_state)
            dst = None
            if self.merge == "dense" and not recording:
                dst = dense_buf[:, c0 : c0 + self.branch_channels[bi]]
                c0 += self.branch_channels[bi]
            elif self.merge == "forward":
                dst = out
            # Residual([[..., Conv, Norm, LIF], [Pass]]): the identity branch is added in the fused
            # Conv->Norm->LIF epilogue of the other branch instead of a separate add_n pass.
            fuse_res = None
            if self.merge == "residual" and len(self.net) == 2 and self._is_identity(self.net[1 - bi]):
                fuse_res = X
            if stems is not None:
                # the branch's opening convolution has already been computed (into its place in the concat buffer)
                rest_state = branch_state[1:]
                Y, consumed = self._run_branch(stems[bi], T, branch[1:], flags[1:], rest_state, dst)
                branch_state = [None] + rest_state
            else:
                Y, consumed = self._run_branch(X, T, branch, flags, branch_state, dst, fuse_res, out,
                                               f32_out=f32_out and self.merge != "dense",
                                               lazy_last=recording and self.merge == "dense")
            out_state.append(branch_state)
            if consumed:
                # the other branch is the identity: its state list is all None
                other = [None] * len(self.net[1 - bi])
                return Y, ([branch_state, other] if bi == 0 else [other, branch_state])
            outs.append(Y)
        if self.merge == "dense" and recording:
            parts = []
            for o in outs:
                if isinstance(o, _CatParts):
                    parts.extend(o)
                else:
                    parts.append(o)
            if lazy_cat:
                return _CatParts(parts), out_state
            return torch.cat(parts, dim=1), out_state   # models/generator.py:158 (nested merges flattened)
        if self.merge == "dense":
            # a concat of spike maps / spike counts is itself one (largest bound): the next Conv may use the
            # bf16-only operand scheme
            bounds = [ops.exact_bound(o) for o in outs]
            if bounds and all(b is not None for b in bounds):
                ops.mark_exact(dense_buf, max(bounds))
            return dense_buf, out_state
        if self.merge == "residual":
            if torch.is_grad_enabled() and any(o.requires_grad for o in outs):
                Y = outs[0]
                for o in outs[1:]:
                    Y = Y + o                  # training: differentiable merge (= torch.stack(out).sum(0) for two branches)
            else:
                Y = ops.add_n([ops.to_f32(o) for o in outs])
            if out is not None:
                _copy_into(Y, out)
                Y = out
            return Y, out_state
        return outs[0], out_state

    def out_hw(self, H: int, W: int) -> Tuple[int, int]:
        """Spatial size of the block's output for an ``H x W`` input (first branch decides)."""
        for layer in self.net[0]:
            if isinstance(layer, nn.Conv2d):
                H, W = ops.conv_out_hw(H, W, layer.kernel_size[0], layer.stride[0])
            elif isinstance(layer, BlockGen):
                H, W = layer.out_hw(H, W)
            elif isinstance(layer, nn.Upsample):
                H, W = int(H * layer.scale_factor), int(W * layer.scale_factor)
            elif hasattr(layer, "kernel_size") and hasattr(layer, "stride"):  # pooling modules
                k, s = layer.kernel_size, layer.stride
                H, W = (H - k) // s + 1, (W - k) // s + 1
        return H, W

    @staticmethod
    def _is_identity(branch) -> bool:
        return len(branch) > 0 and all(type(m) is nn.Identity for m in branch)

    def _identity_slice(self) -> Optional[Tuple[int, int]]:
        """Channel range of the single ``[Pass]`` branch of a Dense block (its input re-appears there)."""
        if self.merge != "dense":
            return None
        ids = [bi for bi, br in enumerate(self.net) if self._is_identity(br)]
        if len(ids) != 1:
            return None
        lo = sum(self.branch_channels[: ids[0]])
        return lo, lo + self.branch_channels[ids[0]]

    def _lookahead_dst(self, branch, end_idx, n, N, C, H, W, dst, device, prealloc):
        """Destination for the output of the layer group ending at ``end_idx`` when the NEXT layer is a Dense
        block with a ``[Pass]`` branch: the group writes straight into that block's concat buffer (allocated
        here, or the caller's ``dst`` when the block closes the branch), so the identity branch costs no copy."""
        if end_idx + 1 >= n or torch.is_grad_enabled():
            return None
        nxt = branch[end_idx + 1]
        if not isinstance(nxt, BlockGen):
            return None
        sl = nxt._identity_slice()
        if sl is None or sl[1] - sl[0] != C or nxt.out_hw(H, W) != (H, W):
            return None
        buf = dst if (end_idx + 1 == n - 1 and dst is not None) else ops.empty_act(
            N, nxt.out_channels, H, W, device, ops.FMT_PAIR if (_packed() and nxt._ch32()) else ops.FMT_F32)
        prealloc[end_idx + 1] = buf
        return buf[:, sl[0] : sl[1]]

    def _next_fmt(self, branch, end_idx: int, n: int, C: int, spikes: bool, f32_out: bool = False) -> int:
        """Format of the output of the layer group ending at ``end_idx`` when no destination buffer dictates it:
        fp32 for a consumer that reads fp32 (or the public caller), else the tensor core's operand format."""
        if end_idx + 1 < n and _wants_f32(branch[end_idx + 1]):
            return ops.FMT_F32
        if f32_out and all(isinstance(m, (nn.Identity, Storage)) for m in branch[end_idx + 1:]):
            return ops.FMT_F32
        return _spike_fmt(C) if spikes else _real_fmt(C)

    def _fused_conv_lif(self, conv, norm, cell, Y, T, v0, i0, res, o, ofmt, Ho, Wo):
        """ONE kernel for Conv -> Norm(eval) -> LIF over all T steps (membrane state on-chip, Residual identity
        added, result stored in the destination's format), or None when no fused kernel takes this case."""
        if Y.dtype == torch.float32:
            pc = _fused_lif_plan(conv, norm, Y, T)
            if pc is None:
                return None
            if pc.kind != _lib.SODA_CONV_TC:
                # direct small-K kernel on the NCHW event frames: writes spikes in any format, takes no residual
                if res is not None:
                    return None
                return ops.conv_lif(Y, pc, T, cell._const, v0, i0, out=o, out_fmt=ofmt)
            if ofmt != ops.FMT_F32 or (res is not None and res.dtype != torch.float32):
                return None
            return ops.conv_lif(Y, pc, T, cell._const, v0, i0, residual=res, out=o)
        shape = (conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.stride[0])
        if not ops.pconv_ok(Y, shape, res, o) or not _fuse_items_ok(conv.out_channels, Ho, Wo, Y.shape[0] // T, T):
            return None
        return ops.pconv_lif(Y, conv.packed_p(norm, ops.fmt_of(Y)), T, cell._const, v0, i0, residual=res, out=o,
                             out_fmt=ofmt)

    def _run_branch(self, X, T, branch, flags, branch_state, dst, residual=None, residual_dst=None, f32_out=False,
                    lazy_last=False):
        """Runs one branch; fuses Conv+Norm(eval)(+LIF) and LI+Tanh; the branch's last layer writes
        straight into ``dst`` (concat slice / caller's buffer) when it can.  ``residual``: tensor the
        caller wants added to the branch result (then written to ``residual_dst`` if given); returns
        ``(Y, residual_consumed)``.  Activations travel in the format their consumer reads (ops.FMT_*)."""
        n = len(branch)
        Y = X
        idx = 0
        prealloc = {}  # index of a Dense block -> its concat buffer, allocated early by _lookahead_dst
        pending_dst = None  # (index of a LIF cell, destination its scan should write) set by an unfused Conv/Norm/LIF
        nograd = not torch.is_grad_enabled()
        while idx < n:
            layer = branch[idx]
            if isinstance(layer, SodaConv2d):
                norm = None
                last = idx
                if idx + 1 < n and isinstance(branch[idx + 1], SodaBatchNorm2d) and not branch[idx + 1].training:
                    norm = branch[idx + 1]
                    last = idx + 1
                cell = branch[last + 1] if last + 1 < n else None
                Cout = layer.out_channels
                Ho, Wo = ops.conv_out_hw(Y.shape[2], Y.shape[3], layer.kernel_size[0], layer.stride[0])
                fuse_on = os.environ.get("SODA_FUSE_LIF", "1") != "0"
                if (norm is not None and type(cell) is LIFCell and nograd and _conv_eligible(layer) and fuse_on
                        and last + 2 < n and isinstance(branch[last + 2], SumPool2d)
                        and branch[last + 2].kernel_size == 2 and branch[last + 2].stride == 2
                        and not (last + 2 == n - 1 and residual is not None)):
                    # Conv, Norm, LIF, Pool("S") block with a small K (README.md:57-64): one kernel, pooled counts out
                    pc = layer.packed(norm, _lib.SODA_CONV_FP32)
                    if ops.conv_lif_pool_ok(Y, pc):
                        lif_idx, pool_idx = last + 1, last + 2
                        is_last = pool_idx == n - 1
                        o = dst if is_last else None
                        if o is not None:
                            ofmt = ops.fmt_of(o)
                        elif (not is_last and isinstance(branch[pool_idx + 1], SodaConv2d) and Cout == 8 and _packed()
                              and branch[pool_idx + 1].kernel_size == (3, 3) and branch[pool_idx + 1].stride == (1, 1)):
                            ofmt = ops.FMT_HALF      # the next block's fused kernel reads 8-channel fp16 counts
                        else:
                            ofmt = self._next_fmt(branch, pool_idx, n, Cout, True, f32_out)
                        st = branch_state[lif_idx]
                        v0, i0 = (None, None) if st is None else (st.v, st.i)
                        Y, v, i = ops.conv_lif_pool(Y, pc, T, cell._const, v0, i0, out=o, out_fmt=ofmt)
                        branch_state[lif_idx] = LIFFeedForwardState(v, i)
                        idx = pool_idx + 1
                        if is_last and o is not None:
                            return Y, False
                        continue
                if _im2col_route(layer, Y):
                    Y = ops.im2col_frames(Y, layer.stride[0])   # [N, 32, Ho, Wo] PAIR: the layer is now a 1x1 one
                    layer = _ColConv(layer)
                Yp = _packed_input(Y, layer)   # one conversion at most; later uses see the packed tensor
                if Yp is not None:
                    Y = Yp
                if (norm is not None and type(cell) is LIFCell and nograd and _conv_eligible(layer) and fuse_on):
                    lif_idx = last + 1
                    is_last = lif_idx == n - 1
                    res = residual if is_last else None
                    if is_last:
                        o = dst if residual is None else residual_dst
                    else:
                        o = self._lookahead_dst(branch, lif_idx, n, Y.shape[0], Cout, Ho, Wo, dst, Y.device, prealloc)
                    ofmt = ops.fmt_of(o) if o is not None else self._next_fmt(branch, lif_idx, n, Cout, res is None, f32_out)
                    st = branch_state[lif_idx]
                    v0, i0 = (None, None) if st is None else (st.v, st.i)
                    done = self._fused_conv_lif(layer, norm, cell, Y, T, v0, i0, res, o, ofmt, Ho, Wo)
                    if done is not None:
                        Y, v, i = done
                        branch_state[lif_idx] = LIFFeedForwardState(v, i)
                        idx = lif_idx + 1
                        if is_last:
                            return Y, residual is not None
                        continue
                    # unfused: fp32 currents, then the cell's own branch below scans them (+ residual, + store into the
                    # look-ahead destination chosen above)
                    Y = _run_conv(layer, norm, Y, None, ops.FMT_F32)
                    pending_dst = (lif_idx, o) if (o is not None and not is_last) else None
                    idx = lif_idx
                    continue
                is_last = last == n - 1
                nxt = branch[last + 1] if not is_last else None
                if (norm is None and nograd and _packed() and isinstance(nxt, BlockGen) and nxt._stem_plan() is not None
                        and layer.kernel_size == (1, 1) and layer.stride == (1, 1) and _conv_eligible(layer)
                        and nxt.net[0][0].in_channels == Cout and os.environ.get("SODA_FOLD_1X1", "1") != "0"):
                    # C2f entry convolution: its only consumers are the 1x1 stems of the next block, which absorb it
                    bidx = last + 1
                    if bidx == n - 1:
                        o = dst
                    else:
                        Hn, Wn = nxt.out_hw(Ho, Wo)
                        o = self._lookahead_dst(branch, bidx, n, Y.shape[0], nxt.out_channels, Hn, Wn, dst, Y.device,
                                                prealloc)
                    Y, branch_state[bidx] = nxt.forward_seq(Y, T, branch_state[bidx], out=o, fold=layer,
                                                            packed_out=not (f32_out and bidx == n - 1 and o is None))
                    idx = bidx + 1
                    if idx == n and o is not None:
                        return Y, False
                    continue
                if is_last:
                    o = dst
                else:
                    o = self._lookahead_dst(branch, last, n, Y.shape[0], Cout, Ho, Wo, dst, Y.device, prealloc)
                ofmt = ops.fmt_of(o) if o is not None else self._next_fmt(branch, last, n, Cout, False, f32_out)
                Y = _run_conv(layer, norm, Y, o, ofmt)
                idx = last + 1
                if is_last and o is not None:
                    return Y, False
                continue
            if isinstance(layer, BlockGen):
                if idx in prealloc:
                    o = prealloc.pop(idx)
                elif idx == n - 1:
                    o = dst
                else:
                    Ho, Wo = layer.out_hw(Y.shape[2], Y.shape[3])
                    o = self._lookahead_dst(branch, idx, n, Y.shape[0], layer.out_channels, Ho, Wo, dst, Y.device,
                                            prealloc)
                Y, branch_state[idx] = layer.forward_seq(Y, T, branch_state[idx], out=o,
                                                         packed_out=not (f32_out and idx == n - 1 and o is None),
                                                         lazy_cat=lazy_last and idx == n - 1 and o is None)
                idx += 1
                if idx == n and o is not None:
                    return Y, False
                continue
            if type(layer) is LIFCell and nograd:
                Yf = ops.to_f32(Y)
                if idx + 1 < n:
                    # Conv, Norm, LIF, Pool block: 2x2 / stride-2 pooling of the spikes fused into the scan's store
                    pool = branch[idx + 1]
                    mode = None
                    if isinstance(pool, SumPool2d) and pool.kernel_size == 2 and pool.stride == 2:
                        mode = "S"
                    elif isinstance(pool, SodaPool2d) and pool.kernel_size == 2 and pool.stride == 2:
                        mode = pool.mode
                    if mode is not None and Yf.is_contiguous(memory_format=torch.channels_last):
                        is_last = idx + 1 == n - 1
                        o = (dst if residual is None else None) if is_last else None
                        if (o is None or o.dtype == torch.float32) and ops.lif_pool_ok(Yf, o):
                            st = branch_state[idx]
                            v0, i0 = (None, None) if st is None else (st.v, st.i)
                            Y, v, i = ops.lif_pool_scan(Yf, T, mode, v0, i0, layer._const, out=o)
                            branch_state[idx] = LIFFeedForwardState(v, i)
                            idx += 2
                            if is_last and o is not None:
                                return Y, False
                            continue
                # the scan adds the Residual identity and stores straight into the destination (concat slice), in
                # the destination's / the consumer's format, in the same pass
                is_last = idx == n - 1
                res = residual if is_last else None
                pend, pending_dst = pending_dst, None
                if is_last:
                    o = dst if residual is None else residual_dst
                elif pend is not None and pend[0] == idx:
                    o = pend[1]
                else:
                    o = self._lookahead_dst(branch, idx, n, Yf.shape[0], Yf.shape[1], Yf.shape[2], Yf.shape[3], dst,
                                            Yf.device, prealloc)
                Yc = Yf if Yf.is_contiguous(memory_format=torch.channels_last) else None
                if Yc is not None and ops.lif_scan_fmt_ok(Yc, res, o):
                    ofmt = ops.fmt_of(o) if o is not None else self._next_fmt(branch, idx, n, Yf.shape[1], res is None, f32_out)
                    st = branch_state[idx]
                    v0, i0 = (None, None) if st is None else (st.v, st.i)
                    Y, v, i = ops.lif_scan_fmt(Yc, T, v0, i0, layer._const, residual=res, out=o, out_fmt=ofmt)
                    branch_state[idx] = LIFFeedForwardState(v, i)
                    idx += 1
                    if is_last and res is not None:
                        return Y, True
                    continue
                # the scan kernel does not take this layout: the generic cell path below produces a new tensor (a
                # look-ahead buffer stays valid - the Dense block's own [Pass] branch then copies into its slice)
                Y = Yf
            if isinstance(layer, (SNNCell, StateStorage)):
                Y = ops.to_f32(Y)
                fuse = (
                    isinstance(layer, LICell)
                    and idx + 1 < n
                    and isinstance(branch[idx + 1], nn.Tanh)
                    and nograd
                )
                if fuse:
                    Y, branch_state[idx] = layer.forward_seq(Y, T, branch_state[idx], fuse_tanh=True)
                    idx += 2
                else:
                    Y, branch_state[idx] = layer.forward_seq(Y, T, branch_state[idx])
                    idx += 1
                continue
            if isinstance(layer, SodaBatchNorm2d):
                Y = layer.forward_seq(ops.to_f32(Y), T)
                idx += 1
                continue
            if isinstance(layer, ConvLSTM):
                Y, branch_state[idx] = layer.forward_seq(ops.to_f32(Y), T, branch_state[idx])
                idx += 1
                continue
            if flags[idx]:
                # unknown stateful module: per-step protocol, one call per time step
                Y = ops.to_f32(Y)
                if T == 1:
                    Y, branch_state[idx] = layer(Y, branch_state[idx])
                else:
                    B = Y.shape[0] // T
                    ys, s = [], branch_state[idx]
                    for t in range(T):
                        y, s = layer(Y[t * B : (t + 1) * B], s)
                        ys.append(y)
                    Y, branch_state[idx] = torch.cat(ys), s
            else:
                if _wants_f32(layer):
                    Y = ops.to_f32(Y)
                b = ops.exact_bound(Y) if isinstance(layer, (nn.Identity, nn.Upsample)) else None
                Y = layer(Y)
                if b is not None and not (isinstance(layer, nn.Upsample) and layer.mode != "nearest"):
                    ops.mark_exact(Y, b)
            idx += 1
        if isinstance(Y, _CatParts):
            return Y, False
        if dst is not None and dst.data_ptr() != Y.data_ptr():
            _copy_into(Y, dst)
            Y = dst
        return Y, False


class ModelGen(nn.Module):
    """Base class of the model generators (reference models/generator.py:206-275)."""

    def __init__(self, cfg, in_channels: int = 2, init_weights: bool = True) -> None:
        super().__init__()
        self.out_channels = 0
        self.net_cfg = self._load_cfg(cfg)
        self._net_generator(in_channels)
        if init_weights:
            for m in self.modules():
                if isinstance(m, nn.Conv2d):
                    nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                    if m.bias is not None:
                        nn.init.constant_(m.bias, 0)
                elif isinstance(m, nn.BatchNorm2d):
                    nn.init.constant_(m.weight, 1)
                    if m.bias is not None:
                        nn.init.constant_(m.bias, 0)

    def _net_generator(self, in_channels: int) -> None:
        self.net = BlockGen(in_channels, self.net_cfg)
        self.out_channels = self.net.out_channels

    def _load_cfg(self, cfg) -> ListGen:
        raise NotImplementedError

    def forward(self, X: torch.Tensor, state: Optional[ListState] = None):
        raise NotImplementedError


class BackboneGen(ModelGen):
    """Returns the tensor from the last layer of the network."""

    def _load_cfg(self, cfg) -> ListGen:
        return cfg()

    def forward(self, X: torch.Tensor, state: Optional[ListState] = None) -> Tuple[torch.Tensor, ListState]:
        return self.net(X, state)

    def forward_seq(self, X: torch.Tensor, T: int, state: Optional[ListState] = None, packed_out: bool = False):
        return self.net.forward_seq(X, T, state, packed_out=packed_out)


class NeckGen(ModelGen):
    """Returns the list of tensors recorded by the ``Return`` taps, in module order."""

    def __init__(self, cfg, in_channels: int = 2, init_weights: bool = False):
        super().__init__(cfg, in_channels, init_weights)
        self.out_shape = self._search_out(self.net_cfg)

    def _search_out(self, cfg) -> List[int]:
        out: List[int] = []
        for module in cfg:
            if isinstance(module, Return):
                out.append(module.out_channels)
            elif isinstance(module, list):
                out += self._search_out(module)
        return out

    def _load_cfg(self, cfg) -> ListGen:
        return cfg()

    def _taps(self) -> List[torch.Tensor]:
        return [m.get_storage() for m in self.net.modules() if isinstance(m, Storage)]

    def forward(self, X: torch.Tensor, state: Optional[ListState] = None):
        _, state = self.net(X, state)
        return [ops.to_f32(t) for t in self._taps()], state

    def forward_seq(self, X: torch.Tensor, T: int, state: Optional[ListState] = None, packed_out: bool = False):
        _, state = self.net.forward_seq(X, T, state, packed_out=True)
        taps = self._taps()
        return (taps if packed_out else [ops.to_f32(t) for t in taps]), state


class Head(nn.Module):
    """One ``HeadGen`` + ``AnchorGenerator`` per pyramid level; flattens and concatenates the
    per-level predictions (reference models/generator.py:359-457)."""

    def __init__(self, cfg, num_classes: int, in_shape: List[int], init_weights: bool = True) -> None:
        super().__init__()
        self.num_classes = num_classes
        hi, lo, size_per_pix = 0.75, 0.08, 3
        sizes = torch.arange(lo, hi, (hi - lo) / (len(in_shape) * size_per_pix), dtype=torch.float32)
        sizes = sizes.reshape((-1, size_per_pix))
        ratios = torch.tensor((0.5, 1.0, 2), dtype=torch.float32)
        num_anchors = size_per_pix * len(ratios)
        num_class_out = num_anchors * (self.num_classes + 1)
        num_box_out = num_anchors * 4
        self.num_levels = len(in_shape)
        for idx, channels in enumerate(in_shape):
            setattr(self, f"anchor_gen_{idx}", AnchorGenerator(sizes=sizes[idx], ratios=ratios))
            setattr(self, f"model_{idx}", HeadGen(cfg, num_box_out, num_class_out, channels, init_weights))

    def forward(self, X: List[torch.Tensor], state: Optional[ListState] = None):
        return self.forward_seq(X, 1, state)

    def forward_seq(self, X: List[torch.Tensor], T: int, state: Optional[ListState] = None):
        """Predictions of the LAST of the T steps (the only ones the detector observes,
        models/soda.py:141-144) + the head state after all T steps."""
        state = [None] * len(X) if state is None else list(state)
        anchors, cls_preds, bbox_preds = [], [], []
        for idx, fmap in enumerate(X):
            anchors.append(getattr(self, f"anchor_gen_{idx}")(fmap))
            boxes, classes, state[idx] = getattr(self, f"model_{idx}").forward_seq(fmap, T, state[idx])
            bbox_preds.append(boxes)
            cls_preds.append(classes)
        anchors = torch.cat(anchors)
        cls_preds = self._concat_preds(cls_preds)
        cls_preds = cls_preds.reshape(cls_preds.shape[0], -1, self.num_classes + 1)
        bbox_preds = self._concat_preds(bbox_preds)
        bbox_preds = bbox_preds.reshape(bbox_preds.shape[0], -1, 4)
        return anchors, cls_preds, bbox_preds, state

    def _flatten_pred(self, pred: torch.Tensor) -> torch.Tensor:
        # pixel-major storage makes permute(0,2,3,1) + flatten a view, not a copy
        return torch.flatten(torch.permute(pred, (0, 2, 3, 1)), start_dim=1)

    def _concat_preds(self, preds: List[torch.Tensor]) -> torch.Tensor:
        return torch.cat([self._flatten_pred(p) for p in preds], dim=1)


class HeadGen(ModelGen):
    """Head for one pyramid level: ``[prepare, box, cls]`` lists (reference generator.py:465-538)."""

    def __init__(self, cfg, box_out: int, cls_out: int, in_channels: int = 2, init_weights=False):
        self.box_out = box_out
        self.cls_out = cls_out
        super().__init__(cfg, in_channels, init_weights)

    def _net_generator(self, in_channels: int) -> None:
        self.base_net = BlockGen(in_channels, [self.net_cfg[0]])
        self.box_net = BlockGen(self.base_net.out_channels, [self.net_cfg[1]])
        self.cls_net = BlockGen(self.base_net.out_channels, [self.net_cfg[2]])

    def _load_cfg(self, cfg) -> ListGen:
        return cfg(self.box_out, self.cls_out)

    def forward(self, X: torch.Tensor, state: Optional[ListState] = None):
        return self.forward_seq(X, 1, state)

    def forward_seq(self, X: torch.Tensor, T: int, state: Optional[ListState] = None):
        state = [None] * 3 if state is None else list(state)
        Y, state[0] = self.base_net.forward_seq(X, T, state[0])   # X: any activation format (a neck tap)
        B = Y.shape[0] // T
        outs = []
        for k, net in ((1, self.box_net), (2, self.cls_net)):
            trains_bn = self.training and any(isinstance(m, nn.modules.batchnorm._BatchNorm) for m in net.modules())
            if T > 1 and not _has_state(net) and not trains_bn:
                # stateless prediction nets: only the last step is observable (a train-mode BatchNorm inside would
                # have to see all T steps to update its running statistics as the reference does)
                o, state[k] = net.forward_seq(Y[(T - 1) * B :], 1, state[k])
            else:
                o, state[k] = net.forward_seq(Y, T, state[k])
                o = o[(T - 1) * B :]
            outs.append(o)
        return outs[0], outs[1], state
