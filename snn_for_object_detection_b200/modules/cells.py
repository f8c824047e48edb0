"""LIF / LI neuron cells backed by the fused CUDA scans.

These take the place of ``norse.torch.LIFCell()`` / ``norse.torch.LICell()`` as the reference
instantiates them (models/modules/layer_gen.py:233,253): same call protocol
``forward(x, state=None) -> (out, new_state)``, same state tuples with ``.v`` / ``.i`` fields, same
default parameters, same initial state (v = v_leak, i = 0).  On top of the per-step protocol each
cell has ``forward_seq(x, T, state)`` which advances T time steps in ONE kernel with the state
held in registers (layer-major execution).
"""

from __future__ import annotations

from typing import NamedTuple, Optional, Tuple

import torch
from torch import nn

from .. import ops
from ..ops import NeuronConst

__all__ = ["LIFParameters", "LIParameters", "LIFFeedForwardState", "LIState", "SNNCell", "LIFCell", "LICell",
           "SLIParameters", "SLIState", "SLICell", "SynapseParameters", "SynapseState", "SynapseCell"]


class LIFParameters(NamedTuple):
    tau_syn_inv: float = 1.0 / 5e-3
    tau_mem_inv: float = 1.0 / 1e-2
    v_leak: float = 0.0
    v_th: float = 1.0
    v_reset: float = 0.0
    method: str = "super"
    alpha: float = 100.0


class LIParameters(NamedTuple):
    tau_syn_inv: float = 1.0 / 5e-3
    tau_mem_inv: float = 1.0 / 1e-2
    v_leak: float = 0.0


class SLIParameters(NamedTuple):
    """Parameters of the saturable leaky integrator (reference ``models/modules/sli.py:25-38``)."""

    tau_syn_inv: float = 1.0 / 5e-3
    tau_mem_inv: float = 1.0 / 1e-2
    v_leak: float = 0.0
    v_st: float = 1.0


class SLIState(NamedTuple):
    v: torch.Tensor
    i: torch.Tensor


class SynapseParameters(NamedTuple):
    """Parameters of the synapse model (reference ``models/modules/synapse.py:22-33``)."""

    tau_med_secretion: float = 1.0 / 1e-3
    tau_med_dissociation: float = 1.0 / 5e-3
    sigma_inhibition: float = 0.0


class SynapseState(NamedTuple):
    p: torch.Tensor


class LIFFeedForwardState(NamedTuple):
    v: torch.Tensor
    i: torch.Tensor


class LIState(NamedTuple):
    v: torch.Tensor
    i: torch.Tensor


def _f32(x: float) -> float:
    return float(torch.tensor(x, dtype=torch.float32))


def _const(p, dt: float, jump_first: bool = True, detach_reset: bool = False) -> NeuronConst:
    # torch evaluates ``dt * tau_inv_tensor`` in fp32: fp32(dt) * fp32(tau_inv), rounded to fp32
    a = float(torch.tensor(dt, dtype=torch.float32) * torch.tensor(p.tau_mem_inv, dtype=torch.float32))
    s = float(torch.tensor(dt, dtype=torch.float32) * torch.tensor(p.tau_syn_inv, dtype=torch.float32))
    return NeuronConst(
        dt_tau_mem=a,
        dt_tau_syn=s,
        v_leak=_f32(p.v_leak),
        v_th=_f32(getattr(p, "v_th", 1.0)),
        v_reset=_f32(getattr(p, "v_reset", 0.0)),
        alpha=_f32(getattr(p, "alpha", 100.0)),
        jump_first=int(jump_first),
        detach_reset=int(detach_reset),
    )


class _LIFScanFn(torch.autograd.Function):
    """Forward scan + fused surrogate-gradient reverse scan (soda_lif_fwd / soda_lif_bwd)."""

    @staticmethod
    def forward(ctx, x, v0, i0, T, const):
        z, v, i, vd, _, _ = ops.lif_scan(x, T, v0, i0, const, want_v_dec=True)
        ctx.save_for_backward(vd)
        ctx.T, ctx.const = T, const
        ctx.has_state = v0 is not None
        return z, v, i

    @staticmethod
    def backward(ctx, gz, gv, gi):
        (vd,) = ctx.saved_tensors
        if gz is None:
            gz = torch.zeros_like(vd)
        gx, gv0, gi0 = ops.lif_scan_bwd(gz, gv, gi, vd, ctx.T, ctx.const)  # re-laid out like v_dec if needed
        if not ctx.has_state:
            gv0 = gi0 = None
        return gx, gv0, gi0, None, None


class _LIScanFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, v0, i0, T, const):
        out, v, i, _ = ops.li_scan(x, T, v0, i0, const, apply_tanh=False)
        ctx.T, ctx.const = T, const
        ctx.has_state = i0 is not None
        return out, v, i

    @staticmethod
    def backward(ctx, gout, gv, gi):
        if gout is None:
            gout = torch.zeros_like(gv).repeat(ctx.T, *([1] * (gv.dim() - 1)))
        gx, gv0, gi0 = ops.li_scan_bwd(gout, gv, gi, ctx.T, ctx.const)
        if not ctx.has_state:
            gv0 = gi0 = None
        return gx, gv0, gi0, None, None


class SNNCell(nn.Module):
    """Base of the stateful neuron cells (role of ``norse.torch.module.snn.SNNCell``)."""

    def __init__(self, p, dt: float = 0.001, jump_first: bool = True, detach_reset: bool = False):
        super().__init__()
        self.p = p
        self.dt = dt
        self.jump_first = jump_first
        self.detach_reset = detach_reset
        self._const = _const(p, dt, jump_first, detach_reset)

    def forward(self, input_tensor: torch.Tensor, state=None):
        return self.forward_seq(input_tensor, 1, state)

    def forward_seq(self, x: torch.Tensor, T: int, state=None, want_trace: bool = False):
        raise NotImplementedError


class LIFCell(SNNCell):
    """Leaky integrate-and-fire cell; returns ``(spikes, LIFFeedForwardState(v, i))``."""

    def __init__(self, p: LIFParameters = LIFParameters(), **kwargs):
        super().__init__(p, **kwargs)

    def forward_seq(self, x, T, state=None, want_trace=False):
        v0, i0 = (None, None) if state is None else (state.v, state.i)
        if torch.is_grad_enabled() and (x.requires_grad or (v0 is not None and v0.requires_grad)):
            z, v, i = _LIFScanFn.apply(x, v0, i0, T, self._const)
            return z, LIFFeedForwardState(v, i)
        z, v, i, _, vt, it = ops.lif_scan(x, T, v0, i0, self._const, want_trace=want_trace)
        if want_trace:
            return z, LIFFeedForwardState(v, i), LIFFeedForwardState(vt, it)
        return z, LIFFeedForwardState(v, i)


class LICell(SNNCell):
    """Leaky integrator cell; returns ``(v, LIState(v, i))``.  ``fuse_tanh`` folds a following Tanh."""

    def __init__(self, p: LIParameters = LIParameters(), **kwargs):
        super().__init__(p, **kwargs)

    def forward_seq(self, x, T, state=None, want_trace=False, fuse_tanh=False):
        v0, i0 = (None, None) if state is None else (state.v, state.i)
        if torch.is_grad_enabled() and (x.requires_grad or (i0 is not None and i0.requires_grad)):
            out, v, i = _LIScanFn.apply(x, v0, i0, T, self._const)
            if fuse_tanh:
                out = torch.tanh(out)
            return out, LIState(v, i)
        out, v, i, it = ops.li_scan(x, T, v0, i0, self._const, apply_tanh=fuse_tanh, want_trace=want_trace)
        if want_trace:
            # per-step v is the (pre-tanh) output itself
            return out, LIState(v, i), LIState(out if not fuse_tanh else None, it)
        return out, LIState(v, i)


class SLICell(SNNCell):
    """Saturable leaky integrator cell of the reference (``models/modules/sli.py:80-126``); returns
    ``(v, SLIState(v, i))``.  Inference: one CUDA scan over all T steps (``soda_sli_fwd``); when a graph is being
    recorded: the reference's own expression step by step in torch ops (differentiable)."""

    def __init__(self, p: SLIParameters = SLIParameters(), **kwargs):
        super().__init__(p, **kwargs)
        self._const = self._const._replace(v_th=_f32(p.v_st))

    def forward_seq(self, x, T, state=None, want_trace=False):
        v0, i0 = (None, None) if state is None else (state.v, state.i)
        if torch.is_grad_enabled() and (x.requires_grad or (i0 is not None and i0.requires_grad)):
            return self._forward_torch(x, T, v0, i0)
        out, v, i, it = ops.sli_scan(x, T, v0, i0, self._const, want_trace=want_trace)
        if want_trace:
            return out, SLIState(v, i), SLIState(out, it)
        return out, SLIState(v, i)

    def _forward_torch(self, x, T, v0, i0):
        B = x.shape[0] // T
        c = self._const
        v = torch.full_like(x[:B], c.v_leak) if v0 is None or v0.dim() == 0 else v0
        i = torch.zeros_like(x[:B]) if i0 is None else i0
        outs = []
        for t in range(T):
            i_jump = i + x[t * B : (t + 1) * B] * torch.sigmoid(c.v_th - torch.abs(v))
            v = v + c.dt_tau_mem * ((c.v_leak - v) + i_jump)
            i = i_jump + (-c.dt_tau_syn) * i_jump
            outs.append(v)
        return (outs[0] if T == 1 else torch.cat(outs)), SLIState(v, i)


class SynapseCell(SNNCell):
    """Synaptic transmission cell of the reference (``models/modules/synapse.py:36-103``); returns
    ``(g, SynapseState(p))``.  Inference: one CUDA scan over all T steps (``soda_synapse_fwd``); when a graph is being
    recorded: the reference's own expression step by step in torch ops (differentiable)."""

    def __init__(self, p: SynapseParameters = SynapseParameters(), dt: float = 0.001, **kwargs):
        nn.Module.__init__(self)
        if p.sigma_inhibition != 0 and p.sigma_inhibition < 0.5:
            raise ValueError("Valid values for sigma_inhibition are 0 or >= 0.5, but received ", p.sigma_inhibition)
        self.p, self.dt = p, dt
        self._c = tuple(_f32(v) for v in (p.tau_med_secretion, p.tau_med_dissociation, p.sigma_inhibition, dt))

    def forward_seq(self, x, T, state=None, want_trace=False):
        p0 = None if state is None else state.p
        if torch.is_grad_enabled() and (x.requires_grad or (p0 is not None and p0.requires_grad)):
            return self._forward_torch(x, T, p0)
        out, p = ops.synapse_scan(x, T, p0, *self._c)
        return out, SynapseState(p)

    def _forward_torch(self, x, T, p0):
        B = x.shape[0] // T
        ts, td, sg, dt = self._c
        p = torch.zeros_like(x[:B]) if p0 is None else p0
        outs = []
        for t in range(T):
            xt = x[t * B : (t + 1) * B]
            tau = torch.where(xt > 0, torch.full_like(xt, ts), torch.full_like(xt, td))
            p = p + (xt - p) * tau * dt
            g = 4 * sg * (p - sg * p.square()) if sg != 0 else p
            outs.append(g.clamp(0.0))
        return (outs[0] if T == 1 else torch.cat(outs)), SynapseState(p)
