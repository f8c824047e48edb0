"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED.

CPU restatement of the parts of ``norse==1.1.0`` (third-party; pinned in the reference's
``environment.yml:222``; NOT present in ``/root/reference`` and not installable here) that the
reference's hot path calls:

* ``norse.torch.LIFCell()``   - reference call site ``models/modules/layer_gen.py:233``
* ``norse.torch.LICell()``    - reference call site ``models/modules/layer_gen.py:253``
* ``norse.torch.module.snn.SNNCell`` / ``SNN`` / ``_merge_states``
                              - ``models/modules/common.py:9,105``, ``sli.py:7``, ``synapse.py:7``
* ``norse.torch.utils.state._is_module_stateful``
                              - ``models/generator.py:21,142``
* ``SuperSpike`` surrogate gradient (norse ``functional/superspike.py``), alpha = 100

Published algorithm restated (norse 1.1.0 ``functional/lif.py::lif_feed_forward_step`` and
``functional/leaky_integrator.py::li_feed_forward_step``), all in the tensor dtype (fp32)::

    i_new     = i + x                                   # current jump first
    dv        = (dt * tau_mem_inv) * ((v_leak - v) + i_new)
    v_decayed = v + dv
    di        = (-dt * tau_syn_inv) * i_new
    i_decayed = i_new + di
    z         = heaviside(v_decayed - v_th)             # strict: x > 0 ;  LI stops before this
    v_new     = (1 - z) * v_decayed + z * v_reset
    return z, (v_new, i_decayed)                        # LI returns v_decayed, (v_decayed, i_decayed)

with dt = 1e-3, tau_syn_inv = 200, tau_mem_inv = 100, v_leak = 0, v_th = 1, v_reset = 0.
The in-tree evidence for this statement order is the reference author's own cell derived from
norse's LI cell, ``models/modules/sli.py:110-126``.  Because norse's source is not on this box the
order is a switch (``order="jump_first"`` default, ``"decay_first"`` = norse <= 0.0.6) and so is the
reset detachment in backward (``detach_reset``), mirrored 1:1 by flags of the CUDA kernels.

The scalar products ``dt * tau_mem_inv`` / ``-dt * tau_syn_inv`` are evaluated the way torch does
for ``python_float * fp32_tensor``: in fp32, giving exactly 0.1f and -0.2f.
"""

from __future__ import annotations

import inspect
from typing import Any, List, NamedTuple, Optional, Tuple

import torch

__all__ = [
    "LIFParameters",
    "LIParameters",
    "LIFFeedForwardState",
    "LIState",
    "SuperSpike",
    "heaviside",
    "threshold",
    "lif_feed_forward_step",
    "li_feed_forward_step",
    "SNNCell",
    "SNN",
    "LIFCell",
    "LICell",
    "_merge_states",
    "_is_module_stateful",
]

ORDER = "jump_first"  # module-level switch read by the step functions (see header)
DETACH_RESET = False


class LIFParameters(NamedTuple):
    tau_syn_inv: torch.Tensor = torch.as_tensor(1.0 / 5e-3)
    tau_mem_inv: torch.Tensor = torch.as_tensor(1.0 / 1e-2)
    v_leak: torch.Tensor = torch.as_tensor(0.0)
    v_th: torch.Tensor = torch.as_tensor(1.0)
    v_reset: torch.Tensor = torch.as_tensor(0.0)
    method: str = "super"
    alpha: float = torch.as_tensor(100.0)


class LIParameters(NamedTuple):
    tau_syn_inv: torch.Tensor = torch.as_tensor(1.0 / 5e-3)
    tau_mem_inv: torch.Tensor = torch.as_tensor(1.0 / 1e-2)
    v_leak: torch.Tensor = torch.as_tensor(0.0)


class LIFFeedForwardState(NamedTuple):
    v: torch.Tensor
    i: torch.Tensor


class LIState(NamedTuple):
    v: torch.Tensor
    i: torch.Tensor


def heaviside(x: torch.Tensor) -> torch.Tensor:
    """norse ``functional/heaviside.py``: strict ``x > 0`` in the input dtype."""
    return torch.gt(x, torch.as_tensor(0.0)).to(x.dtype)


class SuperSpike(torch.autograd.Function):
    """norse ``functional/superspike.py``: forward heaviside, backward g / (alpha*|x| + 1)^2."""

    @staticmethod
    def forward(ctx, input_tensor: torch.Tensor, alpha: float) -> torch.Tensor:
        ctx.save_for_backward(input_tensor)
        ctx.alpha = alpha
        return heaviside(input_tensor)

    @staticmethod
    def backward(ctx, grad_output):
        (inp,) = ctx.saved_tensors
        alpha = ctx.alpha
        grad = grad_output / (alpha * torch.abs(inp) + 1.0).pow(2)
        return grad, None


def threshold(x: torch.Tensor, method: str, alpha) -> torch.Tensor:
    if method == "heaviside":
        return heaviside(x)
    if method == "super":
        return SuperSpike.apply(x, float(alpha))
    raise ValueError(f"unsupported threshold method {method!r} in the oracle")


def lif_feed_forward_step(
    input_spikes: torch.Tensor,
    state: LIFFeedForwardState,
    p: LIFParameters = LIFParameters(),
    dt: float = 0.001,
) -> Tuple[torch.Tensor, LIFFeedForwardState]:
    if ORDER == "jump_first":
        i_new = state.i + input_spikes
        dv = dt * p.tau_mem_inv * ((p.v_leak - state.v) + i_new)
        v_decayed = state.v + dv
        di = -dt * p.tau_syn_inv * i_new
        i_out = i_new + di
    else:  # decay_first (norse <= 0.0.6): integrate old current, jump at the end
        dv = dt * p.tau_mem_inv * ((p.v_leak - state.v) + state.i)
        v_decayed = state.v + dv
        di = -dt * p.tau_syn_inv * state.i
        i_out = state.i + di + input_spikes
    z_new = threshold(v_decayed - p.v_th, p.method, p.alpha)
    zr = z_new.detach() if DETACH_RESET else z_new
    v_new = (1 - zr) * v_decayed + zr * p.v_reset
    return z_new, LIFFeedForwardState(v=v_new, i=i_out)


def li_feed_forward_step(
    input_tensor: torch.Tensor,
    state: LIState,
    p: LIParameters = LIParameters(),
    dt: float = 0.001,
) -> Tuple[torch.Tensor, LIState]:
    if ORDER == "jump_first":
        i_new = state.i + input_tensor
        dv = dt * p.tau_mem_inv * ((p.v_leak - state.v) + i_new)
        v_new = state.v + dv
        di = -dt * p.tau_syn_inv * i_new
        i_out = i_new + di
    else:
        dv = dt * p.tau_mem_inv * ((p.v_leak - state.v) + state.i)
        v_new = state.v + dv
        di = -dt * p.tau_syn_inv * state.i
        i_out = state.i + di + input_tensor
    return v_new, LIState(v_new, i_out)


class SNNCell(torch.nn.Module):
    """norse ``module/snn.py::SNNCell``: ``forward(x, state=None) -> (out, new_state)``."""

    def __init__(self, activation, state_fallback, p: Any, dt: float = 0.001, **kwargs):
        super().__init__()
        self.activation = activation
        self.state_fallback = state_fallback
        self.p = p
        self.dt = dt

    def forward(self, input_tensor: torch.Tensor, state: Optional[Any] = None):
        state = state if state is not None else self.state_fallback(input_tensor)
        return self.activation(input_tensor, state, self.p, self.dt)


class SNN(torch.nn.Module):
    """norse ``module/snn.py::SNN``: iterates a cell over the outer (time) dimension."""

    def __init__(self, activation, state_fallback, p: Any, dt: float = 0.001, **kwargs):
        super().__init__()
        self.activation = activation
        self.state_fallback = state_fallback
        self.p = p
        self.dt = dt

    def forward(self, input_tensor: torch.Tensor, state: Optional[Any] = None):
        state = state if state is not None else self.state_fallback(input_tensor)
        outputs = []
        for ts in range(input_tensor.shape[0]):
            out, state = self.activation(input_tensor[ts], state, self.p, self.dt)
            outputs.append(out)
        return torch.stack(outputs), state


class LIFCell(SNNCell):
    def __init__(self, p: LIFParameters = LIFParameters(), **kwargs):
        super().__init__(
            activation=lif_feed_forward_step,
            state_fallback=self.initial_state,
            p=p,
            **kwargs,
        )

    def initial_state(self, input_tensor: torch.Tensor) -> LIFFeedForwardState:
        state = LIFFeedForwardState(
            v=torch.full(
                input_tensor.shape,
                float(torch.as_tensor(self.p.v_leak).detach()),
                device=input_tensor.device,
                dtype=torch.float32,
            ),
            i=torch.zeros(
                *input_tensor.shape, device=input_tensor.device, dtype=torch.float32
            ),
        )
        state.v.requires_grad = True
        return state


class LICell(SNNCell):
    def __init__(self, p: LIParameters = LIParameters(), **kwargs):
        super().__init__(
            activation=li_feed_forward_step,
            state_fallback=self.initial_state,
            p=p,
            **kwargs,
        )

    def initial_state(self, input_tensor: torch.Tensor) -> LIState:
        state = LIState(
            v=torch.as_tensor(self.p.v_leak).detach().clone(),
            i=torch.zeros(
                *input_tensor.shape, device=input_tensor.device, dtype=input_tensor.dtype
            ),
        )
        state.v.requires_grad = True
        return state


def _merge_states(states: List[Any]):
    """norse ``module/snn.py::_merge_states``: stack every field of a list of state tuples."""
    state_dict = states[0]._asdict()
    cls = states[0].__class__
    keys = list(state_dict.keys())
    output_dict = {}
    for key in keys:
        if isinstance(state_dict[key], tuple):
            output_dict[key] = _merge_states([getattr(s, key) for s in states])
        else:
            output_dict[key] = torch.stack([getattr(s, key) for s in states])
    return cls(**output_dict)


def _is_module_stateful(module: torch.nn.Module) -> bool:
    """norse ``utils/state.py``: a module is stateful iff ``forward`` takes a ``state`` argument."""
    signature = inspect.signature(module.forward)
    if "state" in signature.parameters or isinstance(module, torch.nn.RNNBase):
        return True
    return False
