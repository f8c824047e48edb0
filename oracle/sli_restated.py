"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's saturable leaky
integrator, ``models/modules/sli.py:110-126`` (``sli_feed_forward_step``) with the cell's initial state ``:97-107``
(v = v_leak, i = 0).  Unlike the norse cells this arithmetic lives IN the reference tree, so it is pinned:
``tests/golden/sli.npz`` holds outputs of the reference's own function (``oracle/make_golden.py:golden_sli``).
torch-CPU fp32, same expression order."""
import torch


def sli_scan(x: torch.Tensor, v0=None, i0=None, tau_syn_inv=200.0, tau_mem_inv=100.0, v_leak=0.0, v_st=1.0, dt=0.001):
    """x [T, ...] -> (v per step [T, ...], v_T, i_T)."""
    tsi, tmi = torch.as_tensor(tau_syn_inv), torch.as_tensor(tau_mem_inv)
    vl, vs = torch.as_tensor(v_leak), torch.as_tensor(v_st)
    v = vl if v0 is None else v0
    i = torch.zeros_like(x[0]) if i0 is None else i0
    outs = []
    for t in range(x.shape[0]):
        i_jump = i + x[t] * torch.sigmoid(vs - torch.abs(v))
        dv = dt * tmi * ((vl - v) + i_jump)
        v = v + dv
        di = -dt * tsi * i_jump
        i = i_jump + di
        outs.append(v)
    return torch.stack(outs), v, i
