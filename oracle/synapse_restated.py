"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's synaptic transmission
model, ``models/modules/synapse.py:73-103`` (``synapse_feed_forward_step``) with the cell's initial state ``:58-70``
(p = 0).  The arithmetic lives in the reference tree, so it is pinned: ``tests/golden/synapse.npz`` holds outputs of
the reference's own function (``oracle/make_golden.py:golden_synapse``).  torch-CPU fp32, same expression order."""
import torch


def synapse_scan(x: torch.Tensor, p0=None, tau_med_secretion=1000.0, tau_med_dissociation=200.0, sigma_inhibition=0.0,
                 dt=0.001):
    """x [T, ...] -> (g per step [T, ...], p_T)."""
    ts, td = torch.as_tensor(tau_med_secretion), torch.as_tensor(tau_med_dissociation)
    sg = torch.as_tensor(sigma_inhibition)
    p = torch.zeros_like(x[0]) if p0 is None else p0
    outs = []
    for t in range(x.shape[0]):
        tau = torch.where(x[t] > 0, ts, td)
        p = p + (x[t] - p) * tau * dt
        g = 4 * sg * (p - sg * p.square()) if float(sg) != 0.0 else p
        outs.append(g.clamp(0.0))
    return torch.stack(outs), p
