"""Per-layer spike / membrane mismatch diagnostics (GPU path vs CPU oracle)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from oracle import net_restated as onr
from snn_for_object_detection_b200 import TinyYolo
from snn_for_object_detection_b200.modules.common import StateStorage

torch.manual_seed(5)
model = TinyYolo(num_classes=2, state_storage=True)
T, B, H, W = 6, 2, 64, 96
X = (torch.rand(T, B, 2, H, W) < 0.08).float()
model.eval(); onr.calibrate(model, X, gain=5.0); model.eval()
taps = {}
onr.soda_forward(model, X, taps)
cells = [(n, m) for n, m in model.named_modules() if isinstance(m, StateStorage)]
ref = {n: taps[id(m)] for n, m in cells}
model = model.to("cuda")
with torch.no_grad():
    model(X.cuda())
for n, m in cells:
    outs = torch.stack([o for o, _ in ref[n]])
    v_ref = torch.stack([s.v for _, s in ref[n]])
    got = m.get_spikes().cpu(); st = m.get_state(); v = st.v.cpu()
    if type(m.module).__name__ != "LIFCell":
        print(f"{n:60s} LI  vrel {float((v - v_ref).abs().max() / v_ref.abs().max()):.2e}")
        continue
    flips = got != outs
    nf = int(flips.sum())
    dv = (v - v_ref).abs()
    # neurons that never flipped: pure numerical membrane error
    never = ~flips.any(dim=0, keepdim=True).expand_as(flips)
    clean = float(dv[never].max()) if never.any() else 0.0
    # distance from threshold on the non-spiking side of each flip
    side = torch.where(outs == 0, v_ref, v)[flips]
    dist = float((side - 1).abs().max()) if nf else 0.0
    print(f"{n:60s} LIF flips {nf:5d}/{flips.numel()} ({nf / flips.numel():.1e}) |v|max {float(v_ref.abs().max()):.2f} "
          f"clean-dv-max {clean:.2e} bad>1e-4*max {float((dv > 1e-4 * v_ref.abs().max()).float().mean()):.1e} flip-dist-max {dist:.2e}")
