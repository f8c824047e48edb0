"""Anchor labelling: soda_roi_assign (one CTA per sample) vs the CPU restatement of the reference's utils/roi.py
(oracle/roi_restated.py, checker and reported CPU baseline only).
  python scripts/roi_probe.py [B] [A] [G]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import snn_for_object_detection_b200 as pkg
from oracle import roi_restated as rr

B = int(sys.argv[1]) if len(sys.argv) > 1 else 5
A = int(sys.argv[2]) if len(sys.argv) > 2 else 13545
G = int(sys.argv[3]) if len(sys.argv) > 3 else 12
rng = np.random.default_rng(0)
ctr = rng.random((A, 2), dtype=np.float32)
wh = rng.random((A, 2), dtype=np.float32) * 0.3 + 0.03
anchors = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=1).astype(np.float32)
labels = np.full((B, G, 5), -1.0, np.float32)
for b in range(B):
    for k in range(max(1, G * (b + 1) // (B + 1))):
        labels[b, k, 0] = k % 2
        labels[b, k, 1:] = anchors[rng.integers(A)] + (rng.random(4, dtype=np.float32) - 0.5) * 0.05
da, dl = torch.from_numpy(anchors).cuda(), torch.from_numpy(labels).cuda()
for _ in range(3):
    out = pkg.ops.roi_assign(da, dl, 0.4)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    out = pkg.ops.roi_assign(da, dl, 0.4)
e1.record()
torch.cuda.synchronize()
ms_k = e0.elapsed_time(e1) / 20
t0 = time.perf_counter(); ref = rr.roi(anchors, labels, 0.4); ms_c = (time.perf_counter() - t0) * 1e3
same = bool(np.array_equal(out[2].cpu().numpy(), ref[2]) and np.array_equal(out[1].cpu().numpy(), ref[1]))
print(f"B={B} A={A} G={G}: {int((ref[2] > 0).sum())} object anchors; kernel {ms_k:.3f} ms | CPU restatement (numpy, 1 thread) "
      f"{ms_c:.1f} ms | map identical: {same}; IoU evaluations {B * A * G * (G + 1) / 1e6:.1f} M")
