"""Detection decode + NMS: soda_multibox_detection (one CTA per sample) vs the torch restatement of the reference's
utils/box.py:multibox_detection on the GPU (Python while-loop, one sync per kept box) and on the host cores.
  python scripts/nms_probe.py [B] [A]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import snn_for_object_detection_b200 as pkg
from snn_for_object_detection_b200.utils import box

B = int(sys.argv[1]) if len(sys.argv) > 1 else 5
A = int(sys.argv[2]) if len(sys.argv) > 2 else 13545
torch.manual_seed(0)
ctr, wh = torch.rand(A, 2), torch.rand(A, 2) * 0.2 + 0.02
anchors = torch.cat([ctr - wh / 2, ctr + wh / 2], dim=1)
probs = torch.softmax(torch.randn(B, A, 3) * 2.0, dim=2)
offsets = torch.randn(B, A, 4) * 0.5
dp, do, da = probs.cuda(), offsets.cuda(), anchors.cuda()
for _ in range(3):
    det = pkg.ops.multibox_detection(dp, do, da)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    det = pkg.ops.multibox_detection(dp, do, da)
e1.record()
torch.cuda.synchronize()
ms_k = e0.elapsed_time(e1) / 20
t0 = time.perf_counter(); ref_g = box.multibox_detection(dp, do, da); torch.cuda.synchronize(); ms_g = (time.perf_counter() - t0) * 1e3
t0 = time.perf_counter(); ref_c = box.multibox_detection(probs, offsets, anchors); ms_c = (time.perf_counter() - t0) * 1e3
same = bool(torch.equal(det.cpu()[..., :2], ref_c[..., :2]))
print(f"B={B} A={A}: kept {int((ref_c[..., 0] >= 0).sum())} boxes; kernel {ms_k:.3f} ms | torch loop on GPU {ms_g:.1f} ms | "
      f"torch loop on CPU ({torch.get_num_threads()} threads) {ms_c:.1f} ms | rows identical: {same}")
