import sys, torch
sys.path.insert(0, "/root/repo")
from snn_for_object_detection_b200 import ops
x = (torch.rand(160, 2, 240, 304, device="cuda") < 0.07).float()
for _ in range(3): ops.im2col_frames(x, 2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): ops.im2col_frames(x, 2)
e1.record(); torch.cuda.synchronize()
print("im2col ms", e0.elapsed_time(e1) / 10)
