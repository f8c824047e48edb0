"""Kernel-level breakdown of one BPTT training step (torch.profiler, CUDA activities): which kernels the 63 ms go to.
  python scripts/train_profile.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import snn_for_object_detection_b200 as pkg

torch.manual_seed(0)
dev = torch.device("cuda")
model = pkg.TinyYolo(num_classes=2).to(dev).train()
for m in model.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.weight.data.fill_(3.0)
opt = model.configure_optimizers()
T, B, H, W = 32, 5, 240, 304
X = (torch.rand(T, B, 2, H, W, device=dev) < 0.07).float()
labels = torch.full((B, 10, 5), -1.0, device=dev)
ctr, wh = torch.rand(B, 6, 2, device=dev) * 0.6 + 0.2, torch.rand(B, 6, 2, device=dev) * 0.25 + 0.05
labels[:, :6, 0] = torch.randint(0, 2, (B, 6), device=dev).float()
labels[:, :6, 1:3], labels[:, :6, 3:5] = ctr - wh / 2, ctr + wh / 2

def step():
    opt.zero_grad(set_to_none=True)
    loss = model._loss(model(X), labels)
    loss.backward()
    opt.step()

for _ in range(2):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in rows)
print(f"total device time {tot / 1e3:.2f} ms")
for e in rows[:28]:
    print(f"{e.device_time_total / 1e3:8.3f} ms {100 * e.device_time_total / tot:5.1f} %  x{e.count:<5d} {e.key[:110]}")
