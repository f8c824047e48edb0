"""BPTT training step of TinyYolo (BASELINE config 4 shape per GPU: GEN1 304x240, B=5, T=32): forward in train mode
(our conv / LIF / LI kernels, batch-statistics BatchNorm per time step), backward (our fused LIF / LI reverse scans,
cuDNN dgrad / wgrad), flat gradient all-reduce over NCCL when launched under torchrun, Adamax step
(models/soda.py:135-136).  The loss is the reference's detection loss (models/soda.py:259-281) on synthetic labels
(6 random boxes per sample padded to 10 rows with -1): anchor labelling by soda_roi_assign, cross entropy + masked L1.

  python scripts/train_step_probe.py [--steps 3] [--T 32] [--batch 5] [--hw 240 304]
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/train_step_probe.py
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import snn_for_object_detection_b200 as pkg
from snn_for_object_detection_b200 import sharding

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--T", type=int, default=32)
ap.add_argument("--batch", type=int, default=5)
ap.add_argument("--hw", type=int, nargs=2, default=[240, 304])
ap.add_argument("--profile-last", action="store_true", help="cudaProfilerStart/Stop around the last step (ncu --profile-from-start off)")
ap.add_argument("--sync-bn", action="store_true", help="SyncBatchNorm: per-step statistics over the global batch")
args = ap.parse_args()

rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
model = pkg.TinyYolo(num_classes=2).to(dev).train()
for m in model.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.weight.data.fill_(3.0)  # train-mode BN normalises per step, so gamma alone sets the firing rate
if args.sync_bn:
    sharding.convert_sync_batchnorm(model)
opt = model.configure_optimizers()
params = [p for p in model.parameters() if p.requires_grad]
T, B, (H, W) = args.T, args.batch, args.hw
gen = torch.Generator(device=dev).manual_seed(100 + rank)
X = (torch.rand(T, B, 2, H, W, device=dev, generator=gen) < 0.07).float()
labels = torch.full((B, 10, 5), -1.0, device=dev)
ctr = torch.rand(B, 6, 2, device=dev, generator=gen) * 0.6 + 0.2
wh = torch.rand(B, 6, 2, device=dev, generator=gen) * 0.25 + 0.05
labels[:, :6, 0] = torch.randint(0, 2, (B, 6), device=dev, generator=gen).float()
labels[:, :6, 1:3], labels[:, :6, 3:5] = ctr - wh / 2, ctr + wh / 2
times = []
for it in range(args.steps + 1):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.profile_last and it == args.steps:
        torch.cuda.cudart().cudaProfilerStart()
    e0.record()
    opt.zero_grad(set_to_none=True)
    loss = model._loss(model(X), labels)
    loss.backward()
    sharding.allreduce_flat_gradients(params)
    opt.step()
    e1.record()
    torch.cuda.synchronize()
    if args.profile_last and it == args.steps:
        torch.cuda.cudart().cudaProfilerStop()
    if it > 0:
        times.append(e0.elapsed_time(e1))
ms = sum(times) / len(times)
if world > 1:
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
gn = float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in params if p.grad is not None)))
if rank == 0:
    print(json.dumps({"what": "TinyYolo BPTT training step (detection loss on synthetic labels)", "n_gpus": world, "sync_bn": bool(args.sync_bn), "batch_per_gpu": B, "T": T,
                      "hw": [H, W], "ms_per_step": ms, "frames_per_s": world * T * B / (ms * 1e-3),
                      "loss": float(loss), "grad_norm": gn,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30}))
if world > 1:
    dist.destroy_process_group()
