"""Times single convolution shapes of the TinyYolo GEN1 forward through the C ABI (CUDA events,
rotating input buffers so every launch misses L2) and checks each against an fp64 reference.

  python scripts/conv_probe.py                 # the default shape list
  python scripts/conv_probe.py 160,32,120,152,32,3,1[,exact] ...
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

import snn_for_object_detection_b200 as pkg
from snn_for_object_detection_b200 import _lib, ops

DEFAULT = [
    "160,32,120,152,32,3,1",
    "160,64,60,76,64,3,1",
    "160,128,30,38,128,3,1",
    "160,64,120,152,128,3,2",
    "160,128,60,76,256,3,2",
    "160,128,120,152,64,1,1",
    "160,64,120,152,64,1,1,exact",
    "160,320,60,76,128,1,1",
    "160,768,30,38,256,1,1",
]


def run(spec, iters=10, check=True):
    parts = spec.split(",")
    N, Cin, H, W, Cout, k, s = map(int, parts[:7])
    exact = len(parts) > 7 and parts[7] == "exact"
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    nbuf = max(2, int(400e6 // (N * Cin * H * W * 4)) + 1)
    nbuf = min(nbuf, 6)
    xs = []
    for _ in range(nbuf):
        x = torch.randn(N, Cin, H, W, device=dev, generator=g)
        if exact:
            x = (x > 1.0).float()
        x = x.contiguous(memory_format=torch.channels_last)
        if exact:
            ops.mark_exact(x)
        xs.append(x)
    w = torch.randn(Cout, Cin, k, k, device=dev, generator=g) / (Cin * k * k) ** 0.5
    kind = ops.conv_kind_for(xs[0], Cin, k, s)
    pc = ops.pack_conv(w, s, None, None, kind)
    Ho, Wo = ops.conv_out_hw(H, W, k, s)
    out = ops.empty_pixel_major(N, Cout, Ho, Wo, dev)
    for i in range(3):
        ops.conv2d(xs[i % nbuf], pc, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        ops.conv2d(xs[i % nbuf], pc, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flops = 2.0 * N * Ho * Wo * Cout * Cin * k * k
    byts = 4.0 * (N * Cin * H * W + N * Cout * Ho * Wo)
    err = float("nan")
    if check:
        nchk = min(N, 2)
        ops.conv2d(xs[0], pc, out=out)
        ref = F.conv2d(xs[0][:nchk].double(), w.double(), padding=k // 2, stride=s)
        err = float((out[:nchk].double() - ref).abs().max() / ref.abs().max())
    print(f"{spec:38s} kind{kind} {ms:8.4f} ms  {flops / ms / 1e9:8.1f} TFLOP/s  {byts / ms / 1e6:8.1f} GB/s  relerr {err:.2e}",
          flush=True)


def run_lif(spec, T=8, iters=5):
    """Fused conv+LIF vs conv2d followed by the stand-alone scan (same kernels' arithmetic)."""
    parts = spec.split(",")
    N, Cin, H, W, Cout, k, s = map(int, parts[:7])
    res = len(parts) > 7 and parts[7] == "res"
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(2)
    x = (torch.randn(N, Cin, H, W, device=dev, generator=g) * 0.8).contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, k, k, device=dev, generator=g) / (Cin * k * k) ** 0.5
    shift = torch.randn(Cout, device=dev, generator=g) * 0.1
    pc = ops.pack_conv(w, s, None, shift, _lib.SODA_CONV_TC)
    B = N // T
    y0 = ops.conv2d(x, pc)
    z0, v0, i0, _, _, _ = ops.lif_scan(y0, T)
    if res:
        z0 = z0 + x
    z1, v1, i1 = ops.conv_lif(x, pc, T, ops.LIF_DEFAULT, residual=x if res else None)
    torch.cuda.synchronize()
    mism = float((z0 != z1).float().mean())
    dv = float((v0 - v1).abs().max())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.conv_lif(x, pc, T, ops.LIF_DEFAULT, residual=x if res else None)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"LIF {spec:34s} T={T} B={B} {ms:8.4f} ms  rate {float(z1.mean() if not res else (z1 - x).mean()):.3f} "
          f"mismatch {mism:.2e} dv {dv:.2e}", flush=True)


def run_direct(spec, iters=5):
    """Event-frame layer: direct Conv->BN->LIF kernel vs conv2d(fp32) + lif_scan.  spec = T,B,H,W,Cout,k,s"""
    T, B, H, W, Cout, k, s = map(int, spec.split(","))
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(3)
    x = (torch.rand(T * B, 2, H, W, device=dev, generator=g) < 0.07).float()
    w = torch.randn(Cout, 2, k, k, device=dev, generator=g)
    pc = ops.pack_conv(w, s, None, torch.zeros(Cout, device=dev), _lib.SODA_CONV_FP32)
    y0 = ops.conv2d(x, pc)
    z0, v0, i0, _, _, _ = ops.lif_scan(y0, T)
    z1, v1, i1 = ops.conv_lif(x, pc, T, ops.LIF_DEFAULT)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.conv_lif(x, pc, T, ops.LIF_DEFAULT)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"DIRECT {spec:30s} {ms:8.4f} ms  {z1.numel() * 4 / ms / 1e6:8.1f} GB/s out  rate {float(z1.mean()):.3f} "
          f"equal {bool(torch.equal(z0, z1) and torch.equal(v0, v1))}", flush=True)


if __name__ == "__main__":
    specs = [a for a in sys.argv[1:] if not a.startswith("--")] or DEFAULT
    iters = 10
    for a in sys.argv[1:]:
        if a.startswith("--iters="):
            iters = int(a.split("=")[1])
    with torch.no_grad():
        for sp in specs:
            if sp.startswith("lif:"):
                run_lif(sp[4:])
            elif sp.startswith("direct:"):
                run_direct(sp[7:])
            else:
                run(sp, iters=iters, check="--nocheck" not in sys.argv)

