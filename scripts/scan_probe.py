"""Times soda_lif_fmt_fwd (LIF scan + residual + store in any activation format) on the TinyYolo GEN1 shapes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from snn_for_object_detection_b200 import ops

F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF, "none": None}
DEFAULT = ["32,5,128,60,76,none,half", "32,5,64,60,76,pair,pair", "32,5,256,30,38,none,half", "32,5,128,30,38,pair,pair",
           "32,5,128,15,19,pair,pair", "32,5,64,60,76,f32,f32"]


def run(spec, iters=10):
    T, B, C, H, W = map(int, spec.split(",")[:5])
    res_fmt, out_fmt = spec.split(",")[5:7]
    dev = "cuda"
    xs = [torch.randn(T * B, C, H, W, device=dev).contiguous(memory_format=torch.channels_last) * 3 for _ in range(3)]
    res = None
    if F_[res_fmt] is not None:
        res = ops.act_convert(torch.randn(T * B, C, H, W, device=dev).contiguous(memory_format=torch.channels_last), F_[res_fmt])
    out = ops.empty_act(T * B, C, H, W, dev, F_[out_fmt])
    for i in range(3):
        ops.lif_scan_fmt(xs[i % 3], T, residual=res, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        ops.lif_scan_fmt(xs[i % 3], T, residual=res, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    n = T * B * C * H * W
    byts = n * (4 + (0 if res is None else (2 if res_fmt == "half" else 4)) + (2 if out_fmt == "half" else 4))
    print(f"{spec:34s} {ms:8.4f} ms  {byts / ms / 1e6:8.1f} GB/s  ({n / 1e6:.1f} M neuron-steps)", flush=True)


if __name__ == "__main__":
    for spec in (sys.argv[1:] or DEFAULT):
        run(spec)
