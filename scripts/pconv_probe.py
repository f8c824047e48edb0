"""Times single packed-activation convolutions (soda_pconv_fwd / soda_pconv_lif_fwd) of the TinyYolo GEN1 forward
through the C ABI: CUDA events, rotating input buffers so every launch misses L2.

  python scripts/pconv_probe.py [spec ...]     spec = N,Cin,H,W,Cout,k,s,in,out[,T[,res]]   (in/out: pair|half|f32)
Kernel experiments go through the environment (SODA_TC4_BN / _MT / _MIN_A / _DBG), one process per setting.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from snn_for_object_detection_b200 import ops

DEFAULT = [
    "160,64,120,152,64,1,1,half,pair",
    "160,64,120,152,32,1,1,pair,pair",
    "160,128,120,152,64,1,1,pair,pair",
    "160,32,120,152,32,3,1,pair,pair",
    "160,32,120,152,32,3,1,pair,pair,32,res",
    "160,64,120,152,128,3,2,pair,f32",
    "160,64,120,152,128,3,2,pair,half,32",
    "160,64,60,76,64,3,1,pair,f32",
    "160,64,60,76,64,3,1,pair,pair,32,res",
    "160,320,60,76,128,1,1,pair,pair",
    "160,128,60,76,256,3,2,pair,f32",
    "160,128,30,38,128,3,1,pair,f32",
    "160,768,30,38,256,1,1,pair,pair",
    "160,256,30,38,256,3,2,pair,f32",
    "160,128,15,19,128,3,1,pair,f32",
]
F_ = {"f32": ops.FMT_F32, "pair": ops.FMT_PAIR, "half": ops.FMT_HALF}


def run(spec, iters=10):
    parts = spec.split(",")
    N, Cin, H, W, Cout, k, s = map(int, parts[:7])
    in_fmt, out_fmt = parts[7], parts[8]
    T = int(parts[9]) if len(parts) > 9 else 0
    res = len(parts) > 10 and parts[10] == "res"
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    eb = 2 if in_fmt == "half" else 4
    nbuf = min(6, max(2, int(400e6 // (N * Cin * H * W * eb)) + 1))
    xs = []
    for _ in range(nbuf):
        x = torch.randn(N, Cin, H, W, device=dev, generator=g)
        if in_fmt == "half":
            x = (x > 1.0).float()
        xs.append(ops.act_convert(x.contiguous(memory_format=torch.channels_last), F_[in_fmt]))
    w = torch.randn(Cout, Cin, k, k, device=dev, generator=g) * (3.0 / (Cin * k * k)) ** 0.5
    pc = ops.pack_pconv(w, s, None, None, F_[in_fmt])
    Ho, Wo = ops.conv_out_hw(H, W, k, s)
    out = ops.empty_act(N, Cout, Ho, Wo, dev, F_[out_fmt])
    r = None
    if res:
        r = ops.act_convert(torch.randn(N, Cout, Ho, Wo, device=dev, generator=g).contiguous(memory_format=torch.channels_last),
                            ops.FMT_PAIR)

    def call(i):
        if T:
            ops.pconv_lif(xs[i % nbuf], pc, T, ops.LIF_DEFAULT, residual=r, out=out)
        else:
            ops.pconv(xs[i % nbuf], pc, residual=r, out=out)

    for i in range(3):
        call(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        call(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flops = 2.0 * N * Ho * Wo * Cout * Cin * k * k
    ob = 2 if out_fmt == "half" else 4
    byts = float(eb) * N * Cin * H * W + float(ob) * N * Cout * Ho * Wo * (2 if res else 1)
    print(f"{spec:44s} {ms:8.4f} ms  {flops / ms / 1e9:8.1f} TFLOP/s  {byts / ms / 1e6:8.1f} GB/s", flush=True)


if __name__ == "__main__":
    tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("SODA_"))
    print(f"# {tag or 'defaults'}")
    for spec in (sys.argv[1:] or DEFAULT):
        run(spec)
