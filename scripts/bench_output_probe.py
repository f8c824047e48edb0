"""Is the bench model's output alive?  Prints statistics of the class / box outputs of the bench.py configuration
(eager and from a CUDA graph) and the firing rate of every LIF layer.  python scripts/bench_output_probe.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import bench
import snn_for_object_detection_b200 as pkg
from snn_for_object_detection_b200.utils.calibrate import calibrate

dev = torch.device("cuda", 0)
T, B, H, W = 32, 5, bench.H, bench.W
torch.manual_seed(0)
model = pkg.TinyYolo(num_classes=2).to(dev).eval()
samples = bench.make_event_samples(0, B, T)
ev, offs = pkg.events.pack_samples(samples, pinned=False)
ev, offs = ev.to(dev), offs.to(dev)
t0 = torch.zeros(B, dtype=torch.int64, device=dev)
frames = torch.empty((T, B, 2, H, W), device=dev)


def step():
    pkg.ops.bin_events(ev, offs, t0, bench.STEP_US, T, H, W, clip_x=True, out=frames)
    return model(frames)


def stats(name, t):
    t = t.float()
    print(f"{name:10s} shape {tuple(t.shape)} nonzero {float((t != 0).float().mean()):.4f} mean {float(t.mean()):+.4e} "
          f"absmax {float(t.abs().max()):.4e} std {float(t.std()):.4e}")


with torch.no_grad():
    step()
    calibrate(model, frames, gain=5.0)
    print("input occupancy", float((frames != 0).float().mean()))
    anchors, cls, bbox = step()
    stats("cls eager", cls), stats("bbox eager", bbox)
    c0, b0 = cls.clone(), bbox.clone()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        _, gc, gb = step()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    stats("cls graph", gc), stats("bbox graph", gb)
    print("graph == eager:", bool(torch.equal(gc, c0)), bool(torch.equal(gb, b0)))
