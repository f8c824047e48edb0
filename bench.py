"""Headline benchmark: timestep-frames/s of the SODa spiking forward hot path on B200.

  python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, sm_100a)
  python bench.py --impl reference --gpus N --steps K ...   # the reference algorithm on host cores

Workload = BASELINE.json configs[1]: TinyYolo (YOLOv8-style SODa detector) forward, GEN1 304x240,
2 polarities, batch 5 per GPU, T=32, 16 ms step, synthetic events (10 k events / step / sample,
~7 % occupancy), synthetic calibrated weights.  One "step" = event binning of the raw Event2D
records + the layer-major forward over all T frames -> last-step (cls, bbox).  Samples are
independent, so N GPUs = N ranks each with its own batch of 5 (weak scaling, no data-path collective).

value : frames/s with the raw event records already resident in HBM (CUDA-graph replay).
e2e   : same, through the public API from pinned HOST buffers: H2D of the records, binning,
        forward, D2H of (cls, bbox) inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_STEPS, BATCH, H, W = 32, 5, 240, 304
STEP_US, EVENTS_PER_STEP = 16000, 10000
GFLOP_PER_FRAME = 7.6156  # dense 2*MAC of the 48 convs, TinyYolo GEN1
METRIC = "timestep-frames/s, SODa GEN1 304x240 T=32 fwd"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# --------------------------------------------------------------------------------------------------
# synthetic inputs (seed = 1000*config + 10*rank + sample)
# --------------------------------------------------------------------------------------------------
def make_event_samples(rank: int, batch: int, T: int):
    import numpy as np

    out = []
    for b in range(batch):
        rng = np.random.default_rng(1000 * 2 + 10 * rank + b)
        n = EVENTS_PER_STEP * T
        t = np.sort(rng.integers(0, T * STEP_US, size=n, dtype=np.int64)).astype(np.uint32)
        x = rng.integers(0, W, size=n, dtype=np.uint32)
        y = rng.integers(0, H, size=n, dtype=np.uint32)
        p = rng.integers(0, 2, size=n, dtype=np.uint32)
        out.append(np.stack([t, x | (y << 14) | (p << 28)], axis=1).astype("<u4"))
    return out


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                r = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                    "-i", str(self.index)], capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.rows.append([c.strip() for c in r.stdout.strip().split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in self.rows:
            try:
                sm.append(float(row[0]))
                mx = max(mx, float(row[1]))
            except ValueError:
                continue
            for n, v in zip(names, row[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port) on the host cores
# --------------------------------------------------------------------------------------------------
def cpu_reference_frames_per_s(steps: int, warmup: int, T: int = 2, batch: int = 1):
    """Times the CPU restatement of the reference path (numpy binning + time-major torch-CPU net,
    all host threads) on a bounded sample of the workload: same net, same resolution, B x T smaller."""
    import numpy as np
    import torch

    from oracle import binning as ob
    from oracle import net_restated as onr
    import snn_for_object_detection_b200 as pkg

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = pkg.TinyYolo(num_classes=2).eval()
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.fill_(5.0)
    samples = make_event_samples(0, batch, T)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        frames = ob.bin_batch(samples, [0] * batch, STEP_US, T, H, W, clip_x=True)
        onr.soda_forward(model, torch.from_numpy(frames))
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    per = sum(times) / len(times)
    return (T * batch) / per, per, cores, f"TinyYolo GEN1 304x240, B={batch}, T={T} per step (of B={BATCH}, T={T_STEPS})"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    v, per, cores, sample = cpu_reference_frames_per_s(args.steps, max(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "TinyYolo SODa fwd, GEN1 304x240x2, B=5, T=32 (configs[1])", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    import snn_for_object_detection_b200 as pkg
    from snn_for_object_detection_b200.utils.calibrate import calibrate

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    T, B = args.T, args.batch
    torch.manual_seed(0)
    model = pkg.TinyYolo(num_classes=2).to(dev).eval()
    samples = make_event_samples(rank, B, T)
    ev_host, offs_host = pkg.events.pack_samples(samples, pinned=True)
    t0_host = torch.zeros(B, dtype=torch.int64).pin_memory()
    offs_host = offs_host.pin_memory()
    ev_dev = torch.empty_like(ev_host, device=dev)
    offs_dev = offs_host.to(dev)
    t0_dev = t0_host.to(dev)
    ev_dev.copy_(ev_host)
    frames = torch.empty((T, B, 2, H, W), device=dev)

    def step_device():
        pkg.ops.bin_events(ev_dev, offs_dev, t0_dev, STEP_US, T, H, W, clip_x=True, out=frames)
        anchors, cls, bbox = model(frames)
        return cls, bbox

    with torch.no_grad():
        step_device()  # packs weights, caches anchors
        calibrate(model, frames, gain=5.0)
        cls, bbox = step_device()
        torch.cuda.synchronize()
        rate = None

        # ---- CUDA graph of one step (bin + forward) ------------------------------------------
        use_graph = not args.no_graph
        graph = None
        if use_graph:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    step_device()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            n0 = pkg._lib.launch_count()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                g_cls, g_bbox = step_device()
            launches_per_step = pkg._lib.launch_count() - n0
            run_step = graph.replay
        else:
            n0 = pkg._lib.launch_count()
            g_cls, g_bbox = step_device()
            launches_per_step = pkg._lib.launch_count() - n0
            run_step = step_device

        cls_host = torch.empty(g_cls.shape, dtype=g_cls.dtype).pin_memory()
        bbox_host = torch.empty(g_bbox.shape, dtype=g_bbox.dtype).pin_memory()

        def step_e2e():
            ev_dev.copy_(ev_host, non_blocking=True)
            offs_dev.copy_(offs_host, non_blocking=True)
            t0_dev.copy_(t0_host, non_blocking=True)
            run_step()
            cls_host.copy_(g_cls, non_blocking=True)
            bbox_host.copy_(g_bbox, non_blocking=True)

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        def timed(fn, steps, warmup):
            for _ in range(warmup):
                fn()
            barrier()
            sampler = ClockSampler(local)
            sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            barrier()
            clocks = sampler.stop()
            ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms, clocks

        ms_dev, clocks = timed(run_step, args.steps, args.warmup)
        ms_e2e, _ = timed(step_e2e, args.steps, args.warmup)

        # ---- per-kernel-family device time of one eager step (CUDA events on the launch stream) --
        # The GPU is parked on a spin kernel while the host enqueues the whole step, so that each
        # (start, end) event pair brackets device time only, not host launch latency.
        prof = {}
        step_device()
        torch.cuda.synchronize()
        torch.cuda._sleep(int(1.2e9))
        pkg.ops.PROFILE = prof
        step_device()
        torch.cuda.synchronize()
        pkg.ops.PROFILE = None
        tags = {k: prof.pop(k + "_tags", []) for k in ("conv2d", "conv_lif")}
        meta = {k: prof.pop(k + "_meta", []) for k in ("conv2d", "conv_lif")}
        conv_launches = []  # (family, meta, ms) of every convolution launch of the step
        for k in ("conv2d", "conv_lif"):
            for m, (a, b) in zip(meta[k], prof.get(k, [])):
                conv_launches.append((k, m, a.elapsed_time(b)))
        fam = {k: sum(a.elapsed_time(b) for a, b in v) for k, v in prof.items()}
        if args.per_layer:
            for k in ("conv2d", "conv_lif"):
                for tag, (a, b) in zip(tags[k], prof.get(k, [])):
                    sys.stderr.write(f"{k:8s} {a.elapsed_time(b):8.3f} ms  {tag}\n")
        n_conv = len(prof.get("conv2d", [])) + len(prof.get("conv_lif", []))
        rate = float((g_cls.abs() > 0).float().mean())
        if not rate > 0.5:
            raise SystemExit(f"bench.py: the network is silent (class outputs {rate:.3f} non-zero) - calibration failed")

    pk, pk_src = peaks()
    frames_per_step = T * B
    ms_per_step = ms_dev / args.steps
    value = world * frames_per_step / (ms_per_step * 1e-3)
    e2e_value = world * frames_per_step / (ms_e2e / args.steps * 1e-3)
    conv_ms = fam.get("conv2d", 0.0) + fam.get("conv_lif", 0.0)

    def flops(m):
        pad = m["k"] // 2
        ho = (m["H"] + 2 * pad - m["k"]) // m["stride"] + 1
        wo = (m["W"] + 2 * pad - m["k"]) // m["stride"] + 1
        return 2.0 * m["N"] * ho * wo * m["Cout"] * m["Cin"] * m["k"] * m["k"]

    # Dominant kernel of the step = conv_tc3_kernel<128, false> (tensor-core conv, Cout > 64, plain epilogue):
    # ALGORITHMIC flops of its launches (dense 2*MAC as the reference executes them) / their summed
    # device time, both from this run's CUDA events.
    dom = [(m, ms) for fam_, m, ms in conv_launches if fam_ == "conv2d" and m["kind"] in (1, 2) and m["Cout"] > 64]
    dom_flops, dom_ms = sum(flops(m) for m, _ in dom), sum(ms for _, ms in dom)
    achieved = dom_flops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else None
    peak = pk["bf16_tflops_sustained"]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r1d_traffic.json")
    traffic_launch = None
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic, traffic_launch = tj.get("bytes"), tj.get("launch")
    all_flops = sum(flops(m) for _, m, _ in conv_launches)
    roofline = {
        "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
        "frac": (achieved / peak) if achieved else None,
        "traffic": traffic, "traffic_launch": traffic_launch,
        "kernel": "soda::conv_tc3_kernel<128, false, *> (tcgen05/TMEM/TMA implicit GEMM; tf32 main pass + K-stacked bf16 "
                  "correction pass = 2 tensor passes per product for fp32-level accuracy)",
        "peak_source": f"{pk_src} dense bf16 sustained (kernel timed inside a long step); the 2-pass fp32-accurate "
                       "scheme itself tops out near peak/4 (tf32 rate = bf16/2, two passes)",
        "launches": len(dom), "kernel_ms_per_step": dom_ms, "share_of_step": dom_ms / ms_per_step if dom else None,
        "all_conv": {"launches": n_conv, "ms_per_step": conv_ms, "gflop_per_step": all_flops / 1e9,
                     "tflops": all_flops / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else None},
        "note": "achieved = dense 2*MAC flops of this kernel's launches in one step / their summed CUDA-event time "
                "(eager pass of the same step, GPU parked while the host enqueues); traffic = dram bytes of one "
                "ncu --set full capture of the listed launch shape (profiles/)",
    }
    h2d = ev_host.numel() * ev_host.element_size() + offs_host.numel() * 8 + t0_host.numel() * 8
    d2h = cls_host.numel() * 4 + bbox_host.numel() * 4

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, per, cores, sample = cpu_reference_frames_per_s(2, 1)
            cpu = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample}
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": f"TinyYolo SODa fwd, {'1Mpx 1280x720' if args.mpx else 'GEN1 304x240'}x2, B={B}/GPU, T={T}, "
                            f"16 ms step ({'configs[2] shape' if args.mpx else 'configs[1]'}); "
                            "step = event binning + layer-major forward",
                "events_per_step_per_sample": EVENTS_PER_STEP, "weights": "synthetic, calibrated (gain 5)",
                "l2": "per-step working set (>2 GB of fp32 activations) exceeds the 126 MB L2; no explicit flush",
                "cuda_graph": bool(graph is not None),
            },
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches_per_step * args.steps),
            "clocks": clocks, "roofline": roofline, "kernel_ms_per_step": fam,
            "cls_nonzero_frac": rate,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--T", type=int, default=T_STEPS)
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--mpx", action="store_true",
                    help="1Mpx 1280x720 frames, 150 k events / step (BASELINE configs[2] shape; use --batch 1 per GPU)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--per-layer", action="store_true", help="print per-conv device times to stderr")
    args = ap.parse_args()
    if args.mpx:
        global H, W, EVENTS_PER_STEP, GFLOP_PER_FRAME, METRIC
        H, W, EVENTS_PER_STEP, GFLOP_PER_FRAME = 720, 1280, 150000, 95.974
        METRIC = "timestep-frames/s, SODa 1Mpx 1280x720 T=32 fwd"
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
