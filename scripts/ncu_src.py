"""Summarise an ncu report: key raw metrics + the hottest SASS lines with their stall reasons.
  python scripts/ncu_src.py gpurun_out/prof.ncu-rep [min_samples]"""
import csv, io, re, subprocess, sys

rep = sys.argv[1]
mins = int(sys.argv[2]) if len(sys.argv) > 2 else 150
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct",
        "gpu__dram_throughput.avg.pct", "lts__throughput.avg.pct", "l1tex__throughput.avg.pct", "sm__cycles_elapsed.max",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.sum",
        "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__issue_active.avg.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed_pipe_xu", "lts__t_bytes.sum "]
for k in range(2, len(rows)):
    for a, b, c in zip(rows[0], rows[1], rows[k]):
        if any(a.startswith(x.strip()) for x in keys) and "Triage" not in a:
            print(f"{a:75s} {b:12s} {c}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
tot = sum(int(r[idx["# Samples"]] or 0) for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot)
for r in data:
    s = int(r[idx["# Samples"]] or 0)
    srcl = r[idx["Source"]]
    if s >= mins or re.search(r"UTMALDG|UTCHMMA|UTCBAR|LDTM|STTM|EXIT", srcl):
        st = sorted(((h, int(r[idx[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[:2]
        print(f"{r[idx['Address']][-5:]} {s:6d} {100 * s / tot:5.1f}% ex={r[idx['Instructions Executed']]:>9s} {srcl[:80]:80s} {st}")
