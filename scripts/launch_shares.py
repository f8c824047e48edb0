"""Per-kernel shares of ONE forward step from an ncu launch list (`--metrics gpu__time_duration.sum --csv`).
A step = the launches between the last two bin_bitmap_kernel launches (one eager forward step of bench.py).
  python scripts/launch_shares.py gpurun_out/launches.csv"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
data = [r for r in rows[hi + 1:] if len(r) > vi]
marks = [i for i, r in enumerate(data) if "bin_bitmap_kernel" in r[ki]]
# last COMPLETE step: between the last two binning launches (the capture may end mid-step)
step = data[marks[-2]:marks[-1]] if len(marks) > 1 else data[marks[-1]:]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in step:
    n = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v for _, v in agg.values())
print(f"one step: {len(step)} launches, {tot / 1e3:.3f} ms summed (cold-cache, serialised under ncu)")
print("| share | launches | us | kernel |\n|---|---|---|---|")
for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {100 * v / tot:5.1f} % | {c} | {v:.1f} | `{n[:100]}` |")
