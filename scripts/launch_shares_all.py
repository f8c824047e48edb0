"""Per-kernel shares over a WHOLE ncu launch list (`--metrics gpu__time_duration.sum --csv`), for captures already cut to
one step with cudaProfilerStart/Stop (`ncu --profile-from-start off ... train_step_probe.py --profile-last`).
  python scripts/launch_shares_all.py gpurun_out/train_launches.csv [top]"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
n_launch = 0
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    n = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    n = re.sub(r"<.*", "<…>", n) if not n.startswith("soda::") else n
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    agg[n][0] += 1
    agg[n][1] += v
    n_launch += 1
tot = sum(v for _, v in agg.values())
ours = sum(v for k, (_, v) in agg.items() if k.startswith("soda::"))
print(f"{n_launch} launches, {tot / 1e3:.3f} ms summed (cold-cache, serialised under ncu); soda:: kernels {100 * ours / tot:.1f} % of the time")
print("| share | launches | us | kernel |\n|---|---|---|---|")
for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"| {100 * v / tot:5.1f} % | {c} | {v:.1f} | `{n[:110]}` |")
