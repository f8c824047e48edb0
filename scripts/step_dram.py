"""Whole-step DRAM traffic from an ncu launch list with gpu__time_duration.sum + dram__bytes_{read,write}.sum:
one step = the launches between the last two bin_bitmap_kernel launches.
  python scripts/step_dram.py gpurun_out/r2_config2_dram.csv"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
ki, ni, vi, ui, idi = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
launch = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    d = launch.setdefault(r[idi], {"name": r[ki]})
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    if "time" in r[ni]:
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)
    else:
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    d[r[ni]] = v
L = list(launch.values())
marks = [i for i, d in enumerate(L) if "bin_bitmap_kernel" in d["name"]]
step = L[marks[-2]:marks[-1]] if len(marks) > 1 else L[marks[-1]:]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for d in step:
    n = re.sub(r"\(.*", "", d["name"]).replace("void ", "")
    a = agg[n]
    a[0] += 1
    a[1] += d.get("gpu__time_duration.sum", 0)
    a[2] += d.get("dram__bytes_read.sum", 0)
    a[3] += d.get("dram__bytes_write.sum", 0)
tt = sum(a[1] for a in agg.values()); rd = sum(a[2] for a in agg.values()); wr = sum(a[3] for a in agg.values())
print(f"one step: {len(step)} launches, {tt / 1e3:.3f} ms summed under ncu, DRAM read {rd / 1e9:.3f} GB + write {wr / 1e9:.3f} GB "
      f"= {(rd + wr) / 1e9:.3f} GB")
print("| launches | us | read MB | write MB | GB/s | kernel |\n|---|---|---|---|---|---|")
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {a[0]} | {a[1]:.1f} | {a[2] / 1e6:.1f} | {a[3] / 1e6:.1f} | {(a[2] + a[3]) / a[1] / 1e3 if a[1] else 0:.0f} | `{n[:80]}` |")
