"""Opcode mix of an ncu report (needs --import-source on): warp-instructions executed and stall samples per SASS opcode.
  python scripts/ncu_opmix.py gpurun_out/x.ncu-rep [top]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
ex, smp = collections.Counter(), collections.Counter()
for r in rows[2:]:
    s = r[idx["Source"]].strip()
    s = re.sub(r"^@!?U?P\d+\s+", "", s)
    op = s.split()[0] if s else "?"
    op = op.split(".")[0] + ("." + op.split(".")[1] if "." in op and op.split(".")[0] in ("LDG", "STS", "LDS", "LDL", "STL", "SYNCS", "MOV", "F2FP") else "")
    ex[op] += int(r[idx["Instructions Executed"]] or 0)
    smp[op] += int(r[idx["# Samples"]] or 0)
te, ts = sum(ex.values()), sum(smp.values())
print(f"{te} warp-instructions, {ts} samples")
for op, n in ex.most_common(top):
    print(f"{op:16s} {n:12d} {100 * n / te:5.1f}%   samples {smp[op]:6d} {100 * smp[op] / ts:5.1f}%")
