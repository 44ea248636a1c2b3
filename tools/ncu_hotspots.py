"""Developer helper: aggregates the SASS page of an ncu report (ncu -i X.ncu-rep --page source --csv) by opcode:
executed warp instructions and warp-stall samples. usage: python tools/ncu_hotspots.py report.ncu-rep [kernel-substring]"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ci = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed")}
stall_cols = [(n, i) for i, n in enumerate(hdr) if n.startswith("stall_") and "Not Issued" not in n]
by_op = defaultdict(lambda: [0, 0])
stalls = defaultdict(int)
total_i = total_s = 0
lines = []
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    src = r[ci["Source"]].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    n, s = int(r[ci["Instructions Executed"]] or 0), int(r[ci["# Samples"]] or 0)
    by_op[op.split(".")[0]][0] += n
    by_op[op.split(".")[0]][1] += s
    total_i += n
    total_s += s
    for name, i in stall_cols:
        stalls[name] += int(r[i] or 0)
    lines.append((s, n, src))
print(f"total warp instructions {total_i}, samples {total_s}")
print("by opcode (executed, samples):")
for op, (n, s) in sorted(by_op.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"  {op:12s} {n:10d} {100.0 * n / max(total_i, 1):5.1f}%   samples {s:6d} {100.0 * s / max(total_s, 1):5.1f}%")
print("stall reasons:", {k: v for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]})
print("hottest instructions:")
for s, n, src in sorted(lines, reverse=True)[:25]:
    print(f"  {s:6d} {n:9d}  {src}")
