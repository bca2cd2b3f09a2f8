#!/usr/bin/env python
"""Per-source-line warp-level vs thread-level instruction counts of an .ncu-rep (source page): where a kernel
runs with partly empty warps. Usage: python tools/ncu_divergence.py file.ncu-rep [min_share_pct]"""
import csv, subprocess, sys
from collections import defaultdict
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
cur, hdr = None, None
agg = defaultdict(lambda: [0, 0, 0, ""])
for row in csv.reader(out.splitlines()):
    if not row:
        continue
    if row[0] == "File Path":
        cur = row[1].split("/")[-1]; hdr = None; continue
    if row[0] == "Line No":
        hdr = row; continue
    if hdr is None or not row[0].isdigit():
        continue
    d = dict(zip(hdr, row))
    try:
        a = agg[(cur, int(row[0]))]
        a[0] += int(d.get("Instructions Executed", "0") or 0)
        a[1] += int(d.get("Thread Instructions Executed", "0") or 0)
        a[2] += int(d.get("Divergent Branches", "0") or 0)
        a[3] = row[1][:90]
    except ValueError:
        pass
tot = sum(a[0] for a in agg.values()) or 1
tth = sum(a[1] for a in agg.values())
print(f"total warp-inst {tot}  thread-inst {tth}  avg active threads {tth / tot:.2f}")
for (f, ln), a in sorted(agg.items()):
    if 100.0 * a[0] / tot >= minpct:
        print(f"{f:24s}{ln:5d} inst {100.0 * a[0] / tot:5.1f}%  threads/inst {a[1] / max(a[0], 1):5.1f}  div.br {a[2]:9d} | {a[3]}")
