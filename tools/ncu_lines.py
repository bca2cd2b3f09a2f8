#!/usr/bin/env python
"""Per-source-line aggregation of an .ncu-rep source page: instructions executed, stall samples, shared wavefronts.
Usage: python tools/ncu_lines.py file.ncu-rep [min_share_pct]"""
import csv, subprocess, sys
from collections import defaultdict
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
cur, hdr = None, None
agg = defaultdict(lambda: [0, 0, 0, ""])
for row in csv.reader(out.splitlines()):
    if not row:
        continue
    if row[0] == "File Path":
        cur = row[1].split("/")[-1]; hdr = None; continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Line No":
        hdr = row; continue
    if hdr is None or not row[0].isdigit():
        continue
    d = dict(zip(hdr, row))
    try:
        key = (cur, int(row[0]))
        a = agg[key]
        a[0] += int(d.get("Instructions Executed", "0") or 0)
        a[1] += int(d.get("# Samples", "0") or 0)
        a[2] += int(d.get("L1 Wavefronts Shared", "0") or 0)
        a[3] = row[1][:110]
    except ValueError:
        pass
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
tw = sum(a[2] for a in agg.values()) or 1
print(f"total inst {ti}  samples {ts}  shared wavefronts {tw}")
for key in sorted(agg):
    a = agg[key]
    if 100.0 * a[0] / ti >= minpct or 100.0 * a[1] / ts >= minpct or 100.0 * a[2] / tw >= minpct:
        print(f"{key[0][:22]:22s} {key[1]:4d} inst {100.0*a[0]/ti:5.1f}%  stall {100.0*a[1]/ts:5.1f}%  smem {100.0*a[2]/tw:5.1f}%  | {a[3]}")
