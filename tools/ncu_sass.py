#!/usr/bin/env python
"""SASS-level rows of an .ncu-rep source page filtered by opcode regex: executed count, shared wavefronts, samples.
Usage: python tools/ncu_sass.py file.ncu-rep 'LDS|STS' [B]"""
import csv, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
pat = re.compile(sys.argv[2])
B = float(sys.argv[3]) if len(sys.argv) > 3 else 65536.0
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ix = {n: h.index(n) for n in ("Source", "Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "# Samples")}
tot = [0, 0, 0]
for n, r in enumerate(rows[hi + 1:]):
    if len(r) < len(h) or not pat.search(r[ix["Source"]]):
        continue
    ex, wf, wi, sm = (int(r[ix[k]] or 0) for k in ("Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "# Samples"))
    tot[0] += ex; tot[1] += wf; tot[2] += sm
    print(f"{n:5d} {r[ix['Source']].strip()[:60]:60s} exec/f {ex/B:6.2f} wf/f {wf/B:7.2f} ideal/f {wi/B:7.2f} wf/exec {wf/max(ex,1):5.2f} samples {sm}")
print(f"TOTAL exec/f {tot[0]/B:.1f} wf/f {tot[1]/B:.1f} samples {tot[2]}")
