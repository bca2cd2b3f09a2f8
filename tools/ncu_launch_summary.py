#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`), sorted by total time.
Usage: python tools/ncu_launch_summary.py launches.csv > launches.summary.txt"""
import csv, sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith("==")]
h = rows[0]
kn, mn, mu, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value")
tot = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
        continue
    t = float(r[mv].replace(",", ""))
    t *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[mu], 1e-3)
    e = tot[r[kn]]
    e[0] += 1
    e[1] += t
for name, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{c:4d} launches {t:10.1f} us total {t / c:10.1f} us/launch  {name[:100]}")
