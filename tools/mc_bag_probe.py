#!/usr/bin/env python
"""Diagnostic: one Monte Carlo replay of the surrogate bag (65,536 members) for an ncu launch list."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from slam_duckietown_b200 import events_io as io, monte_carlo as mc
lines, dets = bench._surrogate()
wins = io.windows_from_events([io.parse_event_line(x) for x in lines], dets)
rs = np.random.default_rng(5)
B = 65536
mv = np.exp(rs.uniform(np.log(0.05), np.log(1.0), B)); mm = np.exp(rs.uniform(np.log(0.05), np.log(1.5), B))
out = mc.monte_carlo_replay(wins, B, 16, 6, mv, mm, det_sigma=0.02, seed=7)
print(out["steps"], float(np.median(out["rmse"])))
