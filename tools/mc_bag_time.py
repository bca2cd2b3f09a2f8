#!/usr/bin/env python
"""Diagnostic: wall time of repeated Monte Carlo replays of the surrogate bag (65,536 members), host vs device."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from slam_duckietown_b200 import events_io as io, monte_carlo as mc
lines, dets = bench._surrogate()
wins = io.windows_from_events([io.parse_event_line(x) for x in lines], dets)
rs = np.random.default_rng(5)
B = 65536
mv = np.exp(rs.uniform(np.log(0.05), np.log(1.0), B)); mm = np.exp(rs.uniform(np.log(0.05), np.log(1.5), B))
for i in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = mc.monte_carlo_replay(wins, B, 16, 6, mv, mm, det_sigma=0.02, seed=7)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"pass {i}: returned after {1e3*(t1-t0):.1f} ms, device idle after {1e3*(t2-t0):.1f} ms, rmse median {float(np.median(out['rmse'])):.6f}", flush=True)
