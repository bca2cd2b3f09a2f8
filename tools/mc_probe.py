#!/usr/bin/env python
"""Diagnostic: device time of the pieces of the generated Monte Carlo step at configuration M (1,048,576 x 8 lm):
perturbation generator, fused step, pose moments, ATE — each alone (CUDA events, 20 launches)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl

B, L, m = 1 << 20, 8, 2
capi.lib()
w = wl.make_batched(B, L, m, steps=-(-L // m) + 2, perturb_seed=2)
f = eb.BatchedEKF(B, L)
f.set_noise(w.motion_var, w.meas_var)
f.init(0.1)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(f.dev)
for t in range(w.warm_steps):
    f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
t = w.warm_steps
args = (up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
rec = torch.from_numpy(f.mc_record(w.ang[t, 0], w.lin[t, 0], w.slot[t, 0], w.rng[t, 0], w.brg[t, 0])).to(f.dev)
sse = torch.zeros(B, dtype=torch.float64, device=f.dev)
sig = (0.01, 0.003, 0.02, 0.02)

def timeit(fn, n=20):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.mean([a.elapsed_time(b) for a, b in ev]))

out = {"step_ms": timeit(lambda: f.step(*args)),
       "perturb_ms": timeit(lambda: f.perturb(7, 3, rec, m, sig)),
       "pose_moments_ms": timeit(lambda: f.pose_moments()),
       "ate_ms": timeit(lambda: f.accumulate_ate(0.1, 0.2, sse)),
       "moments_full_ms": timeit(lambda: f.moments())}
print(json.dumps(out))
