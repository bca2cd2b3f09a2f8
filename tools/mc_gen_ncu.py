#!/usr/bin/env python
"""Diagnostic: a few generated Monte Carlo steps at configuration M (1,048,576 x 8 lm) for an ncu capture of the GEN
instantiation of the group kernel."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl
capi.lib()
B, L, m = 1 << 20, 8, 2
w = wl.make_batched(B, L, m, steps=-(-L // m) + 2, perturb_seed=2)
f = eb.BatchedEKF(B, L)
f.set_noise(w.motion_var, w.meas_var)
f.init(0.1)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(f.dev)
for t in range(w.warm_steps):
    f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
t = w.warm_steps
rec = torch.from_numpy(f.mc_record(w.ang[t, 0], w.lin[t, 0], w.slot[t, 0], w.rng[t, 0], w.brg[t, 0])).to(f.dev)
for i in range(8):
    f.step_generated(7, 100 + i, rec, m, (0.01, 0.003, 0.02, 0.02))
torch.cuda.synchronize()
print("ok")
