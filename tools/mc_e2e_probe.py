#!/usr/bin/env python
"""Diagnostic: where the end-to-end Monte Carlo step loses against the device-resident step at the PER-RANK size of
configuration M on 8 GPUs (131,072 filters x 8 lm) and at the 1-GPU size: array-fed step, generated step (device record),
pipelined host entry point, host time per call. One GPU. Usage: python tools/mc_e2e_probe.py [batch ...]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl

capi.lib()
L, m, K = 8, 2, 200
sig = (0.01, 0.003, 0.02, 0.02)
for B in [int(x) for x in sys.argv[1:]] or [131072, 1 << 20]:
    w = wl.make_batched(B, L, m, steps=-(-L // m) + 2, perturb_seed=2)
    f = eb.BatchedEKF(B, L)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(f.dev)
    for t in range(w.warm_steps):
        f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
    t = w.warm_steps
    args = (up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
    rec_np = f.mc_record(w.ang[t, 0], w.lin[t, 0], w.slot[t, 0], w.rng[t, 0], w.brg[t, 0])
    rec_dev = torch.from_numpy(rec_np).to(f.dev)
    rec_pin = torch.from_numpy(rec_np.copy()).pin_memory()
    mom = torch.zeros((2, 12), dtype=torch.float64).pin_memory()

    def loop(fn, sync=None, n=K):
        for i in range(10):
            fn(i)
        (sync or torch.cuda.synchronize)()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        e0.record()
        for i in range(n):
            fn(i)
        host = time.perf_counter() - h0
        if sync:
            sync()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - h0
        return {"gpu_us": 1e3 * e0.elapsed_time(e1) / n, "host_issue_us": 1e6 * host / n, "wall_us": 1e6 * wall / n}

    out = {"batch": B,
           "step_array_fed": loop(lambda i: f.step(*args)),
           "step_generated_device_record": loop(lambda i: f.step_generated(7, 100 + i, rec_dev, m, sig)),
           "generated_plus_pose_moments": loop(lambda i: (f.step_generated(7, 400 + i, rec_dev, m, sig), f.pose_moments())),
           "mc_step_host_async": loop(lambda i: f.mc_step_host_async(7, 700 + i, rec_pin, m, sig, pose_moments_out=mom[i & 1]),
                                      sync=f.host_sync)}
    print(json.dumps(out), flush=True)
    del f
    torch.cuda.empty_cache()
