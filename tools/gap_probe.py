#!/usr/bin/env python
"""Diagnostic: per-launch time of the batched fused step when launches are separated by an idle gap, an L2
flush, or nothing. Usage: python tools/gap_probe.py"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl

B, NLM, M = 65536, 20, 3
capi.lib()
w = wl.make_batched(B, NLM, M, steps=-(-NLM // M) + 8, perturb_seed=2)
f = eb.BatchedEKF(B, NLM, device=0)
f.set_noise(w.motion_var, w.meas_var)
f.init(0.1)
dev = f.dev
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for t in range(w.warm_steps):
    f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
steady = [(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t])) for t in range(w.warm_steps, w.warm_steps + 8)]
scratch = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

def run(n, gap_s=0.0, flush=False, sync_each=False):
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        if flush:
            scratch.zero_()
        if gap_s:
            torch.cuda.synchronize(); time.sleep(gap_s)
        e0[i].record()
        f.step(*steady[i % 8])
        e1[i].record()
        if sync_each:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in zip(e0, e1)])
    return {"mean": round(float(ms[2:].mean()), 4), "min": round(float(ms.min()), 4), "first": [round(float(x), 4) for x in ms[:6]]}

out = {}
out["back_to_back_30"] = run(30)
time.sleep(1.0)
out["sync_each_30"] = run(30, sync_each=True)
time.sleep(1.0)
out["gap_2ms_30"] = run(30, gap_s=0.002)
time.sleep(1.0)
out["gap_20ms_30"] = run(30, gap_s=0.02)
time.sleep(1.0)
out["flush_30"] = run(30, flush=True)
time.sleep(1.0)
out["flush_gap2ms_30"] = run(30, gap_s=0.002, flush=True)
time.sleep(1.0)
out["back_to_back_30_again"] = run(30)
print(json.dumps(out, indent=1))
