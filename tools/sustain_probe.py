#!/usr/bin/env python
"""Diagnostic: per-launch time of the batched fused step over a long back-to-back run, with nvidia-smi clocks,
power and throttle reasons sampled alongside. Usage: python tools/sustain_probe.py [steps] [gap_ms]"""
import json, os, subprocess, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
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
rows = []
q = "clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown"
proc = subprocess.Popen(["nvidia-smi", "-i", "0", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE, text=True)
threading.Thread(target=lambda: [rows.append((time.perf_counter(), l.strip())) for l in proc.stdout], daemon=True).start()
time.sleep(0.5)
torch.cuda.synchronize()
t_start = time.perf_counter()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
ev[0].record()
for i in range(steps):
    f.step(*steady[i % 8])
    ev[i + 1].record()
torch.cuda.synchronize()
t_end = time.perf_counter()
time.sleep(0.3)
proc.terminate()
ms = np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(steps)])
out = {"steps": steps, "mean_ms": float(ms.mean()), "first20": [round(float(x), 4) for x in ms[:20]],
       "by_chunk": [round(float(ms[i:i + steps // 20].mean()), 4) for i in range(0, steps, steps // 20)],
       "min": float(ms.min()), "smi_during": [r for t, r in rows if t_start <= t <= t_end][:60],
       "smi_before": [r for t, r in rows if t < t_start][-3:]}
print(json.dumps(out, indent=1))
