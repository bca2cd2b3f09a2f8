#!/usr/bin/env python
"""Where a CTA of the group kernel spends its life: runs config M (or a smaller batch) on a library built with
-DEKF_PHASE_CLOCKS (clock64 at the phase boundaries, lane 0 of every CTA, atomically summed) and prints the mean
cycles per phase. Usage: python tools/phase_probe.py [batch] (the instrumented library must be csrc/libekf_b200.so)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
nlm, m = 8, 2
lib = capi.lib()
dbg = lib.ekf_debug_phase_clocks
dbg.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
w = wl.make_batched(batch, nlm, m, steps=-(-nlm // m) + 8, perturb_seed=2)
f = eb.BatchedEKF(batch, nlm)
f.set_noise(w.motion_var, w.meas_var)
f.init(0.1)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(f.dev)  # noqa: E731
for t in range(w.warm_steps):
    f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
steady = [(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t])) for t in range(w.warm_steps, w.steps)]
out = (C.c_ulonglong * 8)()
dbg(out, 1)
for i in range(16):
    f.step(*steady[i % len(steady)])
dbg(out, 1)
n = out[7]
names = ["inputs + mu arrived", "augment + motion model", "P arrived (wait)", "predict covariance", "m updates",
         "streaming pass + mu store"]
tot = sum(out[i] for i in range(6))
print(f"batch {batch}: {n} CTAs, mean CTA life {tot / n:.0f} cycles")
for i, nm in enumerate(names):
    print(f"  {nm:28s} {out[i] / n:9.0f} cycles  {100.0 * out[i] / tot:5.1f} %")
