#!/usr/bin/env python
"""Diagnostic: one rank's share of the row-sharded 4096-landmark update on ONE GPU — a row block of 1/8 (or 1/4) of the
rows, fed from a replicated 5-row buffer (no exchange) — for the tiling variants of the streaming kernel
(EKF_LM_VARIANT) and the single-launch form (EKF_LM_SINGLE). Prints us per update, back to back."""
import json, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def run(world):
    import torch
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    from slam_duckietown_b200.sharded import row_partition
    L = 4096
    w = wl.make_large_map(L, updates=130)
    lo, hi = row_partition(L, world)[0]
    f = eb.LargeMapEKF(L, row_range=(lo, hi))
    f.set_noise(0.1, 0.7)
    a = torch.from_numpy(w.factor).cuda()
    f.P[:, :f.n] = a[lo:hi] @ a.T / 64.0
    f.P[torch.arange(hi - lo), torch.arange(lo, hi)] += 0.1
    f.mu[:f.n] = torch.from_numpy(w.mu0).cuda()
    ups = [o for o in w.ops if o[0] == "update" and lo <= 3 + 2 * o[1] < hi - 1][:60] or None
    ups = ups or [o for o in w.ops if o[0] == "update"][:60]
    # replicated buffer: this shard owns rows 0-2; use its own rows for j too when it owns them, else zeros are fine for timing
    for o in ups[:10]:
        f.export_rows(o[1] if lo <= 3 + 2 * o[1] < hi - 1 else 0)
        f.update_from_rows(o[1], o[2], o[3])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for o in ups:
        f.update_from_rows(o[1], o[2], o[3])
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / len(ups), hi - lo

if __name__ == "__main__":
    if len(sys.argv) > 1:
        us, rows = run(int(sys.argv[1]))
        print(json.dumps({"world": int(sys.argv[1]), "rows": rows, "variant": os.environ.get("EKF_LM_VARIANT", "0"),
                          "single": os.environ.get("EKF_LM_SINGLE", "1"), "us_per_update": us}))
    else:
        for world in (8, 4):
            for single in ("1", "0"):
                for var in ("0", "1", "2", "3", "5"):
                    env = dict(os.environ, EKF_LM_VARIANT=var, EKF_LM_SINGLE=single)
                    out = subprocess.run([sys.executable, __file__, str(world)], env=env, capture_output=True, text=True)
                    print(out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:], flush=True)
