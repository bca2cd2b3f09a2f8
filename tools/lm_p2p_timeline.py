#!/usr/bin/env python
"""Diagnostic: where the time of one row-sharded large-map update goes (peer-memory exchange inside the update's
kernels). Runs under torchrun with the DIAGNOSTIC library (make -C slam-duckietown_b200/csrc timeline), whose kernels
leave globaltimer stamps per update; prints the mean phase durations per role (rank 0 / owner of rows j, j+1 / other).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/lm_p2p_timeline.py [L] [updates]

Stamps (per update, per rank): 0 gather entry | 1 owner: acks seen, push starts | 2 owner: stores issued (flag protocol of the
earlier build: flags published) | 3 rows arrived (cells polled) | 4 scalar chain done (S^-1 in shared memory) | 5 last gather CTA done |
6 streaming kernel entry | 7 last streaming CTA done | 8 rank 0: pre-push of the robot rows issued by column tile 0.
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from slam_duckietown_b200 import capi
    capi.LIB_PATH = os.path.join(os.path.dirname(capi.LIB_PATH), "libekf_b200_tl.so")
    from slam_duckietown_b200 import workloads as wl
    from slam_duckietown_b200.sharded import ShardedLargeMapEKF
    import bench
    L = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    timed_n = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    warm = 20
    w = wl.make_large_map(L, updates=warm + timed_n)
    f = ShardedLargeMapEKF(L, rank, world, device=local, exchange="p2p")
    f.set_noise(0.1, 0.7)
    bench._seed_shard(f, w, dev)
    ups = [o for o in w.ops if o[0] == "update"][:warm + timed_n]
    lib = capi.lib()
    lib.ekf_lm_debug_timeline.restype = C.c_int
    lib.ekf_lm_debug_timeline.argtypes = [C.c_void_p, C.c_int]
    for o in ups[:warm]:
        f.update(o[1], o[2], o[3])
    dist.barrier()
    torch.cuda.synchronize()
    assert lib.ekf_lm_debug_timeline(None, 1) == 0
    timed = ups[warm:]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for o in timed:
        f.update(o[1], o[2], o[3])
    e1.record()
    host_us = 1e6 * (time.perf_counter() - t0) / len(timed)
    torch.cuda.synchronize()
    dist.barrier()
    gpu_us = 1e3 * e0.elapsed_time(e1) / len(timed)
    tl = np.zeros((256, 10), dtype=np.uint64)
    assert lib.ekf_lm_debug_timeline(tl.ctypes.data, 0) == 0
    rows = tl[tl[:, 0] > 0]
    rows = rows[np.argsort(rows[:, 0])].astype(np.int64)
    assert len(rows) == len(timed), (len(rows), len(timed))
    owners = [f.owner(o[1]) for o in timed]
    rec = {"rank": rank, "world": world, "L": L, "host_us_per_call": host_us, "gpu_us_per_update": gpu_us,
           "p2p_error": f.local.p2p_error(), "rows": [hi - lo for lo, hi in f.ranges], "owners": owners, "stamps": rows.tolist()}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"tl_{world}_{L}_rank{rank}.json"), "w") as fh:
        json.dump(rec, fh)
    dist.barrier()
    if rank == 0:
        summarise(world, L)
    dist.destroy_process_group()


def summarise(world, L):
    recs = [json.load(open(os.path.join(ROOT, "gpurun_out", f"tl_{world}_{L}_rank{r}.json"))) for r in range(world)]
    print(f"== world {world}, {L} landmarks: gpu us/update {max(r['gpu_us_per_update'] for r in recs):.2f}, "
          f"host us/call {np.mean([r['host_us_per_call'] for r in recs]):.2f}, "
          f"p2p_error {[hex(r['p2p_error']) for r in recs if r['p2p_error']]}, rows per rank {recs[0]['rows']}")
    phases = {}

    def add(role, name, v):
        phases.setdefault(role, {}).setdefault(name, []).append(v / 1e3)

    for r in recs:
        s = np.array(r["stamps"], dtype=np.int64)
        for k in range(3, len(s) - 1):
            owner = r["owners"][k]
            role = "rank0+owner" if (r["rank"] == 0 and owner == 0) else "rank0" if r["rank"] == 0 else \
                "owner" if r["rank"] == owner else "other"
            g0, g1, g2, g3, g4, g5, s0, s1, pp = s[k][:9]
            add(role, "period (gather entry -> next gather entry)", s[k + 1][0] - g0)
            add(role, "prev stream end -> gather entry", g0 - s[k - 1][7])
            if g1:
                add(role, "owner: entry -> acks seen", g1 - g0)
                add(role, "owner: push + publish", g2 - g1)
            add(role, "entry -> rows arrived", g3 - g0)
            add(role, "scalar chain", g4 - g3)
            add(role, "gather tail (dots, stores, last CTA)", g5 - g4)
            add(role, "gather end -> stream entry", s0 - g5)
            add(role, "stream kernel", s1 - s0)
            if pp:
                add(role, "rank 0: stream entry -> pre-push published", pp - s0)
    for role, d in phases.items():
        print(f"-- {role}")
        for name, v in d.items():
            print(f"   {name:48s} mean {np.mean(v):7.2f} us   median {np.median(v):7.2f}   max {np.max(v):7.2f}   (n={len(v)})")


if __name__ == "__main__":
    main()
