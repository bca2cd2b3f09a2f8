#!/usr/bin/env python
"""Large-map single-filter EKF-SLAM update, row-block sharded across the GPUs of one box (BASELINE.json
configs[2] at 1 GPU, configs[4] at 2/4/8 GPUs). Secondary bench; the driver's contract line is bench.py's.

    python bench_large_map.py --landmarks 4096                                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29533 bench_large_map.py --landmarks 32768 --updates 50       # 34 GB of P over 8 GPUs
    ... --check      # additionally: sharded block == the same rows of an unsharded filter, BITWISE (n <= 8195)

One JSON line from rank 0: ms/update (CUDA events, max over ranks; export + NCCL row broadcast + gather +
streaming update all inside the timed region), the HBM roofline of the streaming update per GPU, and the
share of the step spent in the exchange.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--landmarks", type=int, default=4096)
    ap.add_argument("--updates", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=6)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--fused", type=int, default=3, help="also time fused steps of this many updates (0 = skip)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="row exchange: owners push over NVLink peer memory (default) or two NCCL broadcasts")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl
    from slam_duckietown_b200.sharded import ShardedLargeMapEKF, owner_of_slot
    capi.lib()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    L = args.landmarks
    n = 3 + 2 * L
    w = wl.make_large_map(L, updates=args.updates + args.warmup)      # same seeds on every rank
    f = ShardedLargeMapEKF(L, rank, world, device=local, exchange=args.exchange)
    f.set_noise(0.1, 0.7)
    lo, hi = f.ranges[rank]
    a = torch.from_numpy(w.factor).to(dev)                            # (n, 64): P0 = A A^T / 64 + 0.1 I
    blk = a[lo:hi] @ a.T / 64.0                                        # this rank's rows only (never 34 GB at once)
    blk[torch.arange(hi - lo, device=dev), torch.arange(lo, hi, device=dev)] += 0.1
    f.local.P[:, :n] = blk
    f.local.mu[:n] = torch.from_numpy(w.mu0).to(dev)
    del blk
    ref = None
    if args.check:
        if n > 8195:
            raise SystemExit("--check keeps a full unsharded copy per GPU: use <= 4096 landmarks")
        ref = eb.LargeMapEKF(L, device=local)
        ref.set_noise(0.1, 0.7)
        full = a @ a.T / 64.0
        full.diagonal().add_(0.1)
        ref.P[:, :n] = full
        ref.mu[:n] = torch.from_numpy(w.mu0).to(dev)
        del full
    del a

    ups = [o for o in w.ops if o[0] == "update"]

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for o in ups[:args.warmup]:
        f.update(o[1], o[2], o[3])
        if ref:
            ref.update(o[1], o[2], o[3])
    sync()
    timed = ups[args.warmup:]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for o in timed:
        f.update(o[1], o[2], o[3])
    e1.record()
    sync()
    ms = e0.elapsed_time(e1) / len(timed)
    if ref:
        for o in timed:
            ref.update(o[1], o[2], o[3])
    # the NCCL exchange alone (export + two broadcasts), same slots, for comparison
    x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    from slam_duckietown_b200.sharded import exchange_rows
    x0.record()
    for o in timed:
        f.local.export_rows(o[1])
        exchange_rows(f.rows5, owner_of_slot(o[1], L, world))
    x1.record()
    sync()
    xms = x0.elapsed_time(x1) / len(timed)
    # fused steps: --fused updates with ONE exchange and ONE pass over every row block (ekf_lm_*update_multi)
    fms = 0.0
    if args.fused >= 2:
        groups = [timed[i:i + args.fused] for i in range(0, len(timed) - args.fused + 1, args.fused)]
        for g3 in groups[:2]:
            f.update_multi([o[1] for o in g3], [o[2] for o in g3], [o[3] for o in g3])
        sync()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for g3 in groups[2:]:
            f.update_multi([o[1] for o in g3], [o[2] for o in g3], [o[3] for o in g3])
        f1.record()
        sync()
        fms = f0.elapsed_time(f1) / max(len(groups) - 2, 1)
        if ref:
            for g3 in groups:
                for o in g3:
                    ref.update(o[1], o[2], o[3])
    t = torch.tensor([ms, xms, fms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, xms, fms = float(t[0]), float(t[1]), float(t[2])

    check = None
    if ref:
        torch.cuda.synchronize()
        same = bool(torch.equal(f.local.P, ref.P[lo:hi])) and bool(torch.equal(f.local.mu[lo:hi], ref.mu[lo:hi]))
        same &= bool(torch.equal(f.local.mu[:3], ref.mu[:3]))
        c = torch.tensor([1.0 if same else 0.0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(c, op=dist.ReduceOp.MIN)
        check = bool(c.item() == 1.0)
    finite = bool(torch.isfinite(f.local.mu[:n]).all())

    if rank == 0:
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                peak, which = float(json.load(fh)["hbm_gbs"]), "measured"
        except Exception:
            peak, which = 6650.0, "fallback"
        rows_max = max(h - l for l, h in f.ranges)
        per_gpu_bytes = 16 * rows_max * n + 96 * n                    # the largest shard's algorithmic bytes
        ach = per_gpu_bytes / (ms * 1e-3) / 1e9
        print(json.dumps({
            "metric": f"ms/update @{L} lm", "value": ms, "unit": "ms", "n_gpus": world, "higher_is_better": False,
            "updates": len(timed), "warmup": args.warmup, "dtype": "f64", "data": "synthetic",
            "scaling": "strong", "exchange": args.exchange,
            "fused_step": {"updates_per_step": args.fused, "ms_per_step": fms,
                           "ms_per_update_equiv": fms / args.fused} if args.fused >= 2 else None, "nccl_exchange_ms": xms,
            "p2p_error": f.local.p2p_error() if args.exchange == "p2p" else None,
            "config": {"workload": f"single large-map filter, {L} landmarks (n={n}), P {8 * n * n / 1e9:.2f} GB "
                                   f"row-block sharded over {world} GPU(s), one range-bearing update per step",
                       "exchange": ("owners push rows {0,1,2} / {j,j+1} into every rank's buffer over NVLink peer memory "
                                    "(cudaIpc), flag wait, no collective" if args.exchange == "p2p" else
                                    "NCCL broadcast of rows {0,1,2} from rank 0 and {j,j+1} from owner(j)") + ", 40 n bytes"},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "peak_source": which, "unit": "GB/s",
                         "frac": ach / peak, "frac_of_8TBs": ach / 8000.0, "bytes_per_update_per_gpu": per_gpu_bytes,
                         "bytes_per_update_total": 16 * n * n + 96 * n},
            "sharded_equals_unsharded_bitwise": check, "finite": finite}), flush=True)
    f.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
