"""Worker of test_p2p_row_exchange_two_processes: one rank of a row-block sharded large-map filter whose row
exchange runs over CUDA IPC peer memory. Usage: python _p2p_worker.py RANK WORLD PORT DEVICE [joseph|relief]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, port, device = (int(x) for x in sys.argv[1:5])
    joseph = len(sys.argv) > 5 and sys.argv[5] == "joseph"
    relief = 22 if len(sys.argv) > 5 and sys.argv[5] == "relief" else None   # rank 0 owns 22 landmarks fewer
    import torch
    import torch.distributed as dist
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    from slam_duckietown_b200.sharded import ShardedLargeMapEKF
    torch.cuda.set_device(device)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    L = 300
    w = wl.make_large_map(L, updates=20)
    f = ShardedLargeMapEKF(L, rank, world, device=device, exchange="p2p", rank0_fewer=relief)
    ref = eb.LargeMapEKF(L, device=device)
    for g in (f, ref):
        g.set_noise(0.2, 0.5)
        g.set_flags(joseph=joseph)
    p0 = w.p0()
    f.load(w.mu0, p0)
    ref.load(w.mu0, p0)
    pend = []

    def flush():
        if pend:                                   # the step's updates fused: one exchange, one pass per shard
            f.update_multi([o[1] for o in pend], [o[2] for o in pend], [o[3] for o in pend])
            pend.clear()
    for i, op in enumerate(w.ops):
        if op[0] == "update":
            ref.update(op[1], op[2], op[3])
            if i < len(w.ops) // 2:
                f.update(op[1], op[2], op[3])      # first half: one exchange per update
            else:
                pend.append(op)
        else:
            flush()
            f.predict(op[1], op[2])
            ref.predict(op[1], op[2])
    flush()
    torch.cuda.synchronize()
    lo, hi = f.ranges[rank]
    same = bool(torch.equal(f.local.P, ref.P[lo:hi])) and bool(torch.equal(f.local.mu[lo:hi], ref.mu[lo:hi]))
    same = same and bool(torch.equal(f.local.mu[:3], ref.mu[:3]))
    print(json.dumps({"rank": rank, "same": same, "p2p_error": f.local.p2p_error(),
                      "finite": bool(torch.isfinite(f.local.mu).all())}), flush=True)
    dist.barrier()
    f.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
