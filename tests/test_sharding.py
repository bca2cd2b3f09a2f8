"""Row-block sharding of the large-map filter: host logic on CPU (partition, owners, the
per-update row exchange over a world_size-2 gloo group), and — on the GPU — sharded == unsharded BITWISE."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import assert_parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_lm,world", [(4096, 1), (4096, 2), (4096, 8), (32768, 8), (13, 4), (3, 8), (700, 3)])
def test_row_partition_covers_rows_on_landmark_pairs(n_lm, world):
    from slam_duckietown_b200.sharded import owner_of_slot, row_partition
    parts = row_partition(n_lm, world)
    n = 3 + 2 * n_lm
    assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
    for r, (lo, hi) in enumerate(parts):
        assert lo <= hi
        if r:
            assert lo == parts[r - 1][1] and (lo - 3) % 2 == 0 and lo >= 3      # boundary on a landmark pair
    sizes = [(hi - max(lo, 3)) // 2 for lo, hi in parts]
    assert sum(sizes) == n_lm and max(sizes) - min(sizes) <= 1                  # near-equal blocks
    for slot in sorted({0, n_lm - 1, n_lm // 2, n_lm // 3, min(5, n_lm - 1)}):
        o = owner_of_slot(slot, n_lm, world)
        lo, hi = parts[o]
        assert lo <= 3 + 2 * slot and 4 + 2 * slot < hi                          # rows j, j+1 share the owner
    with pytest.raises(ValueError):
        owner_of_slot(n_lm, n_lm, world)


@pytest.mark.parametrize("n_lm,world,fewer", [(4096, 8, 101), (4096, 2, 8), (32768, 8, 101), (300, 3, 22), (13, 4, 5),
                                              (3, 8, 2), (700, 3, 0)])
def test_row_partition_with_rank0_relief(n_lm, world, fewer):
    """Under the peer-memory exchange rank 0 also stores the robot rows to every peer, so it owns fewer landmarks: same
    invariants (cover, landmark-pair boundaries, owner_of_slot consistent), rank 0 below the equal share by `fewer`
    (at least one landmark left), the other ranks near-equal."""
    from slam_duckietown_b200.sharded import owner_of_slot, rank0_relief, row_partition
    parts = row_partition(n_lm, world, fewer)
    n = 3 + 2 * n_lm
    assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
    for r in range(1, world):
        assert parts[r][0] == parts[r - 1][1] and (parts[r][0] - 3) % 2 == 0
    sizes = [(hi - max(lo, 3)) // 2 for lo, hi in parts]
    assert sum(sizes) == n_lm
    if fewer:
        assert sizes[0] == min(n_lm, max(1, n_lm // world - fewer))
        assert max(sizes[1:]) - min(sizes[1:]) <= 1
    for slot in range(0, n_lm, max(1, n_lm // 97)):
        lo, hi = parts[owner_of_slot(slot, n_lm, world, fewer)]
        assert lo <= 3 + 2 * slot and 4 + 2 * slot < hi
    assert owner_of_slot(n_lm - 1, n_lm, world, fewer) == max(r for r in range(world) if sizes[r] > 0)
    # relief only where the row blocks stream from HBM (32 768 landmarks), not for L2-resident blocks (4096 on 8)
    assert rank0_relief(32768, 8) == 101 and rank0_relief(4096, 8) == 0 and rank0_relief(32768, 1) == 0


def test_p2p_exchange_refuses_ranks_without_rows():
    """Checked before anything touches the GPU or the library: 3 landmarks cannot feed 8 ranks."""
    from slam_duckietown_b200.sharded import ShardedLargeMapEKF
    with pytest.raises(ValueError, match="at least one landmark per rank"):
        ShardedLargeMapEKF(3, 0, 8, device=0, exchange="p2p")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, n_lm, slots, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from slam_duckietown_b200.sharded import exchange_rows, owner_of_slot, row_partition
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n = 3 + 2 * n_lm
    ld = (n + 15) // 16 * 16
    lo, hi = row_partition(n_lm, world)[rank]
    # the "covariance" every rank could see if it owned everything; a rank only exports rows it owns
    full = torch.arange(n * ld, dtype=torch.float64).reshape(n, ld)
    ok = True
    for slot in slots:
        j = 3 + 2 * slot
        rows5 = torch.full((5, ld + 4), -1.0 - rank, dtype=torch.float64)
        for k, row in enumerate((0, 1, 2, j, j + 1)):
            if lo <= row < hi:                        # what ekf_lm_export_rows does on the device
                rows5[k, :ld] = full[row]
                rows5[k, ld] = 1000.0 + row           # the mu entry riding along
        exchange_rows(rows5, owner_of_slot(slot, n_lm, world))
        for k, row in enumerate((0, 1, 2, j, j + 1)):
            ok &= bool(torch.equal(rows5[k, :ld], full[row])) and float(rows5[k, ld]) == 1000.0 + row
    # the fused step's exchange: rows {0,1,2} U {j_u, j_u+1} with one owner per landmark pair
    from slam_duckietown_b200.sharded import exchange_rows_multi
    for group in (slots[:2], slots[1:4], slots[1:5]):
        rows = torch.full((3 + 2 * len(group), ld + 4), -1.0 - rank, dtype=torch.float64)
        want = [0, 1, 2] + [r for s_ in group for r in (3 + 2 * s_, 4 + 2 * s_)]
        for k, row in enumerate(want):
            if lo <= row < hi:                        # what ekf_lm_export_rows_multi does on the device
                rows[k, :ld] = full[row]
                rows[k, ld] = 1000.0 + row
        exchange_rows_multi(rows, [owner_of_slot(s_, n_lm, world) for s_ in group])
        for k, row in enumerate(want):
            ok &= bool(torch.equal(rows[k, :ld], full[row])) and float(rows[k, ld]) == 1000.0 + row
    # the bench's max-over-ranks reduction
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok &= float(t) == float(world)
    dist.barrier()
    dist.destroy_process_group()
    with open(os.path.join(out_dir, f"ok{rank}"), "w") as fh:
        fh.write("1" if ok else "0")


def _gloo_moments_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from slam_duckietown_b200.ensemble import allreduce_moments, filter_partition, moments_to_mean_cov
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    batch, n = 101, 7
    x = np.random.default_rng(5).normal(size=(batch, n)) + np.arange(n)          # the whole ensemble, every rank
    lo, hi = filter_partition(batch, world)[rank]
    mine = x[lo:hi]                                                               # what ekf_batch_moments sums on the device
    sums = torch.from_numpy(np.concatenate([mine.sum(0), (mine[:, :, None] * mine[:, None, :]).sum(0).ravel()]))
    count, total = allreduce_moments(hi - lo, sums)
    mean, cov = moments_to_mean_cov(count, total, n)
    ok = count == batch
    ok &= bool(np.allclose(mean.numpy(), x.mean(0), rtol=1e-13, atol=1e-13))
    ok &= bool(np.allclose(cov.numpy(), np.cov(x.T, bias=True), rtol=1e-11, atol=1e-12))
    dist.barrier()
    dist.destroy_process_group()
    with open(os.path.join(out_dir, f"mom{rank}"), "w") as fh:
        fh.write("1" if ok else "0")


@pytest.mark.parametrize("batch,world", [(65536, 8), (1 << 20, 4), (101, 2), (3, 8), (0, 2)])
def test_filter_partition_is_contiguous_and_balanced(batch, world):
    from slam_duckietown_b200.ensemble import filter_partition
    parts = filter_partition(batch, world)
    assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == batch
    assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
    sizes = [e - b for b, e in parts]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_ensemble_moments_over_gloo_world2(tmp_path):
    """world_size 2, gloo, CPU: per-rank moment sums + one all_reduce give the whole ensemble's mean / covariance."""
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_gloo_moments_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert [open(tmp_path / f"mom{r}").read() for r in range(world)] == ["1", "1"]


def test_row_exchange_over_gloo_world2(tmp_path):
    """world_size 2, gloo, CPU: after the exchange every rank holds the owners' five rows and mu entries."""
    import torch.multiprocessing as mp
    n_lm, world = 37, 2
    slots = [0, 5, 18, 19, 36]                        # owners 0,0,0,1,1 (19 landmarks on rank 0)
    mp.spawn(_gloo_worker, args=(world, _free_port(), n_lm, slots, str(tmp_path)), nprocs=world, join=True)
    assert [open(tmp_path / f"ok{r}").read() for r in range(world)] == ["1", "1"]


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_equals_unsharded_bitwise(world):
    """Every shard's block after 30 ops (updates + predicts) is bit-identical to the same rows of the single-GPU
    filter, which is itself oracle-checked (test_gpu_parity). All shards run on this one GPU."""
    from oracle import ekf_oracle as eo
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    from slam_duckietown_b200.sharded import EmulatedShards
    n_lm = 300
    w = wl.make_large_map(n_lm, updates=30)
    single = eb.LargeMapEKF(n_lm)
    single.set_noise(0.1, 0.7)
    single.load(w.mu0, w.p0())
    sh = EmulatedShards(n_lm, world)
    sh.set_noise(0.1, 0.7)
    sh.load(w.mu0, w.p0())
    m_o, c_o = w.mu0.copy(), w.p0()
    for op in w.ops:
        if op[0] == "update":
            single.update(op[1], op[2], op[3])
            sh.update(op[1], op[2], op[3])
            eo.update_rank(m_o, c_o, op[1], op[2], op[3], 0.7)
        else:
            single.predict(op[1], op[2])
            sh.predict(op[1], op[2])
            dx, dy, th, g0, g1 = eo.motion_step(m_o[2], op[1], op[2], eo.Flags())
            m_o[0] += dx; m_o[1] += dy; m_o[2] = th
            eo.predict_cov_rank(c_o, g0, g1, 0.1)
    mu1, p1 = single.state()
    mu2, p2 = sh.state()
    np.testing.assert_array_equal(p2, p1)
    np.testing.assert_array_equal(mu2, mu1)
    assert_parity(mu2, p2, m_o, c_o, what=f"sharded x{world}")
    # every shard keeps the replicated robot pose current
    for s in sh.shards:
        np.testing.assert_array_equal(s.state()[0][:3], mu1[:3])
    sh.close()
    single.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world,group", [(2, 3), (3, 2), (8, 4)])
def test_sharded_fused_multi_update_equals_unsharded_bitwise(world, group):
    """The fused step (m updates, ONE exchange of the rows {0,1,2} U {j_u, j_u+1}, ONE pass over each row block) on
    row-block shards == the unsharded fused step == the sequential single updates, bitwise; every shard keeps the
    replicated pose current."""
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    from slam_duckietown_b200.sharded import EmulatedShards
    n_lm = 300
    w = wl.make_large_map(n_lm, updates=24)
    seq, fus = eb.LargeMapEKF(n_lm), eb.LargeMapEKF(n_lm)
    sh = EmulatedShards(n_lm, world)
    for g in (seq, fus, sh):
        g.set_noise(0.1, 0.7)
        g.load(w.mu0, w.p0())
    pend = []

    def flush():
        if pend:
            args = ([o[1] for o in pend], [o[2] for o in pend], [o[3] for o in pend])
            fus.update_multi(*args)
            if len(pend) == 1:
                sh.update(pend[0][1], pend[0][2], pend[0][3])
            else:
                sh.update_multi(*args)
            pend.clear()
    for op in w.ops:
        if op[0] == "update":
            seq.update(op[1], op[2], op[3])
            pend.append(op)
            if len(pend) == group:
                flush()
        elif group <= 3:
            flush()
            for g in (seq, fus, sh):
                g.predict(op[1], op[2])
    flush()
    mu0, p0 = seq.state()
    mu1, p1 = fus.state()
    mu2, p2 = sh.state()
    np.testing.assert_array_equal(p1, p0)
    np.testing.assert_array_equal(mu1, mu0)
    np.testing.assert_array_equal(p2, p0)
    np.testing.assert_array_equal(mu2, mu0)
    for s in sh.shards:
        np.testing.assert_array_equal(s.state()[0][:3], mu0[:3])
    sh.close()
    seq.close()
    fus.close()
