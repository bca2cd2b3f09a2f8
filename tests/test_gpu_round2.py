"""GPU parity tests added in round 2: long horizons on the DISPATCHED batched kernels, full-size configuration M,
the pipelined host entry point, the literal reference `replay()` over the CUDA core, association overflow,
the non-finite scan, and the XL "tracked rows" leg of the large-map oracle.

Tolerance 1e-9 (BASELINE.json north star) on max-norm mu, max-norm P and covariance-scaled P; integer bookkeeping
bit-exact."""
import os

import numpy as np
import pytest

from conftest import TOL, assert_parity, golden
from oracle import ekf_oracle as eo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eb():
    import torch
    assert torch.cuda.is_available()
    from slam_duckietown_b200 import capi, ekf_bindings
    capi.lib()          # fail loudly if the extension is missing
    return ekf_bindings


def _dev(t, a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(t)


# ---------------------------------------------------------------------------------------------------------
# long horizon on the dispatched kernels (VERDICT r1 "what's weak" #1): 256 filters x 256 steps vs the oracle
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("label,batch,L,m", [("B", 65536, 20, 3), ("M", 8192, 8, 2)])
def test_batched_long_horizon_vs_oracle(eb, label, batch, L, m):
    """Config B at its FULL batch (65,536 x 20 lm, the bench's generator and seed) and config M's record shape
    (8 lm, m = 2; 8,192 filters: the full 1 M x 256-step input set would be 6 GB of host arrays) run 256 consecutive
    steps through the default-dispatched kernel (regw / group8: no EKF_BATCH_KERNEL override). The FIRST 256
    filters are replayed through oracle.step_from_observations and compared at every 32nd step — the recursion the
    home-made fp64 atan2 / sqrt / reciprocal chain could drift in."""
    from slam_duckietown_b200 import workloads as wl
    assert "EKF_BATCH_KERNEL" not in os.environ
    T, NB = 256, 256
    w = wl.make_batched(batch, L, m, steps=T, perturb_seed=2)
    n = 3 + 2 * L
    f = eb.BatchedEKF(batch, L)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)
    snaps = {}
    for t in range(T):
        f.step(_dev(f.dev, w.ang[t]), _dev(f.dev, w.lin[t]), _dev(f.dev, w.slot[t]), _dev(f.dev, w.rng[t]),
               _dev(f.dev, w.brg[t]))
        if t % 32 == 31:
            snaps[t] = f.state(0, NB)
    assert f.fault_count() == 0 and f.nonfinite_count() == 0
    assert int(f.n_known.min()) == L
    worst = 0.0
    for b in range(NB):
        m_o = np.zeros(n)
        c_o = np.zeros((n, n))
        c_o[:3, :3] = np.eye(3) * 0.1
        nk = 0
        fl = eo.Flags(motion_var=float(w.motion_var[b]), meas_var=float(w.meas_var[b]))
        for t in range(T):
            obs = [(int(s), float(r), float(q)) for s, r, q in zip(w.slot[t, b], w.rng[t, b], w.brg[t, b]) if s >= 0]
            m_o, c_o, nk = eo.step_from_observations(m_o, c_o, float(w.ang[t, b]), float(w.lin[t, b]), obs, fl, nk)
            if t in snaps:
                mu, p = snaps[t]
                worst = max(worst, *assert_parity(mu[b], p[b], m_o, c_o, what=f"config {label} t={t} b={b}"))
    print(f"config {label}: 256 filters x 256 steps, worst parity {worst:.3e}")


def test_config_m_full_size_properties(eb):
    """1,048,576 x 8 landmarks through the dispatched group kernel WITH its L2 prefetch (launch-order dependent):
    filters fed identical inputs must end bitwise identical wherever they sit in the batch (first / middle / last
    tile and across the prefetch distance), the first 32 must match the reference golden, the pad stays zero, the
    diagonal positive, nothing non-finite."""
    import torch
    z = golden("batched_b32_l8.npz")
    T, B0, m = z["slot"].shape
    B = 1 << 20
    reps = B // B0

    def tile(a):
        return np.concatenate([a] * reps, axis=0).copy()
    f = eb.BatchedEKF(B, 8)
    f.set_noise(tile(z["motion_var"]), tile(z["meas_var"]))
    f.init(0.1)
    for t in range(T):
        f.step(_dev(f.dev, tile(z["ang"][t])), _dev(f.dev, tile(z["lin"][t])), _dev(f.dev, tile(z["slot"][t])),
               _dev(f.dev, tile(z["rng"][t])), _dev(f.dev, tile(z["brg"][t])))
    assert f.fault_count() == 0 and f.nonfinite_count() == 0
    P = f.P.view(B, 19, 20)
    mu = f.mu
    for tile_idx in (1, 777 // B0 + 1, reps // 2, reps - 1):
        lo = tile_idx * B0
        assert torch.equal(P[:B0], P[lo:lo + B0]) and torch.equal(mu[:B0], mu[lo:lo + B0]), tile_idx
    # every tile at once: fold the batch and compare against tile 0
    assert bool((P.view(reps, B0, 19, 20) == P[:B0].unsqueeze(0)).all())
    mu0, p0 = f.state(0, B0)
    i = len(z["keep_t"]) - 1
    for b in range(B0):
        assert_parity(mu0[b], p0[b], z["mean"][i, b], z["cov"][i, b], what=f"config M full size b={b}")
    assert bool((P[:, :, 19] == 0).all())
    assert bool((torch.diagonal(P[:, :, :19], dim1=1, dim2=2) > 0).all())


# ---------------------------------------------------------------------------------------------------------
# host entry points: ordering against the caller's stream, staging-set reuse
# ---------------------------------------------------------------------------------------------------------
def test_host_step_pipelined_matches_device_steps(eb):
    """init() on the caller's stream followed AT ONCE by 7 pipelined step_host_async calls (both staging sets reused
    three times), then state() on the caller's stream without an explicit device synchronise in between: poses and
    state must be bitwise those of the same steps through ekf_batch_step. Then the synchronous step_host."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    B, L, m, T = 20000, 20, 3, 8
    w = wl.make_batched(B, L, m, steps=T)
    ref = eb.BatchedEKF(B, L)
    ref.set_noise(w.motion_var, w.meas_var)
    ref.init(0.1)
    want_pose = []
    for t in range(T):
        ref.step(_dev(ref.dev, w.ang[t]), _dev(ref.dev, w.lin[t]), _dev(ref.dev, w.slot[t]), _dev(ref.dev, w.rng[t]),
                 _dev(ref.dev, w.brg[t]))
        want_pose.append(ref.mu[:, :3].cpu().numpy().copy())
    pin = [tuple(torch.from_numpy(np.ascontiguousarray(a[t])).pin_memory() for a in (w.ang, w.lin, w.slot, w.rng, w.brg))
           for t in range(T)]
    poses = [torch.empty((B, 3), dtype=torch.float64).pin_memory() for _ in range(T)]
    f = eb.BatchedEKF(B, L)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)                                   # enqueued on the caller's stream; no synchronise before the host steps
    for t in range(T - 1):
        f.step_host_async(*pin[t], poses[t])
    mu_mid, p_mid = f.state()                     # caller's stream again: must see step T-2's state (no host_sync yet)
    f.host_sync()
    for t in range(T - 1):
        np.testing.assert_array_equal(poses[t].numpy(), want_pose[t], err_msg=f"pose of pipelined step {t}")
    f.step_host(*pin[T - 1], poses[T - 1])        # synchronous form on top of the pipelined ones
    np.testing.assert_array_equal(poses[T - 1].numpy(), want_pose[T - 1])
    mu, p = f.state()
    mu_r, p_r = ref.state()
    np.testing.assert_array_equal(mu, mu_r)
    np.testing.assert_array_equal(p, p_r)
    assert f.fault_count() == 0
    # the state read between the pipelined steps and host_sync: that of step T-2
    ref2 = eb.BatchedEKF(B, L)
    ref2.set_noise(w.motion_var, w.meas_var)
    ref2.init(0.1)
    for t in range(T - 1):
        ref2.step(_dev(ref2.dev, w.ang[t]), _dev(ref2.dev, w.lin[t]), _dev(ref2.dev, w.slot[t]),
                  _dev(ref2.dev, w.rng[t]), _dev(ref2.dev, w.brg[t]))
    mu2, p2 = ref2.state()
    np.testing.assert_array_equal(mu_mid, mu2)
    np.testing.assert_array_equal(p_mid, p2)


# ---------------------------------------------------------------------------------------------------------
# the literal north-star boundary: the reference's UNCHANGED replay() drives the CUDA core
# ---------------------------------------------------------------------------------------------------------
def test_unchanged_reference_replay_over_cuda_core(eb, surrogate, tmp_path):
    """src/replay_no_ros.py:replay() itself (loaded from oracle/_ref bytecode on the GPU box, from /root/reference
    where that exists) with its module globals EKF_pose_estimation / delta_phi / displacement rebound to
    ekf_bindings: 70 EKF steps on the surrogate bag, each equal to what the reference's own arithmetic produced
    (tests/golden/surrogate/replay_ref.npz)."""
    from oracle import ref_loader
    from slam_duckietown_b200 import replay
    if not ref_loader.reference_available():
        pytest.skip("reference neither at /root/reference nor compiled into oracle/_ref")
    lines, dets, ref = surrogate
    with open(tmp_path / "events.csv", "w") as fh:
        fh.writelines(lines)
    mod = ref_loader.load_replay_module()
    recs = replay.drive_reference_replay(mod, str(tmp_path), dets)
    assert len(recs) == len(ref["n"]) == 70
    worst = 0.0
    for i, r in enumerate(recs):
        n = int(ref["n"][i])
        assert len(r["mean"]) == n
        assert r["order"] == [int(x) for x in ref["order"][i] if x >= 0]
        assert r["ang"] == ref["ang"][i] and r["lin"] == ref["lin"][i]
        worst = max(worst, *assert_parity(r["mean"], r["cov"], ref["mean"][i][:n], ref["cov"][i][:n, :n], what=f"step {i}"))
    assert list(recs[-1]["tag_index"].keys()) == list(ref["tag_ids"])
    assert list(recs[-1]["tag_index"].values()) == list(ref["tag_slots"])
    print(f"unchanged replay() over the CUDA core ({ref_loader.reference_kind()} reference): worst parity {worst:.3e}")


# ---------------------------------------------------------------------------------------------------------
# association overflow and the non-finite scan
# ---------------------------------------------------------------------------------------------------------
def test_association_overflow_keeps_slot_order(eb):
    """A window with MORE distinct new tags than the step block holds: the extra detections are dropped before a slot
    is handed out, so slot order stays gap-free, the dropped tags are first-seen in the next window, the step raises
    no fault, and the overflow is counted."""
    import torch
    B, L, m = 5, 8, 2
    f = eb.BatchedEKF(B, L)
    f.init(0.1)
    tag_slot, n_seen = f.new_slot_table(6)
    tags = np.array([3, 1, 4, 0, 1], dtype=np.int32)          # four distinct new tags, m = 2
    tx = np.array([0.1, -0.2, 0.3, 0.0, -0.21])
    tz = np.array([0.9, 1.0, 0.8, 1.1, 1.01])
    slot, rng, brg = f.associate(_dev(f.dev, tags), _dev(f.dev, tx), _dev(f.dev, tz), tag_slot, n_seen, m)
    assert f.associate_overflow_count() == B
    s = slot.cpu().numpy()
    assert np.all(s == np.array([0, 1]))                       # tags 3, 1 in first-seen order; 4 and 0 dropped
    ts = tag_slot.cpu().numpy()
    assert np.all(ts[3] == 0) and np.all(ts[1] == 1) and np.all(ts[4] == -1) and np.all(ts[0] == -1)
    assert np.all(n_seen.cpu().numpy() == 2)
    z = torch.zeros(B, dtype=torch.float64, device=f.dev)
    f.step(z + 0.05, z + 0.03, slot, rng, brg)
    # next window: the dropped tags come back and get the NEXT slots, in first-seen order
    tags2 = np.array([0, 4], dtype=np.int32)
    slot2, rng2, brg2 = f.associate(_dev(f.dev, tags2), _dev(f.dev, tx[:2]), _dev(f.dev, tz[:2]), tag_slot, n_seen, m)
    assert np.all(slot2.cpu().numpy() == np.array([2, 3]))
    f.step(z + 0.05, z + 0.03, slot2, rng2, brg2)
    assert f.fault_count() == 0
    assert np.all(f.n_known.cpu().numpy() == 4)
    assert f.nonfinite_count() == 0


def test_nonfinite_count_flags_poisoned_filters(eb):
    f = eb.BatchedEKF(1000, 8)
    f.init(0.1)
    assert f.nonfinite_count() == 0
    f.P[17, 5] = float("nan")
    f.mu[400, 2] = float("inf")
    f.P[999, f.p_stride - 1] = float("-inf")
    assert f.nonfinite_count() == 3


# ---------------------------------------------------------------------------------------------------------
# XL leg (iii): the rank-2 restatement on tracked rows (oracle/sampled_rows.py) against row-block shards
# ---------------------------------------------------------------------------------------------------------
def test_tracked_rows_leg_on_emulated_shards(eb):
    """The check bench.py runs at 32,768 landmarks on real shards, here at 4096 landmarks (n = 8195) on 3 emulated
    shards: 10 updates + predicts; 64 sampled rows of every shard plus the gathered rows vs TrackedRows."""
    from oracle import sampled_rows as sr
    from slam_duckietown_b200 import workloads as wl
    from slam_duckietown_b200.sharded import EmulatedShards
    L, world = 4096, 3
    w = wl.make_large_map(L, updates=10)
    sh = EmulatedShards(L, world)
    sh.set_noise(0.1, 0.7)
    p0 = w.p0()
    sh.load(w.mu0, p0)
    slots = [op[1] for op in w.ops if op[0] == "update"]
    rs = np.random.default_rng(9)
    rows = sr.required_rows(slots)
    for lo, hi in sh.ranges:
        rows += [int(r) for r in rs.choice(np.arange(lo, hi), 64, replace=False) if r not in rows]
    tr = sr.TrackedRows(w.mu0, rows, p0[rows, :])
    fl = eo.Flags()
    for op in w.ops:
        if op[0] == "update":
            sh.update(op[1], op[2], op[3])
            tr.update(op[1], op[2], op[3], 0.7)
        else:
            sh.predict(op[1], op[2])
            tr.predict(op[1], op[2], fl)
    mu, p = sh.state()
    e = tr.errors(mu[rows], p[rows, :], np.diag(p))
    print(f"tracked rows on 3 emulated shards, n = {sh.n}: {e}")
    assert max(e) <= TOL, e
    sh.close()


@pytest.mark.parametrize("n_lm,group", [(700, 3), (256, 4), (1024, 2)])
def test_large_map_fused_joseph_bitwise(eb, n_lm, group):
    """Joseph form: ekf_lm_update_multi (one pass over P, two rank-2 terms per update) == the same updates through the
    sequential Joseph kernels, BITWISE, and both match the oracle's Joseph update."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_large_map(n_lm, updates=18)
    seq, fus = eb.LargeMapEKF(n_lm), eb.LargeMapEKF(n_lm)
    m_o, c_o = w.mu0.copy(), w.p0()
    for g in (seq, fus):
        g.set_noise(0.2, 0.5)
        g.set_flags(joseph=True)
        g.load(w.mu0, w.p0())
    fl = eo.Flags(motion_var=0.2, meas_var=0.5, joseph=True)
    pend = []

    def flush():
        if pend:
            fus.update_multi([o[1] for o in pend], [o[2] for o in pend], [o[3] for o in pend])
            pend.clear()
    for op in w.ops:
        if op[0] == "update":
            seq.update(op[1], op[2], op[3])
            eo.update_rank(m_o, c_o, op[1], op[2], op[3], 0.5, joseph=True)
            pend.append(op)
            if len(pend) == group:
                flush()
        else:
            flush()
            for g in (seq, fus):
                g.predict(op[1], op[2])
            dx, dy, th, g0, g1 = eo.motion_step(m_o[2], op[1], op[2], fl)
            m_o[0] += dx; m_o[1] += dy; m_o[2] = th
            eo.predict_cov_rank(c_o, g0, g1, 0.2)
    flush()
    assert torch.equal(seq.mu, fus.mu), "mu differs between sequential and fused Joseph updates"
    assert torch.equal(seq.P, fus.P), "P differs between sequential and fused Joseph updates"
    mu, p = fus.state()
    assert_parity(mu, p, m_o, c_o, what=f"fused Joseph n_lm={n_lm} group={group}")


@pytest.mark.parametrize("L,m,joseph", [(40, 5, False), (56, 3, True)])
def test_batched_large_capacity_vs_oracle(eb, L, m, joseph):
    """Capacities above 28 landmarks (VERDICT r1 missing #9) run on the general staged kernel: 90 filters, every
    landmark initialised then steady state, sampled filters against the oracle."""
    from slam_duckietown_b200 import workloads as wl
    B = 90
    w = wl.make_batched(B, L, m, steps=-(-L // m) + 6)
    f = eb.BatchedEKF(B, L)
    f.set_noise(w.motion_var, w.meas_var)
    f.set_flags(joseph=joseph)
    f.init(0.1)
    for t in range(w.steps):
        f.step(_dev(f.dev, w.ang[t]), _dev(f.dev, w.lin[t]), _dev(f.dev, w.slot[t]), _dev(f.dev, w.rng[t]),
               _dev(f.dev, w.brg[t]))
    assert f.fault_count() == 0 and f.nonfinite_count() == 0 and int(f.n_known.min()) == L
    mu, p = f.state()
    n = 3 + 2 * L
    for b in (0, 31, B - 1):
        m_o, c_o, nk = np.zeros(n), np.zeros((n, n)), 0
        c_o[:3, :3] = np.eye(3) * 0.1
        fl = eo.Flags(motion_var=float(w.motion_var[b]), meas_var=float(w.meas_var[b]), joseph=joseph)
        for t in range(w.steps):
            obs = [(int(s), float(r), float(q)) for s, r, q in zip(w.slot[t, b], w.rng[t, b], w.brg[t, b]) if s >= 0]
            m_o, c_o, nk = eo.step_from_observations(m_o, c_o, float(w.ang[t, b]), float(w.lin[t, b]), obs, fl, nk)
        assert_parity(mu[b], p[b], m_o, c_o, what=f"capacity {L} m={m} joseph={joseph} b={b}")
    with pytest.raises(Exception):
        eb.BatchedEKF(4, 36)                                   # not a built capacity: refused loudly
