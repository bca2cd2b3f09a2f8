"""The counter-based Monte Carlo perturbation generator: numpy twin (oracle/mc_noise.py) on CPU — Random123 known
answers, accuracy of the single-precision Box-Muller — and, on the GPU, the device generator == the twin BIT FOR BIT, per-member
parity of a generated Monte Carlo replay against the oracle fed the same generated numbers, and invariance of every
member under re-sharding of the ensemble."""
import numpy as np
import pytest

from conftest import assert_parity
from oracle import ekf_oracle as eo, mc_noise as mn


def test_philox4x32_10_known_answers():
    """Random123 kat_vectors (philox4x32, 10 rounds)."""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = mn.philox4x32_10(*ctr, *key)
        assert tuple(int(x) for x in got) == want


def test_uniforms_strictly_inside_unit_interval():
    u = mn.uniform23(np.array([0, 0xffffffff, 0x1ff, 0x200], dtype=np.uint32))
    assert u.dtype == np.float32
    assert 0.0 < u[0] == 2.0 ** -24 and u[1] == 1.0 - 2.0 ** -24 < 1.0 and u[2] == u[0] and u[3] == 1.5 * 2.0 ** -23
    # the radial uniform: a second word refines the lowest cell, everything else is uniform23
    t = mn.uniform23_tail(np.array([0, 0x1ff, 0x200, 0xffffffff], dtype=np.uint32),
                          np.array([0, 0xffffffff, 123, 456], dtype=np.uint32))
    assert t.dtype == np.float32
    assert t[0] == 2.0 ** -47 and t[1] == (2.0 ** 23 - 0.5) * 2.0 ** -46 < 2.0 ** -23 and t[2] == u[3] and t[3] == u[1]


def test_log_and_sincos_accuracy():
    """Single precision is the design (noise deviates): ~2e-7 relative, and every value a float32."""
    x = np.concatenate([np.logspace(-14.1, 0, 5000), np.linspace(1e-3, 1.0, 5000), [2.0 ** -47, 1 - 2.0 ** -24]])
    x = x.astype(np.float32)
    x64 = x.astype(np.float64)
    got = mn.det_log(x)
    assert got.dtype == np.float32
    assert np.max(np.abs(got - np.log(x64)) / np.maximum(np.abs(np.log(x64)), 1e-3)) < 5e-7
    t = np.concatenate([np.linspace(0, 1, 200001)[:-1], [2.0 ** -24, 1 - 2.0 ** -24, 0.125, 0.25, 0.375, 0.5, 0.625, 0.75, 0.875]])
    t = t.astype(np.float32)
    s, c = mn.sincos_2pi(t)
    t64 = t.astype(np.float64)
    assert s.dtype == np.float32 and c.dtype == np.float32
    assert np.max(np.abs(s - np.sin(2 * np.pi * t64))) < 3e-7 and np.max(np.abs(c - np.cos(2 * np.pi * t64))) < 3e-7
    assert np.max(np.abs(s.astype(np.float64) ** 2 + c.astype(np.float64) ** 2 - 1.0)) < 5e-7


def test_box_muller_reaches_the_tails():
    u1 = np.array([2.0 ** -47, 0.5, 1 - 2.0 ** -24], dtype=np.float32)
    u2 = np.array([2.0 ** -24, 0.25, 0.5], dtype=np.float32)
    z0, z1 = (z.astype(np.float64) for z in mn.box_muller(u1, u2))
    assert abs(np.hypot(z0[0], z1[0]) - np.sqrt(-2 * np.log(2.0 ** -47))) < 1e-5      # 8.07 sigma
    assert abs(z0[1]) < 1e-7 and abs(z1[1] - np.sqrt(2 * np.log(2))) < 1e-6
    assert np.hypot(z0[2], z1[2]) < 4e-4
    # without the refinement the radial uniform would stop at 2^-24: 5.77 sigma
    assert np.sqrt(-2 * np.log(2.0 ** -24)) < 5.8


def test_normal_pairs_look_normal_and_streams_are_independent():
    f = np.arange(400000)
    z0, z1 = mn.normal_pair(11, f, 5, 0)
    y0, _ = mn.normal_pair(11, f, 5, 1)
    w0, _ = mn.normal_pair(11, f, 6, 0)
    for z in (z0, z1, y0, w0):
        assert abs(z.mean()) < 0.01 and abs(z.std() - 1.0) < 0.01
        assert abs(np.mean(z ** 4) - 3.0) < 0.1
    for a, b in ((z0, z1), (z0, y0), (z0, w0)):
        assert abs(np.corrcoef(a, b)[0, 1]) < 0.01
    a0, _ = mn.normal_pair(11, f + (1 << 33), 5, 0)          # the high word of the filter index is part of the counter
    assert not np.array_equal(a0, z0)


@pytest.fixture(scope="module")
def eb():
    import torch
    assert torch.cuda.is_available()
    from slam_duckietown_b200 import capi, ekf_bindings
    capi.lib()
    return ekf_bindings


@pytest.mark.gpu
def test_device_generator_equals_numpy_twin_bitwise(eb):
    import torch
    B, m = 50021, 3
    f = eb.BatchedEKF(B, 8)
    sigma = (0.01, 0.003, 0.02, 0.025)
    for seed, step, first in ((7, 0, 0), (2 ** 40 + 5, 123456, 3 * 2 ** 32 + 17)):
        rec = f.mc_record(0.08, 0.05, [4, -1, 2], [0.9, 0.0, 1.2], [0.3, 0.0, -0.4])
        ang, lin, slot, rng, brg = f.perturb(seed, step, torch.from_numpy(rec).to(f.dev), m, sigma, first)
        a, l_, s_, r_, b_ = mn.perturb_step(seed, first, B, step, 0.08, 0.05, [4, -1, 2], [0.9, 0.0, 1.2], [0.3, 0.0, -0.4], sigma)
        assert np.array_equal(ang.cpu().numpy(), a) and np.array_equal(lin.cpu().numpy(), l_)
        assert np.array_equal(slot.cpu().numpy(), s_)
        assert np.array_equal(rng.cpu().numpy(), r_) and np.array_equal(brg.cpu().numpy(), b_)
        ntx, ntz = f.detection_noise(seed, step, 5, 0.02, first)
        tx, tz = mn.detection_noise(seed, first, B, step, 5, 0.02)
        assert np.array_equal(ntx.cpu().numpy(), tx) and np.array_equal(ntz.cpu().numpy(), tz)


@pytest.mark.gpu
def test_pose_moments_match_numpy_and_are_deterministic(eb):
    import torch
    from slam_duckietown_b200 import workloads as wl
    for B in (37, 70001, 400000):
        w = wl.make_batched(B, 8, 2, steps=3)
        f = eb.BatchedEKF(B, 8)
        f.set_noise(w.motion_var, w.meas_var)
        f.init(0.1)
        for t in range(3):
            f.step(*(torch.from_numpy(np.ascontiguousarray(a[t])).to(f.dev) for a in (w.ang, w.lin, w.slot, w.rng, w.brg)))
        s1, s2 = f.pose_moments().cpu().numpy(), f.pose_moments().cpu().numpy()
        assert np.array_equal(s1, s2)
        pose = f.mu[:, :3].cpu().numpy()
        want = np.concatenate([pose.sum(0), (pose.T @ pose).ravel()])
        assert np.max(np.abs(s1 - want)) <= 1e-12 * max(1.0, np.abs(want).max()) * np.log2(B + 1)
        full = f.moments().cpu().numpy()
        n = f.n
        np.testing.assert_allclose(s1[:3], full[:3], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(s1[3:].reshape(3, 3), full[n:].reshape(n, n)[:3, :3], rtol=1e-12, atol=1e-12)


@pytest.mark.gpu
def test_generated_monte_carlo_replay_per_member_parity_and_sharding_invariance(eb):
    """B members replay one common trajectory through ekf_batch_mc_step_host_async (record H2D, on-device
    perturbations, fused step, ATE, pose moments). (1) sampled members == the oracle fed the numbers the numpy twin
    generates for them; (2) the same ensemble split into two handles with first_filter offsets gives every member the
    identical bits; (3) ATE and the per-step pose moments agree with host recomputation."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    B, L, m, seed = 6000, 8, 2, 20261105
    sigma = (0.01, 0.003, 0.02, 0.02)
    base = wl.make_batched(1, L, m, steps=24)                       # the common trajectory: filter 0's unperturbed inputs
    rs = np.random.default_rng(1)
    mv, mm = np.exp(rs.uniform(np.log(0.05), np.log(1.0), B)), np.exp(rs.uniform(np.log(0.05), np.log(1.5), B))
    gts = [(0.01 * t, -0.005 * t) for t in range(base.steps)]

    def run(lo, hi):
        f = eb.BatchedEKF(hi - lo, L)
        f.set_noise(mv[lo:hi], mm[lo:hi])
        f.init(0.1)
        sse = torch.zeros(hi - lo, dtype=torch.float64, device=f.dev)
        recs = [torch.from_numpy(f.mc_record(base.ang[t, 0], base.lin[t, 0], base.slot[t, 0], base.rng[t, 0],
                                             base.brg[t, 0])).pin_memory() for t in range(base.steps)]
        moms = [torch.zeros(12, dtype=torch.float64).pin_memory() for _ in range(base.steps)]
        for t in range(base.steps):
            f.mc_step_host_async(seed, t, recs[t], m, sigma, first_filter=lo, gt=gts[t], sse=sse, pose_moments_out=moms[t])
        f.host_sync()
        assert f.fault_count() == 0 and f.nonfinite_count() == 0
        mu, p = f.state()
        return mu, p, sse.cpu().numpy(), np.stack([x.numpy().copy() for x in moms]), f
    mu, p, sse, moms, f = run(0, B)
    # (1) per-member parity against the oracle fed the twin's numbers
    n = 3 + 2 * L
    gen = [mn.perturb_step(seed, 0, B, t, base.ang[t, 0], base.lin[t, 0], base.slot[t, 0], base.rng[t, 0],
                           base.brg[t, 0], sigma) for t in range(base.steps)]
    sse_ref = np.zeros(B)
    for b in list(range(0, B, 97)) + [B - 1]:
        m_o, c_o, nk = np.zeros(n), np.zeros((n, n)), 0
        c_o[:3, :3] = np.eye(3) * 0.1
        fl = eo.Flags(motion_var=float(mv[b]), meas_var=float(mm[b]))
        for t in range(base.steps):
            a, l_, s_, r_, b_ = gen[t]
            obs = [(int(s_[b, k]), float(r_[b, k]), float(b_[b, k])) for k in range(m) if s_[b, k] >= 0]
            m_o, c_o, nk = eo.step_from_observations(m_o, c_o, float(a[b]), float(l_[b]), obs, fl, nk)
            sse_ref[b] += (m_o[0] - gts[t][0]) ** 2 + (m_o[1] - gts[t][1]) ** 2
        assert_parity(mu[b], p[b], m_o, c_o, what=f"generated MC member {b}")
        assert abs(sse[b] - sse_ref[b]) <= 1e-9 * max(1.0, sse_ref[b])
    # (3) the last step's pose moments == numpy over the final state
    pose = mu[:, :3]
    want = np.concatenate([pose.sum(0), (pose.T @ pose).ravel()])
    assert np.max(np.abs(moms[-1] - want)) <= 1e-11 * max(1.0, np.abs(want).max())
    assert np.all(np.isfinite(moms)) and not np.array_equal(moms[0], moms[-1])
    f.close()
    # (2) re-sharded: two handles, global filter indices preserved
    half = B // 2 + 7
    mu_a, p_a, sse_a, _, fa = run(0, half)
    mu_b, p_b, sse_b, _, fb = run(half, B)
    assert np.array_equal(np.concatenate([mu_a, mu_b]), mu) and np.array_equal(np.concatenate([p_a, p_b]), p)
    assert np.array_equal(np.concatenate([sse_a, sse_b]), sse)
    fa.close()
    fb.close()


@pytest.mark.gpu
def test_generated_association_and_common_ticks_equal_array_forms_bitwise(eb):
    """ekf_batch_associate_generated (noise drawn inside the kernel) == ekf_batch_detection_noise arrays fed to
    ekf_batch_associate; ekf_batch_step_ticks_common == ekf_batch_step_ticks with four constant tick arrays."""
    import torch
    B, L, m, D = 3001, 16, 6, 9
    rs = np.random.default_rng(4)
    res = []
    for mode in ("arrays", "generated"):
        f = eb.BatchedEKF(B, L)
        f.init(0.1)
        tag_slot, n_seen = f.new_slot_table(12)
        for step in range(5):
            st = np.random.default_rng(100 + step)
            tags = torch.from_numpy(st.integers(0, 12, D).astype(np.int32)).to(f.dev)
            tx = torch.from_numpy(st.normal(0, 0.3, D)).to(f.dev)
            tz = torch.from_numpy(st.uniform(0.4, 1.45, D)).to(f.dev)
            if mode == "arrays":
                ntx, ntz = f.detection_noise(99, step, D, 0.03, first_filter=12345)
                slot, rng, brg = f.associate(tags, tx, tz, tag_slot, n_seen, m, ntx, ntz)
                full = [torch.full((B,), v, dtype=torch.int32, device=f.dev) for v in (1010 + 7 * step, 1012 + 6 * step,
                                                                                      1003 + 7 * step, 1006 + 6 * step)]
                f.step_ticks(*full, 0.0318, 0.1, 135, slot, rng, brg)
            else:
                slot, rng, brg = f.associate(tags, tx, tz, tag_slot, n_seen, m, generated=(99, step, 12345, 0.03))
                f.step_ticks_common(1010 + 7 * step, 1012 + 6 * step, 1003 + 7 * step, 1006 + 6 * step, 0.0318, 0.1, 135,
                                    slot, rng, brg)
        assert f.fault_count() == 0
        res.append((f.mu.clone(), f.P.clone(), tag_slot.clone(), n_seen.clone(), f.associate_overflow_count()))
    for a, b in zip(res[0][:4], res[1][:4]):
        assert torch.equal(a, b)
    assert res[0][4] == res[1][4]
    del rs


@pytest.mark.gpu
@pytest.mark.parametrize("L,m,variant", [(20, 3, "auto"), (8, 2, "auto"), (16, 6, "auto"), (20, 3, "staged"),
                                         (8, 4, "group16"), (12, 1, "regw")])
def test_in_kernel_generation_equals_perturb_then_step_bitwise(eb, monkeypatch, L, m, variant):
    """ekf_batch_step_generated (inputs drawn by m + 1 lanes inside the step kernel, under the TMA load) ==
    ekf_batch_perturb followed by ekf_batch_step, BITWISE, on every kernel family."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    monkeypatch.setenv("EKF_BATCH_KERNEL", variant)
    B = 2003
    base = wl.make_batched(1, L, m, steps=-(-L // m) + 5)
    rs = np.random.default_rng(8)
    mv, mm = rs.uniform(0.05, 0.5, B), rs.uniform(0.1, 1.0, B)
    sigma = (0.01, 0.003, 0.02, 0.02)
    outs = []
    for mode in ("perturb", "generated"):
        f = eb.BatchedEKF(B, L)
        f.set_noise(mv, mm)
        f.init(0.1)
        for t in range(base.steps):
            rec = torch.from_numpy(f.mc_record(base.ang[t, 0], base.lin[t, 0], base.slot[t, 0], base.rng[t, 0],
                                               base.brg[t, 0])).to(f.dev)
            if mode == "perturb":
                f.step(*f.perturb(77, t, rec, m, sigma, first_filter=10 ** 10))
            else:
                f.step_generated(77, t, rec, m, sigma, first_filter=10 ** 10)
        assert f.fault_count() == 0 and int(f.n_known.min()) == L
        outs.append((f.mu.clone(), f.P.clone()))
        f.close()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
