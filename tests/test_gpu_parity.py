"""GPU parity tests: the CUDA path (through the C-ABI, via ekf_bindings) against the golden vectors dumped
from the unmodified reference and against the numpy oracle on the same seeded inputs.

Tolerance: 1e-9 (BASELINE.json north star) on max-norm mu, max-norm P and covariance-scaled P; landmark
association / slot bookkeeping must be identical."""
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


def _run_batched(eb, z, joseph=False, ticks=False):
    L = int(z["n_landmarks"])
    T, B, m = z["slot"].shape
    f = eb.BatchedEKF(B, L)
    f.set_noise(z["motion_var"], z["meas_var"])
    f.set_flags(joseph=joseph)
    f.init(0.1)
    out = {}
    keep = {int(t): i for i, t in enumerate(z["keep_t"])}
    for t in range(T):
        f.step(_dev(f.dev, z["ang"][t]), _dev(f.dev, z["lin"][t]), _dev(f.dev, z["slot"][t]),
               _dev(f.dev, z["rng"][t]), _dev(f.dev, z["brg"][t]))
        if t in keep:
            out[t] = f.state()
    assert f.fault_count() == 0
    assert np.all(f.n_known.cpu().numpy() == L)
    return out, keep


@pytest.mark.parametrize("name", ["batched_b48_l20.npz", "batched_b32_l8.npz"])
@pytest.mark.parametrize("joseph", [False, True])
def test_batched_vs_reference_golden(eb, name, joseph):
    z = golden(name)
    out, keep = _run_batched(eb, z, joseph)
    worst = 0.0
    for t, (mu, p) in out.items():
        for b in range(mu.shape[0]):
            worst = max(worst, *assert_parity(mu[b], p[b], z["mean"][keep[t], b], z["cov"][keep[t], b],
                                              what=f"{name} joseph={joseph} t={t} b={b}"))
    print(f"{name} joseph={joseph}: worst parity {worst:.3e}")


def test_batched_pack_unpack_roundtrip_and_pad(eb):
    rng = np.random.default_rng(0)
    f = eb.BatchedEKF(7, 20)
    mean = rng.normal(size=(7, 43))
    cov = rng.normal(size=(7, 43, 43))
    f.load(mean, cov)
    m2, c2 = f.state()
    np.testing.assert_array_equal(m2, mean)
    np.testing.assert_array_equal(c2, cov)
    P = f.P.cpu().numpy().reshape(7, 43, 44)
    assert np.all(P[:, :, 43] == 0) and np.all(f.mu.cpu().numpy()[:, 43] == 0)


def test_batched_ticks_odometry_equals_displacement(eb):
    """a1/a2 on the device (ticks) must give the same step as host delta_phi/displacement + step."""
    import torch
    z = golden("batched_b32_l8.npz")
    B, L = 32, 8
    rng = np.random.default_rng(3)
    prev_l, prev_r = rng.integers(900, 1100, B).astype(np.int32), rng.integers(900, 1100, B).astype(np.int32)
    cur_l, cur_r = prev_l + rng.integers(-3, 12, B).astype(np.int32), prev_r + rng.integers(-3, 12, B).astype(np.int32)
    ang = np.zeros(B)
    lin = np.zeros(B)
    for b in range(B):
        ang[b], lin[b] = eb.displacement(0.0318, 0.1, eb.delta_phi(int(cur_l[b]), int(prev_l[b]), 135),
                                         eb.delta_phi(int(cur_r[b]), int(prev_r[b]), 135))
        a_o, l_o = eo.displacement(0.0318, 0.1, eo.delta_phi(int(cur_l[b]), int(prev_l[b]), 135),
                                   eo.delta_phi(int(cur_r[b]), int(prev_r[b]), 135))
        assert (ang[b], lin[b]) == (a_o, l_o)            # integer ticks -> identical doubles
    res = []
    for mode in ("disp", "ticks"):
        f = eb.BatchedEKF(B, L)
        f.set_noise(z["motion_var"], z["meas_var"])
        f.init(0.1)
        sl, r, q = _dev(f.dev, z["slot"][0]), _dev(f.dev, z["rng"][0]), _dev(f.dev, z["brg"][0])
        if mode == "disp":
            f.step(_dev(f.dev, ang), _dev(f.dev, lin), sl, r, q)
        else:
            f.step_ticks(_dev(f.dev, cur_l), _dev(f.dev, cur_r), _dev(f.dev, prev_l), _dev(f.dev, prev_r),
                         0.0318, 0.1, 135, sl, r, q)
        torch.cuda.synchronize()
        res.append(f.state())
    np.testing.assert_array_equal(res[0][0], res[1][0])
    np.testing.assert_array_equal(res[0][1], res[1][1])


def test_batched_faults_and_skips(eb):
    """slot -1 = no observation; a slot that skips first-seen order is refused and counted."""
    f = eb.BatchedEKF(4, 8)
    f.init(0.1)
    z = np.zeros(4)
    slot = np.array([[0, -1], [-1, -1], [2, 0], [0, 1]], dtype=np.int32)
    r = np.full((4, 2), 1.0)
    b = np.array([[0.1, 0.0], [0.0, 0.0], [0.2, -0.2], [0.3, -0.3]])
    f.step(_dev(f.dev, z + 0.05), _dev(f.dev, z + 0.03), _dev(f.dev, slot), _dev(f.dev, r), _dev(f.dev, b))
    assert f.fault_count() == 1
    assert list(f.n_known.cpu().numpy()) == [1, 0, 1, 2]
    mu, p = f.state()
    for k, obs in enumerate([[(0, 1.0, 0.1)], [], [(0, 1.0, -0.2)], [(0, 1.0, 0.3), (1, 1.0, -0.3)]]):
        m0 = np.zeros(19)
        c0 = np.zeros((19, 19))
        c0[:3, :3] = 0.1 * np.eye(3)
        m_o, c_o, _ = eo.step_from_observations(m0, c0, 0.05, 0.03, obs, eo.Flags(), 0)
        assert_parity(mu[k], p[k], m_o, c_o, what=f"filter {k}")


@pytest.mark.parametrize("flags", [dict(circular=False), dict(disable_motion=True), dict(measurement=False)])
def test_batched_flag_variants_vs_oracle(eb, flags):
    z = golden("batched_b32_l8.npz")
    T, B, m = z["slot"].shape
    f = eb.BatchedEKF(B, 8)
    f.set_noise(z["motion_var"], z["meas_var"])
    f.set_flags(**flags)
    f.init(0.1)
    for t in range(6):
        f.step(_dev(f.dev, z["ang"][t]), _dev(f.dev, z["lin"][t]), _dev(f.dev, z["slot"][t]),
               _dev(f.dev, z["rng"][t]), _dev(f.dev, z["brg"][t]))
    mu, p = f.state()
    for b in range(0, B, 7):
        fl = eo.Flags(motion_var=float(z["motion_var"][b]), meas_var=float(z["meas_var"][b]),
                      circular=flags.get("circular", True), disable_motion=flags.get("disable_motion", False),
                      enable_measurement=flags.get("measurement", True))
        m_o = np.zeros(19)
        c_o = np.zeros((19, 19))
        c_o[:3, :3] = 0.1 * np.eye(3)
        nk = 0
        for t in range(6):
            obs = [(int(s), float(r), float(q)) for s, r, q in zip(z["slot"][t, b], z["rng"][t, b], z["brg"][t, b]) if s >= 0]
            m_o, c_o, nk = eo.step_from_observations(m_o, c_o, float(z["ang"][t, b]), float(z["lin"][t, b]), obs, fl, nk)
        assert_parity(mu[b], p[b], m_o, c_o, what=f"{flags} b={b}")


# ---------------------------------------------------------------------------------------------------------
# large map
# ---------------------------------------------------------------------------------------------------------
def _run_large(eb, name, joseph=False):
    from slam_duckietown_b200 import workloads as wl
    z = golden(name)
    L = int(z["n_landmarks"])
    w = wl.make_large_map(L, updates=2 * len(z["sched"]))
    g = eb.LargeMapEKF(L)
    g.set_noise(0.1, 0.7)
    g.set_flags(joseph=joseph)
    g.load(w.mu0, w.p0())
    for row in z["sched"]:
        g.predict(float(row[0]), float(row[1]))
        for i in range(int(row[2])):
            g.update(int(row[3 + 3 * i]), float(row[4 + 3 * i]), float(row[5 + 3 * i]))
    mu, p = g.state()
    return z, mu, p


@pytest.mark.parametrize("name", ["large_l256.npz", "large_l1024.npz", "large_l4096.npz"])
@pytest.mark.parametrize("joseph", [False, True])
def test_large_map_vs_reference_golden(eb, name, joseph):
    z, mu, p = _run_large(eb, name, joseph)
    rows = z["rows"]
    d = np.sqrt(z["p_diag"])
    e_mu = np.max(np.abs(mu - z["mean"])) / np.max(np.abs(z["mean"]))
    e_rows = np.max(np.abs(p[rows] - z["p_rows"]) / (d[rows][:, None] * d[None, :]))
    e_cols = np.max(np.abs(p[:, rows[:8]] - z["p_cols"]) / (d[:, None] * d[rows[:8]][None, :]))
    e_diag = np.max(np.abs(np.diag(p) - z["p_diag"]) / z["p_diag"])
    e_sum = abs(p.sum() - z["p_sum"][0]) / z["p_sum"][1]
    print(f"{name} joseph={joseph}: mu {e_mu:.2e} rows {e_rows:.2e} cols {e_cols:.2e} diag {e_diag:.2e} sum {e_sum:.2e}")
    assert max(e_mu, e_rows, e_cols, e_diag, e_sum) <= TOL


def test_large_map_50_steps_vs_oracle_and_pad(eb):
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_large_map(700, updates=50)          # n = 1403: column tail + row tail both non-trivial
    g = eb.LargeMapEKF(700)
    g.set_noise(0.2, 0.5)
    g.load(w.mu0, w.p0())
    m_o, c_o = w.mu0.copy(), w.p0()
    fl = eo.Flags(motion_var=0.2, meas_var=0.5)
    for op in w.ops:
        if op[0] == "update":
            g.update(op[1], op[2], op[3])
            eo.update_rank(m_o, c_o, op[1], op[2], op[3], 0.5)
        else:
            g.predict(op[1], op[2])
            dx, dy, th, g0, g1 = eo.motion_step(m_o[2], op[1], op[2], fl)
            m_o[0] += dx; m_o[1] += dy; m_o[2] = th
            eo.predict_cov_rank(c_o, g0, g1, 0.2)
    mu, p = g.state()
    assert_parity(mu, p, m_o, c_o, what="50 ops n=1403")
    full = g.P.cpu().numpy()
    assert np.all(full[:, g.n:] == 0.0), "pad columns must stay exactly zero"


@pytest.mark.parametrize("n_lm,group", [(700, 3), (700, 6), (256, 2), (1024, 4)])
def test_large_map_fused_multi_update_bitwise(eb, n_lm, group):
    """ekf_lm_update_multi (one pass over P for the step's updates) == the same updates applied one at a time,
    BITWISE, and both match the oracle; groups above 4 are split, repeated slots inside a group are allowed."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_large_map(n_lm, updates=24)
    seq, fus = eb.LargeMapEKF(n_lm), eb.LargeMapEKF(n_lm)
    m_o, c_o = w.mu0.copy(), w.p0()
    for g in (seq, fus):
        g.set_noise(0.2, 0.5)
        g.load(w.mu0, w.p0())
    fl = eo.Flags(motion_var=0.2, meas_var=0.5)
    pend = []

    def flush():
        if pend:
            fus.update_multi([o[1] for o in pend], [o[2] for o in pend], [o[3] for o in pend])
            pend.clear()
    for op in w.ops:
        if op[0] == "update":
            seq.update(op[1], op[2], op[3])
            eo.update_rank(m_o, c_o, op[1], op[2], op[3], 0.5)
            pend.append(op)
            if len(pend) == group:
                flush()
        elif group <= 4:                                # group > 4: no predicts, so that groups really get split
            flush()
            for g in (seq, fus):
                g.predict(op[1], op[2])
            dx, dy, th, g0, g1 = eo.motion_step(m_o[2], op[1], op[2], fl)
            m_o[0] += dx; m_o[1] += dy; m_o[2] = th
            eo.predict_cov_rank(c_o, g0, g1, 0.2)
    flush()
    assert torch.equal(seq.mu, fus.mu), "mu differs between sequential and fused updates"
    assert torch.equal(seq.P, fus.P), "P differs between sequential and fused updates"
    mu, p = fus.state()
    assert_parity(mu, p, m_o, c_o, what=f"fused multi-update n_lm={n_lm} group={group}")


def test_large_map_augment_matches_growing_arrays(eb):
    """capacity pre-allocation == the reference's zero-padded growth (:341-360)."""
    g = eb.LargeMapEKF(16)
    g.set_noise(0.1, 0.7)
    g.load(np.array([0.1, 0.2, 0.3]), np.eye(3) * 0.1)
    m_o = np.zeros(35)
    m_o[:3] = [0.1, 0.2, 0.3]
    c_o = np.zeros((35, 35))
    c_o[:3, :3] = np.eye(3) * 0.1
    nk = 0
    rng = np.random.default_rng(2)
    for t in range(8):
        obs = [(s, float(rng.uniform(0.3, 1.4)), float(rng.uniform(-1, 1))) for s in ([2 * t, 2 * t + 1] if t < 8 else [])]
        pose = m_o[:3].copy()
        for s, r, b in obs:
            g.augment(s, pose[0] + r * np.cos(b + pose[2]), pose[1] + r * np.sin(b + pose[2]))
        g.predict(0.1, 0.04)
        for s, r, b in obs:
            g.update(s, r, b)
        m_o, c_o, nk = eo.step_from_observations(m_o, c_o, 0.1, 0.04, obs, eo.Flags(), nk)
    mu, p = g.state()
    assert_parity(mu, p, m_o, c_o, what="augment")


# ---------------------------------------------------------------------------------------------------------
# reference call surface
# ---------------------------------------------------------------------------------------------------------
def test_scalar_surface_unit_cases(eb):
    z = golden("unit_cases.npz")
    names = ("ENABLE_CIRCULAR_INTERPOLATION", "DISABLE_MOTION_MODEL", "ENABLE_MEASUREMENT_MODEL", "IGNORE_TAGS",
             "MOTION_MODEL_VARIANCE", "MEASUREMENT_MODEL_VARIANCE")
    defaults = {k: getattr(eb, k) for k in names}
    for i, name in enumerate(z["names"]):
        p = f"c{i}_"
        flags = {str(k): eval(str(v)) for k, v in zip(z[p + "flags_keys"], z[p + "flags_vals"])}
        dets = [[], []]
        for img, tag, pt, err in zip(z[p + "det_img"], z[p + "det_tag"], z[p + "det_pose_t"], z[p + "det_err"]):
            dets[int(img)].append(eo.Detection(int(tag), pt.reshape(3, 1), float(err)))
        ti = {int(a): int(b) for a, b in zip(z[p + "ti_in_ids"], z[p + "ti_in_slots"])}
        mean_in, cov_in = z[p + "mean_in"].copy(), z[p + "cov_in"].copy()
        try:
            for k, v in flags.items():
                setattr(eb, k, v)
            m, c, tags = eb.EKF_pose_estimation(float(z[p + "ang"]), float(z[p + "lin"]), mean_in, cov_in, 0.7,
                                                [("0", dets[0]), ("1", dets[1])], ti)
        finally:
            for k, v in defaults.items():
                setattr(eb, k, v)
        np.testing.assert_array_equal(mean_in, z[p + "mean_in"])       # inputs untouched, new arrays returned
        assert list(ti.keys()) == list(z[p + "ti_out_ids"]) and list(ti.values()) == list(z[p + "ti_out_slots"])
        assert list(tags.keys()) == list(z[p + "order"]), name
        got = np.array([[float(x) for x in v] for v in tags.values()]).reshape(-1, 6)
        np.testing.assert_array_equal(got[:, 3], z[p + "tags"][:, 3])           # tag ids
        np.testing.assert_allclose(got, z[p + "tags"], rtol=1e-13, atol=1e-13)
        assert m.shape == z[p + "mean_out"].shape
        assert_parity(m, c, z[p + "mean_out"], z[p + "cov_out"], what=str(name))


def test_scalar_surface_closed_loop(eb):
    z = golden("closed_loop_n20.npz")
    mean, cov, ti = np.zeros(3), np.eye(3) * 0.1, {}
    keep = {int(t): i for i, t in enumerate(z["keep_t"])}
    ds = z["det_step"]
    worst = 0.0
    for t in range(len(z["ang"])):
        sel = np.flatnonzero(ds == t)
        pairs = [(str(t), [eo.Detection(int(z["det_tag"][k]), z["det_pose_t"][k].reshape(3, 1), float(z["det_err"][k]))
                           for k in sel if z["det_img"][k] == im]) for im in range(3)]
        mean, cov, _ = eb.EKF_pose_estimation(float(z["ang"][t]), float(z["lin"][t]), mean, cov, 0.7, pairs, ti)
        if t in keep:
            i, n = keep[t], int(z["keep_n"][keep[t]])
            assert len(mean) == n
            worst = max(worst, *assert_parity(mean, cov, z["keep_mean"][i][:n], z["keep_cov"][i][:n, :n], what=f"step {t}"))
    assert list(ti.keys()) == list(z["tag_ids"]) and list(ti.values()) == list(z["tag_slots"])
    print(f"closed loop 300 steps: worst parity {worst:.3e}")


def test_surrogate_bag_replay(eb, surrogate):
    """Configuration 1: the decoded-bag event loop drives the CUDA core; every one of the 70 EKF steps must match
    what the reference's own replay() produced on the same events."""
    from slam_duckietown_b200 import replay
    lines, dets, ref = surrogate
    recs = replay.replay_events(lines, dets)
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
    print(f"surrogate replay: worst parity {worst:.3e}")


def test_ekf3_surface(eb):
    import torch
    z = golden("ekf3.npz")
    K = len(z["v"])
    dev = torch.device("cuda")
    st = _dev(dev, z["state"].T.copy())
    cv = _dev(dev, z["cov"].reshape(K, 9).T.copy())
    eb.predict3_batched(st, cv, _dev(dev, z["v"]), _dev(dev, z["omega"]), _dev(dev, z["dt"]))
    np.testing.assert_allclose(st.cpu().numpy().T, z["pred_state"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(cv.cpu().numpy().T.reshape(K, 3, 3), z["pred_cov"], rtol=1e-11, atol=1e-13)
    eb.update3_batched(st, cv, _dev(dev, z["obs"][:, 0].copy()), _dev(dev, z["obs"][:, 1].copy()),
                       _dev(dev, z["lm"][:, 0].copy()), _dev(dev, z["lm"][:, 1].copy()))
    np.testing.assert_allclose(st.cpu().numpy().T, z["upd_state"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(cv.cpu().numpy().T.reshape(K, 3, 3), z["upd_cov"], rtol=1e-9, atol=1e-11)
    s, c = eb.predict(z["state"][3], z["cov"][3], (z["v"][3], z["omega"][3]), z["dt"][3])
    np.testing.assert_allclose(s, z["pred_state"][3], atol=1e-12)
    s, c = eb.update(s, c, z["obs"][3], z["lm"][3])
    np.testing.assert_allclose(c, z["upd_cov"][3], rtol=1e-9, atol=1e-11)


@pytest.mark.parametrize("name", ["batched_b48_l20.npz", "batched_b32_l8.npz"])
def test_batched_joseph_variants_bitwise_equal(eb, name, monkeypatch):
    """One-pass Joseph form: the dispatched kernels (W-in-registers at 20 landmarks, sub-warp groups at 8) apply the
    staged kernel's per-element operation sequence (two rank-2 terms per update: K W, then AHt K^T): BITWISE equal."""
    z = golden(name)
    outs = {}
    variants = ("auto", "staged", "regw", "group8")
    for variant in variants:
        monkeypatch.setenv("EKF_BATCH_KERNEL", variant)
        outs[variant], _ = _run_batched(eb, z, joseph=True)
    for variant in variants[1:]:
        for t, (mu, p) in outs["auto"].items():
            mu2, p2 = outs[variant][t]
            assert np.array_equal(mu, mu2), f"{name} joseph {variant} mu differs at t={t}"
            assert np.array_equal(p, p2), f"{name} joseph {variant} P differs at t={t}"


@pytest.mark.parametrize("name", ["batched_b48_l20.npz", "batched_b32_l8.npz"])
def test_batched_kernel_variants_bitwise_equal(eb, name, monkeypatch):
    """The W-in-registers kernel (default at 20 landmarks), the sub-warp group kernels (default at <= 8 landmarks),
    and the fully staged kernel apply the same per-element operation sequence: their states
    must be BITWISE equal at every kept step."""
    import torch
    z = golden(name)
    outs = {}
    variants = ("auto", "staged", "regw", "group8", "group16")
    for variant in variants:
        monkeypatch.setenv("EKF_BATCH_KERNEL", variant)      # read by ekf_batch_create
        outs[variant], _ = _run_batched(eb, z)
    for variant in variants[1:]:
        for t, (mu, p) in outs["auto"].items():
            mu2, p2 = outs[variant][t]
            assert np.array_equal(mu, mu2), f"{name} {variant} mu differs at t={t}"
            assert np.array_equal(p, p2), f"{name} {variant} P differs at t={t}"



# ---------------------------------------------------------------------------------------------------------
# full-size properties (BASELINE.json sizes; no oracle at this scale)
# ---------------------------------------------------------------------------------------------------------
def test_batched_full_size_properties(eb):
    """65,536 x 20 landmarks: filters fed identical inputs must end bitwise identical wherever they sit in the
    batch; the first 48 filters must still match the reference golden; P stays symmetric to rounding and
    its diagonal positive; an update never increases a variance."""
    import torch
    z = golden("batched_b48_l20.npz")
    T, B0, m = z["slot"].shape
    B = 65536
    reps = B // B0 + 1

    def tile(a):
        return np.concatenate([a] * reps, axis=0)[:B].copy()
    f = eb.BatchedEKF(B, 20)
    f.set_noise(tile(z["motion_var"]), tile(z["meas_var"]))
    f.init(0.1)
    for t in range(T):
        f.step(_dev(f.dev, tile(z["ang"][t])), _dev(f.dev, tile(z["lin"][t])), _dev(f.dev, tile(z["slot"][t])),
               _dev(f.dev, tile(z["rng"][t])), _dev(f.dev, tile(z["brg"][t])))
    assert f.fault_count() == 0
    P = f.P.view(B, 43, 44)
    mu = f.mu
    last = (B // B0 - 1) * B0
    assert torch.equal(P[:B0], P[last:last + B0]) and torch.equal(mu[:B0], mu[last:last + B0])
    assert torch.equal(P[:B0], P[B0 * 700:B0 * 701])
    mu0, p0 = f.state(0, B0)
    i = len(z["keep_t"]) - 1
    for b in range(B0):
        assert_parity(mu0[b], p0[b], z["mean"][i, b], z["cov"][i, b], what=f"full-size b={b}")
    Pm = P[:, :, :43]
    d = torch.diagonal(Pm, dim1=1, dim2=2)
    assert bool((d > 0).all())
    asym = (Pm - Pm.transpose(1, 2)).abs().amax(dim=(1, 2)) / d.amax(dim=1)
    assert float(asym.max()) < 1e-9
    assert bool((P[:, :, 43] == 0).all())


def test_large_map_full_size_properties(eb):
    """4096 landmarks (n = 8195): update properties that hold at any size — the updated landmark's variance
    drops, trace(P) does not increase, P stays symmetric to rounding, a second filter fed the same ops is
    bitwise identical, and plain vs Joseph agree to 1e-9 covariance-scaled."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_large_map(4096, updates=6)
    outs = []
    for joseph in (False, True, False):
        g = eb.LargeMapEKF(4096)
        g.set_noise(0.1, 0.7)
        g.set_flags(joseph=joseph)
        g.load(w.mu0, w.p0())
        tr0 = float(torch.diagonal(g.P[:, :g.n]).sum())
        for op in w.ops:
            if op[0] == "update":
                j = 3 + 2 * op[1]
                before = float(g.P[j, j])
                g.update(op[1], op[2], op[3])
                assert float(g.P[j, j]) < before
            else:
                g.predict(op[1], op[2])
        Pm = g.P[:, :g.n]
        assert float(torch.diagonal(Pm).sum()) < tr0 + 3 * 0.1 ** 2 * 2
        d = torch.sqrt(torch.diagonal(Pm))
        asym = ((Pm - Pm.T).abs() / (d[:, None] * d[None, :])).max()
        assert float(asym) < 1e-9
        assert bool((g.P[:, g.n:] == 0).all())
        outs.append((g.mu.clone(), Pm.clone()))
        g.close()
    assert torch.equal(outs[0][1], outs[2][1]) and torch.equal(outs[0][0], outs[2][0])
    d = torch.sqrt(torch.diagonal(outs[0][1]))
    rel = ((outs[0][1] - outs[1][1]).abs() / (d[:, None] * d[None, :])).max()
    assert float(rel) < 1e-9


# ---------------------------------------------------------------------------------------------------------
# device-side association (a3/a4 per filter) and ensemble trajectory error
# ---------------------------------------------------------------------------------------------------------
def test_device_association_matches_host_and_feeds_step(eb):
    """ekf_batch_associate on per-filter perturbed detections == the host association (ekf_bindings.
    associate_window, itself pinned to replay_no_ros.py:280-337) run filter by filter: gate decisions, slot tables
    and update order bit-exact, range/bearing to 1e-12; the (slot, rng, brg) block then drives the fused step and
    the state matches the oracle fed by the host association."""
    import torch
    from slam_duckietown_b200.replay import make_detection
    rs = np.random.default_rng(11)
    B, L, m, T, n_tags = 40, 12, 8, 12, 10      # m = 8 >= detections per window: no overflow
    tag_ids = [80, 44, 41, 26, 110, 106, 101, 31, 65, 23]
    ignore = (41,)
    world = rs.uniform([-0.2, -1.2], [1.6, 1.2], size=(n_tags, 2))     # camera-frame (x fwd, y left), near the gate
    f = eb.BatchedEKF(B, L)
    mv, mm = rs.uniform(0.05, 0.3, B), rs.uniform(0.1, 0.8, B)
    f.set_noise(mv, mm)
    f.init(0.1)
    tag_slot, n_seen = f.new_slot_table(n_tags)
    host_index = [dict() for _ in range(B)]
    n = 3 + 2 * L
    c0 = np.zeros((n, n)); c0[:3, :3] = np.eye(3) * 0.1
    st = [(np.zeros(n), c0.copy(), 0) for _ in range(B)]
    sse = torch.zeros(B, dtype=torch.float64, device=f.dev)
    sse_ref = np.zeros(B)
    for t in range(T):
        D = int(rs.integers(0, 9))
        which = rs.integers(0, n_tags, D)
        tx = -(world[which, 1] + rs.normal(0, 0.05, D))               # pose_t[0] = -y_rel
        tz = world[which, 0] + rs.normal(0, 0.05, D)                  # pose_t[2] = x_rel
        if D:
            tz[0] = np.sqrt(max(2.25 - tx[0] * tx[0], 0.0))            # a detection sitting right on the 1.5 m gate
        ntx, ntz = rs.normal(0, 0.03, (D, B)), rs.normal(0, 0.03, (D, B))
        if D:
            ntx[0, ::2] = 0.0
            ntz[0, ::2] = 0.0
        compact = np.array([-1 if tag_ids[w] in ignore else w for w in which], dtype=np.int32)
        slot, rng, brg = f.associate(_dev(f.dev, compact), _dev(f.dev, tx), _dev(f.dev, tz), tag_slot, n_seen, m,
                                     _dev(f.dev, ntx), _dev(f.dev, ntz))
        ang, lin = rs.normal(0.08, 0.05, B), np.abs(rs.normal(0.05, 0.02, B))
        slot_h, rng_h, brg_h = slot.cpu().numpy(), rng.cpu().numpy(), brg.cpu().numpy()
        for b in range(B):
            dets = [make_detection(tag_ids[which[d]], [tx[d] + ntx[d, b], 0.0, tz[d] + ntz[d, b]]) for d in range(D)]
            mean, cov, nk = st[b]
            tags = eb.associate_window(mean, [("0", dets)], host_index[b], ignore)
            order = list(tags.keys())[:m]
            assert list(slot_h[b, :len(order)]) == order and np.all(slot_h[b, len(order):] == -1), (t, b)
            for k, s in enumerate(order):
                assert abs(rng_h[b, k] - tags[s][4]) <= 1e-12 * max(1.0, abs(tags[s][4]))
                assert abs(brg_h[b, k] - tags[s][5]) <= 1e-12
            obs = [(s, tags[s][4], tags[s][5]) for s in order]
            fl = eo.Flags(motion_var=float(mv[b]), meas_var=float(mm[b]))
            mean2, cov2, nk = eo.step_from_observations(mean, cov, float(ang[b]), float(lin[b]), obs, fl, nk)
            st[b] = (mean2, cov2, nk)
            sse_ref[b] += (mean2[0] - 0.1 * t) ** 2 + (mean2[1] + 0.05 * t) ** 2
        f.step(_dev(f.dev, ang), _dev(f.dev, lin), slot, rng, brg)
        f.accumulate_ate(0.1 * t, -0.05 * t, sse)
    assert f.fault_count() == 0 and f.associate_overflow_count() == 0
    ts = tag_slot.cpu().numpy()
    for b in range(B):
        want = np.full(n_tags, -1)
        for tid, s in host_index[b].items():
            want[tag_ids.index(tid)] = s
        assert np.array_equal(ts[:, b], want), f"slot table of filter {b}"
        assert int(n_seen[b]) == len(host_index[b])
    mu, p = f.state()
    for b in range(B):
        assert_parity(mu[b], p[b], st[b][0], st[b][1], what=f"device association pipeline b={b}")
    assert np.allclose(sse.cpu().numpy(), sse_ref, rtol=1e-9, atol=1e-12)


def test_monte_carlo_replay_of_decoded_bag(eb, surrogate):
    """events.csv -> windows -> device association -> tick-odometry fused step, 64 ensemble members. With zero
    perturbation and the reference's noise flags EVERY member must reproduce the reference's own replay() of the
    same events at each of the 70 steps (state, slot order); with perturbation the members differ, stay finite and
    fault-free, and the ATE accumulator equals a host recomputation."""
    import torch
    from slam_duckietown_b200 import events_io as io, monte_carlo as mc
    lines, dets, ref = surrogate
    wins = io.windows_from_events([io.parse_event_line(l) for l in lines], dets)
    B, L, m = 64, 16, 6
    step = {"i": 0, "worst": 0.0}
    track = []

    def check(f, w):
        i = step["i"]
        n = int(ref["n"][i])
        mu, p = f.state(0, B)
        for b in (0, B - 1):
            e = assert_parity(mu[b][:n], p[b][:n, :n], ref["mean"][i][:n], ref["cov"][i][:n, :n], what=f"MC step {i} b={b}")
            step["worst"] = max(step["worst"], *e)
            assert np.all(mu[b][n:] == 0.0)
        track.append((mu[:, 0].copy(), mu[:, 1].copy(), w.ground_truth))
        step["i"] += 1
    out = mc.monte_carlo_replay(wins, B, L, m, eb._flag("MOTION_MODEL_VARIANCE"), eb._flag("MEASUREMENT_MODEL_VARIANCE"),
                                det_sigma=0.0, on_step=check)
    assert step["i"] == 70 and out["faults"] == 0 and out["overflows"] == 0
    for b in (0, B - 1):                                                           # per-member TAG_INDEX == reference's
        got = {tid: int(out["tag_slot"][c, b]) for tid, c in out["tag_table"].items() if out["tag_slot"][c, b] >= 0}
        assert got == {int(t): int(s) for t, s in zip(ref["tag_ids"], ref["tag_slots"])}
    assert np.all(out["n_landmarks_seen"] == len(ref["tag_ids"]))
    sse = np.zeros(B)
    cnt = 0
    for x, y, gt in track:
        if gt is not None:
            sse += (x - gt[0]) ** 2 + (y - gt[1]) ** 2
            cnt += 1
    assert cnt == out["scored_steps"] > 0
    assert np.allclose(out["rmse"], np.sqrt(sse / cnt), rtol=1e-12, atol=0)
    assert np.all(out["rmse"] == out["rmse"][0])                                   # identical members
    print(f"monte carlo replay: worst parity {step['worst']:.3e}, ATE rmse {out['rmse'][0]:.4f} m")
    rs = np.random.default_rng(5)
    out2 = mc.monte_carlo_replay(wins, B, L, m, np.exp(rs.uniform(np.log(0.05), np.log(1.0), B)),
                                 np.exp(rs.uniform(np.log(0.05), np.log(1.5), B)), det_sigma=0.02, seed=7)
    assert out2["faults"] == 0 and np.all(np.isfinite(out2["rmse"])) and np.all(np.isfinite(out2["pose"]))
    assert len(np.unique(out2["rmse"])) > B // 2


def test_group_kernel_ragged_batch_and_mixed_observations(eb, monkeypatch):
    """Batch sizes that do not fill the last warp of the sub-warp group kernels (4 or 2 filters per warp), filters
    of one warp with DIFFERENT observation patterns (some slots -1, so groups diverge), and a predict-only step:
    group8 / group16 / warp-per-filter must agree bitwise."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    B, L, m = 30, 8, 3
    w = wl.make_batched(B, n_landmarks=L, m=m, steps=10)
    slot = w.slot.copy()
    rs = np.random.default_rng(3)
    steady = slice(w.warm_steps, None)
    drop = rs.random(slot[steady].shape) < 0.3          # after every landmark is known: drop observations at random
    slot[steady][drop] = -1
    res = {}
    for variant in ("regw", "group8", "group16"):
        monkeypatch.setenv("EKF_BATCH_KERNEL", variant)
        f = eb.BatchedEKF(B, L)
        f.set_noise(w.motion_var, w.meas_var)
        f.init(0.1)
        for t in range(w.steps):
            f.step(_dev(f.dev, w.ang[t]), _dev(f.dev, w.lin[t]), _dev(f.dev, slot[t]), _dev(f.dev, w.rng[t]),
                   _dev(f.dev, w.brg[t]))
        f.step(_dev(f.dev, w.ang[0]), _dev(f.dev, w.lin[0]))                     # m = 0: predict only
        assert f.fault_count() == 0
        res[variant] = (f.mu.clone(), f.P.clone(), f.n_known.clone())
    for variant in ("group8", "group16"):
        for a, b in zip(res["regw"], res[variant]):
            assert torch.equal(a, b), variant
    assert bool(torch.isfinite(res["regw"][0]).all())


@pytest.mark.parametrize("world,form", [(2, "plain"), (3, "plain"), (2, "joseph"), (3, "joseph"), (3, "relief")])
def test_p2p_row_exchange_two_processes(eb, world, form):
    """Row-block shards in TWO processes (one per rank, both on cuda:0 here; one per GPU on a multi-GPU box) exchange
    their rows over CUDA IPC peer memory — owners store tagged cells, consumers poll them, no collective: every shard must equal
    the same rows of an unsharded filter BITWISE and no spin may have timed out. form "joseph": the one-pass Joseph
    form, whose W needs K at every column — all-gathered through the same exchange inside the gather kernel; "relief": the
    uneven partition of large maps (rank 0 owns fewer landmarks), plain form."""
    import json
    import socket
    import subprocess
    import sys
    import os
    import torch
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_p2p_worker.py")
    ndev = torch.cuda.device_count()
    for attempt in range(4):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        procs = [subprocess.Popen([sys.executable, worker, str(r), str(world), str(port), str(r % ndev), form],
                                  stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
        res = []
        for p in procs:
            try:
                o, e = p.communicate(timeout=240)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            res.append((p.returncode, o, e))
            if p.returncode != 0:                      # do not wait for the peers of a rank that is gone
                for q in procs:
                    if q.poll() is None:
                        q.kill()
        # the probed port can be taken by someone else before rank 0 listens on it: rendezvous plumbing, try another
        if attempt < 3 and any(rc != 0 and "EADDRINUSE" in e for rc, _, e in res):
            continue
        break
    outs = []
    for rc, o, e in res:
        assert rc == 0, e[-2000:]
        outs.append(json.loads(o.strip().splitlines()[-1]))
    for o in outs:
        assert o["same"] and o["p2p_error"] == 0 and o["finite"], o

@pytest.mark.gpu
@pytest.mark.parametrize("batch,n_lm", [(5000, 20), (37, 8), (65536, 20)])
def test_ensemble_moments_match_numpy(eb, batch, n_lm):
    """ekf_batch_moments (fixed-order device reduction) == numpy sums of mu and mu mu^T over the filters; the
    ensemble mean / covariance derived from it == numpy's; two calls give bit-identical sums (no atomics)."""
    import torch
    from slam_duckietown_b200 import workloads as wl
    from slam_duckietown_b200.ensemble import ensemble_mean_cov
    w = wl.make_batched(batch, n_landmarks=n_lm, m=2, steps=4)
    f = eb.BatchedEKF(batch, n_lm)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)
    dev = f.dev
    for t in range(w.steps):
        f.step(torch.from_numpy(w.ang[t]).to(dev), torch.from_numpy(w.lin[t]).to(dev),
               torch.from_numpy(w.slot[t]).to(dev), torch.from_numpy(w.rng[t]).to(dev),
               torch.from_numpy(w.brg[t]).to(dev))
    mu, _ = f.state()
    n = f.n
    s1 = f.moments().cpu().numpy()
    s2 = f.moments().cpu().numpy()
    np.testing.assert_array_equal(s1, s2)
    want = np.concatenate([mu.sum(0), (mu.T @ mu).ravel()])
    scale = np.abs(want).max()
    assert np.abs(s1 - want).max() <= 1e-12 * scale * max(1.0, np.log2(batch))
    mean, cov = ensemble_mean_cov(f)                      # one rank: no process group, identity all-reduce
    np.testing.assert_allclose(mean.cpu().numpy(), mu.mean(0), rtol=0, atol=1e-12 * max(1.0, np.abs(mu).max()))
    np.testing.assert_allclose(cov.cpu().numpy(), np.cov(mu.T, bias=True), rtol=0, atol=1e-10 * max(1.0, np.abs(mu).max()) ** 2)


@pytest.mark.gpu
@pytest.mark.parametrize("L,m,variants", [(8, 3, ("regw", "group8", "group16", "staged")),
                                           (8, 4, ("regw", "group8", "staged")),
                                           (20, 4, ("regw", "staged")),
                                           (16, 6, ("auto", "regw", "staged")),      # the decoded-bag shape: K history re-read
                                           (20, 5, ("regw", "staged"))])             # from shared memory for m > 4
def test_batched_random_observation_patterns(eb, monkeypatch, L, m, variants):
    """Steady-state steps with RANDOM observation blocks per filter: any subset of the m entries empty (-1), the
    entries in random order, and the SAME landmark observed more than once in one step (the C-ABI takes any slot
    block; sequential semantics must hold). Every kernel variant must agree bitwise and sampled filters must match
    the oracle's sequential updates."""
    import torch
    from oracle import ekf_oracle as eo
    from slam_duckietown_b200 import workloads as wl
    B = 203                                                # not a multiple of the filters per warp
    w = wl.make_batched(B, n_landmarks=L, m=m, steps=-(-L // m) + 6)
    rs = np.random.default_rng(11)
    slot, rng_, brg = w.slot.copy(), w.rng.copy(), w.brg.copy()
    for t in range(w.warm_steps, w.steps):
        for b in range(B):
            k = rs.integers(0, 4)
            if k == 0:                                     # duplicate: entry 1 re-observes entry 0's landmark
                slot[t, b, 1] = slot[t, b, 0]
                rng_[t, b, 1] = rng_[t, b, 0] * (1.0 + 0.01 * rs.normal())
                brg[t, b, 1] = brg[t, b, 0] + 0.01 * rs.normal()
            elif k == 1:                                   # random holes
                slot[t, b, rs.random(m) < 0.5] = -1
            elif k == 2:                                   # random order
                perm = rs.permutation(m)
                slot[t, b], rng_[t, b], brg[t, b] = slot[t, b, perm], rng_[t, b, perm], brg[t, b, perm]
    res = {}
    for variant in variants:
        monkeypatch.setenv("EKF_BATCH_KERNEL", variant)
        f = eb.BatchedEKF(B, L)
        f.set_noise(w.motion_var, w.meas_var)
        f.init(0.1)
        for t in range(w.steps):
            f.step(_dev(f.dev, w.ang[t]), _dev(f.dev, w.lin[t]), _dev(f.dev, slot[t]), _dev(f.dev, rng_[t]),
                   _dev(f.dev, brg[t]))
        assert f.fault_count() == 0, variant
        res[variant] = (f.mu.clone(), f.P.clone(), f.n_known.clone())
        if variant == variants[0]:
            mu, p = f.state()
    for variant in variants[1:]:
        for a, b_ in zip(res[variants[0]], res[variant]):
            assert torch.equal(a, b_), variant
    n = 3 + 2 * L
    for b in (0, 1, 2, 3, 101, B - 1):
        m_o, c_o, nk = np.zeros(n), np.zeros((n, n)), 0
        c_o[:3, :3] = np.eye(3) * 0.1
        fl = eo.Flags(motion_var=float(w.motion_var[b]), meas_var=float(w.meas_var[b]))
        for t in range(w.steps):
            obs = [(int(s_), float(r), float(q)) for s_, r, q in zip(slot[t, b], rng_[t, b], brg[t, b]) if s_ >= 0]
            m_o, c_o, nk = eo.step_from_observations(m_o, c_o, float(w.ang[t, b]), float(w.lin[t, b]), obs, fl, nk)
        assert_parity(mu[b], p[b], m_o, c_o, what=f"random observation patterns L={L} m={m} b={b}")
