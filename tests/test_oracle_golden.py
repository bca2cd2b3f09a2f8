"""CPU: the oracle restatement against the golden vectors dumped from the unmodified reference."""
import numpy as np
import pytest

from conftest import assert_parity, golden
from oracle import ekf_oracle as eo


def _unit_cases():
    z = golden("unit_cases.npz")
    for i, name in enumerate(z["names"]):
        p = f"c{i}_"
        flags = {str(k): eval(str(v)) for k, v in zip(z[p + "flags_keys"], z[p + "flags_vals"])}
        dets = [[], []]
        for img, tag, pt, err in zip(z[p + "det_img"], z[p + "det_tag"], z[p + "det_pose_t"], z[p + "det_err"]):
            dets[int(img)].append(eo.Detection(int(tag), pt.reshape(3, 1), float(err)))
        ti = {int(a): int(b) for a, b in zip(z[p + "ti_in_ids"], z[p + "ti_in_slots"])}
        yield str(name), z, p, flags, [("0", dets[0]), ("1", dets[1])], ti


def _flags(d):
    return eo.Flags(motion_var=d.get("MOTION_MODEL_VARIANCE", 0.1), meas_var=d.get("MEASUREMENT_MODEL_VARIANCE", 0.7),
                    enable_measurement=d.get("ENABLE_MEASUREMENT_MODEL", True),
                    circular=d.get("ENABLE_CIRCULAR_INTERPOLATION", True),
                    disable_motion=d.get("DISABLE_MOTION_MODEL", False), ignore_tags=tuple(d.get("IGNORE_TAGS", ())))


def test_unit_cases_both_forms():
    n = 0
    for name, z, p, flags, pairs, ti in _unit_cases():
        for form in ("rank", "dense"):
            ti2 = dict(ti)
            m, c, tags = eo.ekf_pose_estimation(float(z[p + "ang"]), float(z[p + "lin"]), z[p + "mean_in"],
                                                z[p + "cov_in"], 0.7, pairs, ti2, _flags(flags), form=form)
            assert list(ti2.keys()) == list(z[p + "ti_out_ids"]) and list(ti2.values()) == list(z[p + "ti_out_slots"])
            assert list(tags.keys()) == list(z[p + "order"]), name
            got = np.array([[float(x) for x in v] for v in tags.values()]).reshape(-1, 6)
            np.testing.assert_array_equal(got[:, 2:], z[p + "tags"][:, 2:])      # err, id, range, bearing: exact
            np.testing.assert_allclose(got[:, :2], z[p + "tags"][:, :2], rtol=0, atol=1e-13)
            assert_parity(m, c, z[p + "mean_out"], z[p + "cov_out"], what=f"{name}/{form}")
        n += 1
    assert n == 11


def test_joseph_matches_plain_to_rounding():
    for name, z, p, flags, pairs, ti in _unit_cases():
        f = _flags(flags)
        f.joseph = True
        m, c, _ = eo.ekf_pose_estimation(float(z[p + "ang"]), float(z[p + "lin"]), z[p + "mean_in"],
                                         z[p + "cov_in"], 0.7, pairs, dict(ti), f)
        assert_parity(m, c, z[p + "mean_out"], z[p + "cov_out"], what=f"{name}/joseph")


def test_closed_loop_300_steps():
    z = golden("closed_loop_n20.npz")
    mean, cov, ti = np.zeros(3), np.eye(3) * 0.1, {}
    keep = {int(t): i for i, t in enumerate(z["keep_t"])}
    ds = z["det_step"]
    worst = 0.0
    for t in range(len(z["ang"])):
        sel = np.flatnonzero(ds == t)
        pairs = [(str(t), [eo.Detection(int(z["det_tag"][k]), z["det_pose_t"][k].reshape(3, 1), float(z["det_err"][k]))
                           for k in sel if z["det_img"][k] == im]) for im in range(3)]
        mean, cov, _ = eo.ekf_pose_estimation(float(z["ang"][t]), float(z["lin"][t]), mean, cov, 0.7, pairs, ti)
        if t in keep:
            i, n = keep[t], int(z["keep_n"][keep[t]])
            assert len(mean) == n
            worst = max(worst, max(assert_parity(mean, cov, z["keep_mean"][i][:n], z["keep_cov"][i][:n, :n],
                                                 what=f"step {t}")))
    assert list(ti.keys()) == list(z["tag_ids"]) and list(ti.values()) == list(z["tag_slots"])
    assert worst < 1e-10


def test_batched_step_form_vs_reference():
    for name in ("batched_b48_l20.npz", "batched_b32_l8.npz"):
        z = golden(name)
        L = int(z["n_landmarks"])
        n = 3 + 2 * L
        T, B, _ = z["slot"].shape
        keep = {int(t): i for i, t in enumerate(z["keep_t"])}
        for b in range(0, B, 5):
            mean = np.zeros(n)
            cov = np.zeros((n, n))
            cov[:3, :3] = np.eye(3) * 0.1
            nk = 0
            fl = eo.Flags(motion_var=float(z["motion_var"][b]), meas_var=float(z["meas_var"][b]))
            for t in range(T):
                obs = [(int(s), float(r), float(q)) for s, r, q in zip(z["slot"][t, b], z["rng"][t, b], z["brg"][t, b])
                       if s >= 0]
                mean, cov, nk = eo.step_from_observations(mean, cov, float(z["ang"][t, b]), float(z["lin"][t, b]),
                                                          obs, fl, nk)
                if t in keep:
                    assert_parity(mean, cov, z["mean"][keep[t], b], z["cov"][keep[t], b], what=f"{name} b={b} t={t}")


def test_large_map_rank_form_vs_reference():
    from slam_duckietown_b200 import workloads as wl
    for name in ("large_l256.npz", "large_l1024.npz"):
        z = golden(name)
        L = int(z["n_landmarks"])
        w = wl.make_large_map(L, updates=2 * len(z["sched"]))
        mean, cov = w.mu0.copy(), w.p0()
        fl = eo.Flags()
        for row in z["sched"]:
            ang, lin, k = float(row[0]), float(row[1]), int(row[2])
            obs = [(int(row[3 + 3 * i]), float(row[4 + 3 * i]), float(row[5 + 3 * i])) for i in range(k)]
            mean, cov, _ = eo.step_from_observations(mean, cov, ang, lin, obs, fl)
        rows = z["rows"]
        d = np.sqrt(z["p_diag"])
        assert np.max(np.abs(mean - z["mean"])) / np.max(np.abs(z["mean"])) <= 1e-9
        assert np.max(np.abs(cov[rows] - z["p_rows"]) / (d[rows][:, None] * d[None, :])) <= 1e-9
        assert np.max(np.abs(cov[:, rows[:8]] - z["p_cols"]) / (d[:, None] * d[rows[:8]][None, :])) <= 1e-9
        assert abs(cov.sum() - z["p_sum"][0]) <= 1e-9 * z["p_sum"][1]


def test_ekf3_vectors():
    z = golden("ekf3.npz")
    for i in range(len(z["v"])):
        s, c = eo.predict3(z["state"][i], z["cov"][i], (z["v"][i], z["omega"][i]), z["dt"][i])
        np.testing.assert_allclose(s, z["pred_state"][i], rtol=0, atol=1e-14)
        np.testing.assert_allclose(c, z["pred_cov"][i], rtol=1e-13, atol=1e-15)
        s, c = eo.update3(s, c, z["obs"][i], z["lm"][i])
        np.testing.assert_allclose(s, z["upd_state"][i], rtol=0, atol=1e-12)
        np.testing.assert_allclose(c, z["upd_cov"][i], rtol=1e-11, atol=1e-13)


def test_odometry_helpers():
    assert eo.delta_phi(1010, 1000, 135) == 10 * (2 * np.pi / 135)
    ang, lin = eo.displacement(0.0318, 0.1, 0.2, 0.5)
    assert ang == (0.0318 * 0.5 - 0.0318 * 0.2) / 0.1 and lin == (0.0318 * 0.2 + 0.0318 * 0.5) / 2
    assert eo.wrap_pi(np.pi) == -np.pi and abs(eo.wrap_pi(-np.pi) + np.pi) < 1e-15


def test_tracked_rows_oracle_equals_full_oracle():
    """oracle/sampled_rows.TrackedRows (the XL leg: the rank-2 restatement on a subset of rows) == the full rank-2
    oracle restricted to those rows, through a mixed sequence of updates and predicts."""
    from oracle import sampled_rows as sr
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_large_map(100, updates=30)
    m_o, c_o = w.mu0.copy(), w.p0()
    fl = eo.Flags(motion_var=0.2, meas_var=0.5)
    slots = [op[1] for op in w.ops if op[0] == "update"]
    rs = np.random.default_rng(0)
    rows = sr.required_rows(slots)
    rows += [int(r) for r in rs.permutation(w.n) if r not in rows][:17]
    tr = sr.TrackedRows(w.mu0, rows, c_o[rows, :])
    for op in w.ops:
        if op[0] == "update":
            eo.update_rank(m_o, c_o, op[1], op[2], op[3], 0.5)
            tr.update(op[1], op[2], op[3], 0.5)
        else:
            dx, dy, th, g0, g1 = eo.motion_step(m_o[2], op[1], op[2], fl)
            m_o[0] += dx; m_o[1] += dy; m_o[2] = th
            eo.predict_cov_rank(c_o, g0, g1, 0.2)
            tr.predict(op[1], op[2], fl)
    e = tr.errors(m_o[rows], c_o[rows, :], np.diag(c_o))
    assert max(e) <= 1e-12, e
    with pytest.raises(KeyError):
        sr.TrackedRows(w.mu0, [0, 1, 2], w.p0()[[0, 1, 2], :]).update(5, 1.0, 0.1, 0.5)
