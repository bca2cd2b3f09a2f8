"""The oracle (oracle/ekf_oracle.py) against the UNMODIFIED reference functions, live, on randomised inputs — the
direct pin that the golden fixtures only give second-hand. Runs wherever the reference can be loaded: from
/root/reference (this container) or from the bytecode oracle/build_ref.py compiled out of it (oracle/_ref, which
travels to the GPU box). Also re-generates two golden fixtures and compares them with the checked-in files, and
checks that oracle/_ref is a faithful compile of the reference sources when both are present."""
import marshal
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import ekf_oracle as eo, ref_loader

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not ref_loader.reference_available(), reason="reference not loadable here")]


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load_replay_module()


def test_delta_phi_and_displacement_exact(ref):
    rs = np.random.default_rng(0)
    for _ in range(500):
        t, p = int(rs.integers(-5000, 5000)), int(rs.integers(-5000, 5000))
        assert ref.delta_phi(t, p, 135) == eo.delta_phi(t, p, 135)
        dl, dr = rs.normal(0, 2), rs.normal(0, 2)
        assert ref.displacement(0.0318, 0.1, dl, dr) == eo.displacement(0.0318, 0.1, dl, dr)


FLAG_NAMES = ("ENABLE_CIRCULAR_INTERPOLATION", "DISABLE_MOTION_MODEL", "ENABLE_MEASUREMENT_MODEL", "IGNORE_TAGS",
              "MOTION_MODEL_VARIANCE", "MEASUREMENT_MODEL_VARIANCE")


@pytest.mark.parametrize("variant", ["default", "linear", "nomotion", "nomeas", "ignore", "noise"])
def test_pose_estimation_closed_loop_vs_reference(ref, variant):
    """60 closed-loop calls with raw detections (gate edge, ignored ids, growth of the state): every call's outputs of
    the oracle's restatement == the reference's, to rounding (both forms), with identical integer bookkeeping."""
    over = {"default": {}, "linear": {"ENABLE_CIRCULAR_INTERPOLATION": False}, "nomotion": {"DISABLE_MOTION_MODEL": True},
            "nomeas": {"ENABLE_MEASUREMENT_MODEL": False}, "ignore": {"IGNORE_TAGS": [17, 3]},
            "noise": {"MOTION_MODEL_VARIANCE": 0.31, "MEASUREMENT_MODEL_VARIANCE": 1.1}}[variant]
    saved = {k: getattr(ref, k) for k in FLAG_NAMES}
    fl = eo.Flags(motion_var=over.get("MOTION_MODEL_VARIANCE", 0.1), meas_var=over.get("MEASUREMENT_MODEL_VARIANCE", 0.7),
                  enable_measurement=over.get("ENABLE_MEASUREMENT_MODEL", True),
                  circular=over.get("ENABLE_CIRCULAR_INTERPOLATION", True),
                  disable_motion=over.get("DISABLE_MOTION_MODEL", False), ignore_tags=tuple(over.get("IGNORE_TAGS", ())))
    rs = np.random.default_rng(7)
    L = 9
    lms = np.stack([rs.uniform(-0.8, 1.6, L), rs.uniform(-1.2, 1.2, L)], 1)
    ids = [17, 3, 44, 80, 101, 5, 26, 300, 65]
    pose = np.zeros(3)
    m_r, c_r, ti_r = np.zeros(3), np.eye(3) * 0.1, {}
    m_o, c_o, ti_o = np.zeros(3), np.eye(3) * 0.1, {}
    m_d, c_d, ti_d = np.zeros(3), np.eye(3) * 0.1, {}
    try:
        for k, v in over.items():
            setattr(ref, k, v)
        for t in range(60):
            ang = float(rs.normal(0.08, 0.05)) if t % 9 else float(rs.uniform(-9e-3, 9e-3))
            lin = float(abs(rs.normal(0.05, 0.02)))
            pairs = []
            for im in range(3):
                dets = []
                for q in range(L):
                    d = lms[q] - pose[:2]
                    r = np.hypot(*d)
                    b = (np.arctan2(d[1], d[0]) - pose[2] + np.pi) % (2 * np.pi) - np.pi
                    if r < 1.7 and abs(b) < 1.2:
                        xr, yr = r * np.cos(b) + rs.normal(0, 0.02), r * np.sin(b) + rs.normal(0, 0.02)
                        dets.append(eo.Detection(ids[q], np.array([[-yr], [0.0], [xr]]), float(rs.uniform())))
                pairs.append((str(t), dets))
            m_r, c_r, tg_r = ref.EKF_pose_estimation(ang, lin, m_r, c_r, 0.7, pairs, ti_r)
            m_o, c_o, tg_o = eo.ekf_pose_estimation(ang, lin, m_o, c_o, 0.7, pairs, ti_o, fl)
            m_d, c_d, _ = eo.ekf_pose_estimation(ang, lin, m_d, c_d, 0.7, pairs, ti_d, fl, form="dense")
            assert list(ti_r.items()) == list(ti_o.items()) == list(ti_d.items())
            assert list(tg_r.keys()) == list(tg_o.keys())
            for s in tg_r:
                assert tg_r[s][3] == tg_o[s][3]
                np.testing.assert_allclose([float(x) for x in tg_r[s]][4:], [float(x) for x in tg_o[s]][4:], rtol=1e-13, atol=1e-13)
                # the global guess adds the running pose estimate, which carries the filters' own rounding difference
                np.testing.assert_allclose([float(x) for x in tg_r[s]][:3], [float(x) for x in tg_o[s]][:3], rtol=0, atol=1e-9)
            assert m_r.shape == m_o.shape == m_d.shape
            e = eo.parity_errors(m_o, c_o, m_r, c_r)
            assert max(e) <= 1e-10, (variant, t, e)
            e = eo.parity_errors(m_d, c_d, m_r, c_r)
            assert max(e) <= 1e-10, (variant, "dense", t, e)
            if abs(ang) <= 1e-2:
                pose = pose + np.array([lin * np.cos(pose[2]), lin * np.sin(pose[2]), 0])
            else:
                r = lin / ang
                pose = np.array([pose[0] - r * np.sin(pose[2]) + r * np.sin(pose[2] + ang),
                                 pose[1] + r * np.cos(pose[2]) - r * np.cos(pose[2] + ang), pose[2] + ang])
    finally:
        for k, v in saved.items():
            setattr(ref, k, v)
    assert len(m_r) > 3


def test_three_state_surface_vs_reference():
    e3 = ref_loader.load_ekf3_functions()
    rs = np.random.default_rng(3)
    for i in range(200):
        st = rs.uniform(-2, 2, 3)
        a = rs.normal(size=(3, 3))
        cv = a @ a.T / 3 + 0.05 * np.eye(3)
        v, om, dt = rs.uniform(0, 0.5), (rs.normal(0, 0.4) if i % 7 else rs.uniform(-1e-6, 1e-6)), rs.uniform(0.05, 1.0)
        s_r, c_r = e3.predict(st, cv, (v, om), dt)
        s_o, c_o = eo.predict3(st, cv, (v, om), dt)
        np.testing.assert_allclose(s_o, s_r, rtol=0, atol=1e-14)
        np.testing.assert_allclose(c_o, c_r, rtol=1e-13, atol=1e-15)
        obs, lm = (rs.uniform(0.2, 3), rs.uniform(-3.1, 3.1)), rs.uniform(-3, 3, 2)
        u_r, uc_r = e3.update(s_r, c_r, obs, lm)
        u_o, uc_o = eo.update3(s_r, c_r, obs, lm)
        np.testing.assert_allclose(u_o, u_r, rtol=0, atol=1e-12)
        np.testing.assert_allclose(uc_o, uc_r, rtol=1e-11, atol=1e-13)


def test_golden_fixtures_regenerate_identically(ref, tmp_path, monkeypatch):
    """The committed generator run again gives the committed bytes' CONTENT (array by array, bitwise)."""
    from oracle import gen_golden
    monkeypatch.setattr(gen_golden, "OUT", str(tmp_path))
    gen_golden.gen_units(ref)
    gen_golden.gen_batched(ref, "batched_b32_l8.npz", B=32, L=8, m=2, steps=12, keep_every=4)
    gen_golden.gen_ekf3()
    for name in ("unit_cases.npz", "batched_b32_l8.npz", "ekf3.npz"):
        new, old = np.load(tmp_path / name), np.load(os.path.join(GOLDEN, name))
        assert sorted(new.files) == sorted(old.files), name
        for k in new.files:
            assert np.array_equal(new[k], old[k]), (name, k)


@pytest.mark.skipif(not (ref_loader.source_available() and ref_loader.compiled_available()),
                    reason="needs both /root/reference and oracle/_ref")
def test_compiled_reference_is_the_unmodified_source():
    """oracle/_ref/replay_no_ros.bytecode == compile(/root/reference/src/replay_no_ros.py): the travelling build output
    is the reference's own code, nothing added or removed."""
    src = os.path.join(ref_loader.REFERENCE_ROOT, "src", "replay_no_ros.py")
    with open(src) as fh:
        want = compile(fh.read(), src, "exec", dont_inherit=True)
    with open(os.path.join(ref_loader.REF_BUILD_DIR, "replay_no_ros.bytecode"), "rb") as fh:
        fh.read(16)
        got = marshal.load(fh)
    assert marshal.dumps(got) == marshal.dumps(want)
