"""CPU: host-side logic of the product (association bookkeeping, the restated replay event loop, workload
generators) checked against the oracle and the reference's golden replay. The EKF arithmetic itself is
injected from the oracle here — on the product path it is always the CUDA library."""
import numpy as np

from conftest import assert_parity, golden
from oracle import ekf_oracle as eo


def test_association_bookkeeping_bit_exact():
    from slam_duckietown_b200 import ekf_bindings as eb
    z = golden("closed_loop_n20.npz")
    ti_a, ti_b = {}, {}
    ds = z["det_step"]
    mean = np.array([0.3, -0.2, 0.7])
    for t in range(0, 120):
        sel = np.flatnonzero(ds == t)
        pairs = [(str(t), [eo.Detection(int(z["det_tag"][k]), z["det_pose_t"][k].reshape(3, 1), float(z["det_err"][k]))
                           for k in sel if z["det_img"][k] == im]) for im in range(3)]
        a = eb.associate_window(mean, pairs, ti_a, ())
        b = eo.associate_and_average(mean, pairs, ti_b, ())
        assert ti_a == ti_b and list(ti_a.items()) == list(ti_b.items())
        assert list(a.keys()) == list(b.keys())
        for s in a:
            assert a[s][3] == b[s][3]
            np.testing.assert_allclose(a[s][:3] + a[s][4:], [float(x) for x in b[s][:3] + b[s][4:]], rtol=1e-14, atol=1e-15)


def test_gate_and_ignore_edges():
    from slam_duckietown_b200 import ekf_bindings as eb
    d = [eo.Detection(4, np.array([[1.5], [0.0], [0.0]])),             # exactly 1.5 m: kept (strict >)
         eo.Detection(5, np.array([[1.2], [0.0], [0.9000001]])),       # outside
         eo.Detection(6, np.array([[0.1], [0.0], [0.5]])),             # ignored
         eo.Detection(7, np.array([[0.0], [0.0], [1.0]]))]
    ti = {}
    out = eb.associate_window(np.zeros(3), [("0", d)], ti, (6,))
    assert ti == {4: 0, 7: 1} and list(out.keys()) == [0, 1]
    assert out[0][3] == 4 and out[0][4] == 1.5


def test_replay_event_loop_matches_reference_replay(surrogate):
    """The restated event loop (windowing, tick latching, sentinels) with the ORACLE arithmetic injected
    reproduces the reference's own replay() outputs step by step."""
    from slam_duckietown_b200 import replay
    lines, dets, ref = surrogate

    def ekf(ang, lin, mean, cov, dt, pairs, ti):
        return eo.ekf_pose_estimation(ang, lin, mean, cov, dt, pairs, ti)

    class HostOdom:      # a1/a2 without touching the CUDA library
        delta_phi = staticmethod(eo.delta_phi)
        displacement = staticmethod(eo.displacement)
        _flag = staticmethod(lambda name: 0.1)
        EKF_pose_estimation = None
    saved = replay.eb
    replay.eb = HostOdom
    try:
        recs = replay.replay_events(lines, dets, ekf=ekf)
    finally:
        replay.eb = saved
    assert len(recs) == 70
    for i, r in enumerate(recs):
        n = int(ref["n"][i])
        assert r["ang"] == ref["ang"][i] and r["lin"] == ref["lin"][i]
        assert r["order"] == [int(x) for x in ref["order"][i] if x >= 0]
        assert_parity(r["mean"], r["cov"], ref["mean"][i][:n], ref["cov"][i][:n, :n], what=f"step {i}")
    assert list(recs[-1]["tag_index"].items()) == list(zip(ref["tag_ids"].tolist(), ref["tag_slots"].tolist()))


def test_workloads_are_deterministic_and_gate_compliant():
    from slam_duckietown_b200 import workloads as wl
    a, b = wl.make_batched(16, 20, 3, 12), wl.make_batched(16, 20, 3, 12)
    for k in ("ang", "lin", "slot", "rng", "brg", "motion_var", "meas_var"):
        np.testing.assert_array_equal(getattr(a, k), getattr(b, k))
    assert a.rng.max() < 1.5 and a.warm_steps == 7
    z = golden("batched_b48_l20.npz")
    w = wl.make_batched(48, 20, 3, 20)
    np.testing.assert_array_equal(w.slot, z["slot"])
    np.testing.assert_array_equal(w.ang, z["ang"])
    lw = wl.make_large_map(64, updates=9)
    assert sum(o[0] == "update" for o in lw.ops) == 9 and sum(o[0] == "predict" for o in lw.ops) == 3
    p = lw.p0()
    assert p.shape == (131, 131) and np.all(np.linalg.eigvalsh(p) > 0) and np.count_nonzero(p) == p.size
