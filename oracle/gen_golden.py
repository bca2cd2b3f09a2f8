"""TEST INFRASTRUCTURE — dumps golden vectors from the UNMODIFIED reference into tests/golden/.

    python oracle/gen_golden.py            # needs /root/reference; ~2 min (the 4096-landmark steps dominate)

The reference has no tests or golden vectors of its own, so these vectors — outputs of the
reference's own functions on seeded inputs — are what pins parity on machines where /root/reference does not
exist (the GPU box). Every file records the inputs next to the outputs.
"""
from __future__ import annotations

import gzip
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader, surrogate_bag  # noqa: E402
from oracle.ekf_oracle import Detection  # noqa: E402
import slam_duckietown_b200.workloads as wl  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def det_from_rb(tag_id, r, b, err=0.0):
    x_rel, y_rel = r * np.cos(b), r * np.sin(b)
    return Detection(int(tag_id), np.array([[-y_rel], [0.0], [x_rel]]), float(err))


def gen_surrogate(ref):
    lines, dets, lms, tag_ids = surrogate_bag.build(os.path.join(ref_loader.REFERENCE_ROOT, "bags"))
    d = os.path.join(OUT, "surrogate")
    os.makedirs(d, exist_ok=True)
    with gzip.open(os.path.join(d, "events.csv.gz"), "wt") as fh:
        fh.writelines(lines)
    names = sorted(dets)
    off, tid, pt, pe = [0], [], [], []
    for nm in names:
        for (t, p, e) in dets[nm]:
            tid.append(t)
            pt.append(p)
            pe.append(e)
        off.append(len(tid))
    np.savez_compressed(os.path.join(d, "detections.npz"), names=np.array(names), offsets=np.array(off),
                        tag_id=np.array(tid, dtype=np.int64), pose_t=np.array(pt), pose_err=np.array(pe),
                        landmarks=lms, landmark_tag_ids=np.array(tag_ids))
    det_objs = {nm: [Detection(t, np.asarray(p).reshape(3, 1), e) for (t, p, e) in dets[nm]] for nm in names}
    with tempfile.TemporaryDirectory() as tmp:
        with open(os.path.join(tmp, "events.csv"), "w") as fh:
            fh.writelines(lines)
        log = ref_loader.run_reference_replay(tmp, det_objs)
    nmax = max(len(r["mean"]) for r in log)
    T = len(log)
    mean = np.zeros((T, nmax))
    cov = np.zeros((T, nmax, nmax))
    n = np.zeros(T, dtype=np.int64)
    order = np.full((T, 16), -1, dtype=np.int64)
    for i, r in enumerate(log):
        k = len(r["mean"])
        n[i] = k
        mean[i, :k] = r["mean"]
        cov[i, :k, :k] = r["cov"]
        order[i, :len(r["order"])] = r["order"]
    final = log[-1]["tag_index"]
    np.savez_compressed(os.path.join(d, "replay_ref.npz"), mean=mean, cov=cov, n=n, order=order,
                        ang=np.array([r["ang"] for r in log]), lin=np.array([r["lin"] for r in log]),
                        tag_ids=np.array(list(final.keys())), tag_slots=np.array(list(final.values())))
    print(f"surrogate: {len(lines)} events, {len(names)} images, {T} EKF steps, final n={n[-1]}, "
          f"max updates/step={int((order >= 0).sum(1).max())}")


def gen_closed_loop(ref, steps=300, L=20, seed=5):
    """Seeded synthetic closed loop through the reference with RAW detections (gate, ignore list, 3 images
    per window). Inputs are stored flattened; outputs every 25 steps."""
    rng = np.random.default_rng(seed)
    lms = np.stack([rng.uniform(-1, 3, L), rng.uniform(-2, 2, L)], 1)
    ids = rng.permutation(400)[:L]
    pose = np.zeros(3)
    mean, cov, ti = np.zeros(3), np.eye(3) * 0.1, {}
    d_step, d_img, d_tag, d_pt, d_err = [], [], [], [], []
    angs, lins, keep_t, keep_mean, keep_cov, keep_n = [], [], [], [], [], []
    for t in range(steps):
        ang, lin = float(rng.normal(0.08, 0.05)), float(abs(rng.normal(0.05, 0.02)))
        if t % 17 == 5:
            ang = float(rng.uniform(-9e-3, 9e-3))      # exercise the straight branch (:376)
        pairs = []
        for im in range(3):
            dets = []
            for k in range(L):
                d = lms[k] - pose[:2]
                r = np.hypot(*d)
                b = (np.arctan2(d[1], d[0]) - pose[2] + np.pi) % (2 * np.pi) - np.pi
                if r < 1.8 and abs(b) < 1.0:
                    xr, yr = r * np.cos(b) + rng.normal(0, 0.02), r * np.sin(b) + rng.normal(0, 0.02)
                    pt = np.array([[-yr], [0.0], [xr]])
                    e = float(rng.uniform())
                    dets.append(Detection(int(ids[k]), pt, e))
                    d_step.append(t); d_img.append(im); d_tag.append(int(ids[k])); d_pt.append(pt.ravel()); d_err.append(e)
            pairs.append((str(t), dets))
        mean, cov, _ = ref.EKF_pose_estimation(ang, lin, mean, cov, 0.7, pairs, ti)
        angs.append(ang); lins.append(lin)
        if t % 25 == 24 or t == steps - 1:
            keep_t.append(t); keep_n.append(len(mean))
            m = np.zeros(3 + 2 * L); m[:len(mean)] = mean
            c = np.zeros((3 + 2 * L,) * 2); c[:len(mean), :len(mean)] = cov
            keep_mean.append(m); keep_cov.append(c)
        if abs(ang) <= 1e-2:
            pose = pose + np.array([lin * np.cos(pose[2]), lin * np.sin(pose[2]), 0])
        else:
            r = lin / ang
            pose = np.array([pose[0] - r * np.sin(pose[2]) + r * np.sin(pose[2] + ang),
                             pose[1] + r * np.cos(pose[2]) - r * np.cos(pose[2] + ang), pose[2] + ang])
    np.savez_compressed(os.path.join(OUT, "closed_loop_n20.npz"), ang=np.array(angs), lin=np.array(lins),
                        det_step=np.array(d_step), det_img=np.array(d_img), det_tag=np.array(d_tag),
                        det_pose_t=np.array(d_pt), det_err=np.array(d_err), keep_t=np.array(keep_t),
                        keep_n=np.array(keep_n), keep_mean=np.array(keep_mean), keep_cov=np.array(keep_cov),
                        tag_ids=np.array(list(ti.keys())), tag_slots=np.array(list(ti.values())))
    print(f"closed loop: {steps} steps, {len(d_tag)} detections, final n={len(mean)}")


def gen_batched(ref, name, B, L, m, steps, keep_every):
    """First B filters of the batched workload replayed through the reference, one filter at a time."""
    w = wl.make_batched(B, n_landmarks=L, m=m, steps=steps)
    n = 3 + 2 * L
    keep_t = [t for t in range(steps) if t % keep_every == keep_every - 1 or t == steps - 1]
    means = np.zeros((len(keep_t), B, n))
    covs = np.zeros((len(keep_t), B, n, n))
    rr = np.zeros_like(w.rng)
    bb = np.zeros_like(w.brg)
    for b in range(B):
        ref.MOTION_MODEL_VARIANCE = float(w.motion_var[b])
        ref.MEASUREMENT_MODEL_VARIANCE = float(w.meas_var[b])
        mean, cov, ti = np.zeros(3), np.eye(3) * 0.1, {}
        for t in range(steps):
            dets = [det_from_rb(100 + int(w.slot[t, b, k]), w.rng[t, b, k], w.brg[t, b, k])
                    for k in range(m) if w.slot[t, b, k] >= 0]
            mean, cov, tags = ref.EKF_pose_estimation(float(w.ang[t, b]), float(w.lin[t, b]), mean, cov, 0.7,
                                                      [("0", dets)], ti)
            got = [int(s) for s in tags.keys()]
            want = [int(s) for s in w.slot[t, b] if s >= 0]
            assert got == want, (t, b, got, want)        # nothing gated, slots in first-seen order
            for k, s in enumerate(got):
                rr[t, b, k], bb[t, b, k] = tags[s][4], tags[s][5]
            if t in keep_t:
                i = keep_t.index(t)
                means[i, b, :len(mean)] = mean
                covs[i, b, :len(mean), :len(mean)] = cov
    ref.MOTION_MODEL_VARIANCE, ref.MEASUREMENT_MODEL_VARIANCE = 0.1, 0.7
    np.savez_compressed(os.path.join(OUT, name), n_landmarks=L, ang=w.ang, lin=w.lin, slot=w.slot, rng=rr, brg=bb,
                        motion_var=w.motion_var, meas_var=w.meas_var, keep_t=np.array(keep_t), mean=means,
                        cov=covs)
    print(f"{name}: B={B} L={L} m={m} steps={steps}")


def gen_units(ref):
    """Single calls covering the three motion branches, flags, gate / ignore edges and wrap edges."""
    rng = np.random.default_rng(11)
    cases = []

    def spd(n):
        a = rng.normal(size=(n, n))
        return a @ a.T / n + 0.05 * np.eye(n)

    base_ti = {7: 0, 9: 1, 4: 2}
    for name, ang, lin, flags in [
        ("straight", 5e-3, 0.07, {}), ("straight_edge", 1e-2, 0.07, {}), ("arc", 0.3, 0.11, {}),
        ("arc_wrap", 1.2, 0.2, {}), ("arc_neg", -0.7, 0.05, {}),
        ("linear", 0.4, 0.09, {"ENABLE_CIRCULAR_INTERPOLATION": False}),
        ("nomotion", 0.4, 0.09, {"DISABLE_MOTION_MODEL": True}),
        ("nomeas", 0.2, 0.09, {"ENABLE_MEASUREMENT_MODEL": False}),
        ("ignore", 0.2, 0.09, {"IGNORE_TAGS": [9]}),
        ("noise", 0.2, 0.09, {"MOTION_MODEL_VARIANCE": 0.33, "MEASUREMENT_MODEL_VARIANCE": 1.2}),
    ]:
        n = 9
        mean = np.concatenate([[0.2, -0.1, 3.0 if name == "arc_wrap" else 0.4], rng.uniform(-1, 1, 6)])
        cov = spd(n)
        dets = [det_from_rb(7, 0.8, 0.3, 0.1), det_from_rb(9, 1.2, -0.5, 0.2), det_from_rb(7, 0.82, 0.28, 0.3),
                det_from_rb(55, 1.1, 0.9, 0.4),                       # new landmark -> augment
                Detection(4, np.array([[1.5], [0.0], [0.0]]), 0.5),    # exactly on the gate: kept (strict >)
                Detection(4, np.array([[1.2], [0.0], [0.9000001]]), 0.6),   # just outside: dropped
                det_from_rb(66, 1.49, 3.1, 0.7)]                       # bearing near pi -> wrap of the innovation
        cases.append((name, ang, lin, mean, cov, dict(base_ti), dets, flags))
    # empty window on a 3-state filter (the null first step of replay())
    cases.append(("null_step", 0.0, 0.0, np.zeros(3), np.eye(3) * 0.1, {}, [], {}))
    out = {"names": np.array([c[0] for c in cases])}
    defaults = {k: getattr(ref, k) for k in ("ENABLE_CIRCULAR_INTERPOLATION", "DISABLE_MOTION_MODEL",
                                             "ENABLE_MEASUREMENT_MODEL", "IGNORE_TAGS", "MOTION_MODEL_VARIANCE",
                                             "MEASUREMENT_MODEL_VARIANCE")}
    for i, (name, ang, lin, mean, cov, ti, dets, flags) in enumerate(cases):
        for k, v in flags.items():
            setattr(ref, k, v)
        ti_in = dict(ti)
        m2, c2, tags = ref.EKF_pose_estimation(ang, lin, mean.copy(), cov.copy(), 0.7, [("0", dets[:4]), ("1", dets[4:])], ti)
        for k, v in defaults.items():
            setattr(ref, k, v)
        p = f"c{i}_"
        out[p + "ang"], out[p + "lin"] = ang, lin
        out[p + "mean_in"], out[p + "cov_in"] = mean, cov
        out[p + "ti_in_ids"], out[p + "ti_in_slots"] = np.array(list(ti_in.keys()), dtype=np.int64), np.array(list(ti_in.values()), dtype=np.int64)
        out[p + "det_img"] = np.array([0] * len(dets[:4]) + [1] * len(dets[4:]), dtype=np.int64)
        out[p + "det_tag"] = np.array([d.tag_id for d in dets], dtype=np.int64)
        out[p + "det_pose_t"] = np.array([d.pose_t.ravel() for d in dets]).reshape(-1, 3)
        out[p + "det_err"] = np.array([d.pose_err for d in dets])
        out[p + "flags_keys"] = np.array(list(flags.keys()))
        out[p + "flags_vals"] = np.array([repr(v) for v in flags.values()])
        out[p + "mean_out"], out[p + "cov_out"] = m2, c2
        out[p + "ti_out_ids"], out[p + "ti_out_slots"] = np.array(list(ti.keys()), dtype=np.int64), np.array(list(ti.values()), dtype=np.int64)
        out[p + "order"] = np.array(list(tags.keys()), dtype=np.int64)
        out[p + "tags"] = np.array([[float(x) for x in v] for v in tags.values()]).reshape(-1, 6)
    np.savez_compressed(os.path.join(OUT, "unit_cases.npz"), **out)
    print(f"unit cases: {len(cases)}")


def gen_ekf3():
    e3 = ref_loader.load_ekf3_functions()
    rng = np.random.default_rng(21)
    K = 256
    st = rng.uniform(-2, 2, (K, 3)); st[:, 2] = rng.uniform(-3.5, 3.5, K)
    cv = np.zeros((K, 3, 3))
    for i in range(K):
        a = rng.normal(size=(3, 3)); cv[i] = a @ a.T / 3 + 0.05 * np.eye(3)
    v, om, dt = rng.uniform(0, 0.5, K), rng.normal(0, 0.4, K), rng.uniform(0.05, 1.0, K)
    om[::7] = rng.uniform(-1e-6, 1e-6, len(om[::7]))          # straight branch (:38-41)
    obs = np.stack([rng.uniform(0.2, 3, K), rng.uniform(-3.1, 3.1, K)], 1)
    lm = rng.uniform(-3, 3, (K, 2))
    ps, pc, us, uc = (np.zeros_like(st), np.zeros_like(cv), np.zeros_like(st), np.zeros_like(cv))
    for i in range(K):
        ps[i], pc[i] = e3.predict(st[i], cv[i], (v[i], om[i]), dt[i])
        us[i], uc[i] = e3.update(ps[i], pc[i], obs[i], lm[i])
    np.savez_compressed(os.path.join(OUT, "ekf3.npz"), state=st, cov=cv, v=v, omega=om, dt=dt, obs=obs, lm=lm,
                        pred_state=ps, pred_cov=pc, upd_state=us, upd_cov=uc)
    print("ekf3: 256 predict+update vectors")


def gen_large(ref, L, steps, rows_keep, name):
    """`steps` calls (predict + 2 updates each) on a pre-initialised L-landmark map through the reference.
    Stores mu and a sample of rows + the diagonal of the final P (the full P of 4096 landmarks is 537 MB)."""
    w = wl.make_large_map(L, updates=2 * steps)
    ups = [o for o in w.ops if o[0] == "update"]
    prs = [o for o in w.ops if o[0] == "predict"]
    rng = np.random.default_rng(31)
    mean, cov = w.mu0.copy(), w.p0()
    ti = {100 + s: s for s in range(L)}
    sched = []
    for t in range(steps):
        ang, lin = (prs[t][1], prs[t][2]) if t < len(prs) else (float(rng.normal(0.08, 0.05)), 0.05)
        two = ups[2 * t: 2 * t + 2]
        if two[0][1] == two[1][1]:
            two = two[:1]
        # ranges are recomputed from the CURRENT estimate so the innovation stays small and inside the gate
        dets = []
        for (_, s, r, b) in two:
            dets.append(det_from_rb(100 + s, min(r, 1.45), b))
        mean, cov, tags = ref.EKF_pose_estimation(ang, lin, mean, cov, 0.7, [("0", dets)], ti)
        assert len(tags) == len(two)
        sched.append((ang, lin, [(int(s), float(tags[s][4]), float(tags[s][5])) for s in tags]))
    n = len(mean)
    rows = np.unique(np.concatenate([[0, 1, 2], rng.integers(0, n, rows_keep),
                                     [3 + 2 * s for _, _, o in sched for (s, _, _) in o]]))
    flat = []
    for (ang, lin, o) in sched:
        flat.append([ang, lin, len(o)] + [x for (s, r, b) in o for x in (s, r, b)] + [0.0] * (3 * (2 - len(o))))
    np.savez_compressed(os.path.join(OUT, name), n_landmarks=L, sched=np.array(flat), mean=mean, rows=rows,
                        p_rows=cov[rows, :], p_cols=cov[:, rows[:8]], p_diag=np.diag(cov).copy(),
                        p_sum=np.array([cov.sum(), np.abs(cov).sum()]))
    print(f"{name}: L={L} n={n} steps={steps}")


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = ref_loader.load_replay_module()
    gen_surrogate(ref)
    gen_closed_loop(ref)
    gen_units(ref)
    gen_ekf3()
    gen_batched(ref, "batched_b48_l20.npz", B=48, L=20, m=3, steps=20, keep_every=5)
    gen_batched(ref, "batched_b32_l8.npz", B=32, L=8, m=2, steps=12, keep_every=4)
    gen_large(ref, 256, 4, 24, "large_l256.npz")
    gen_large(ref, 1024, 3, 24, "large_l1024.npz")
    if "--skip-4096" not in sys.argv:
        gen_large(ref, 4096, 3, 24, "large_l4096.npz")


if __name__ == "__main__":
    main()
