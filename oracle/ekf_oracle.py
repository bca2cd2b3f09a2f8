"""TEST INFRASTRUCTURE — CPU (numpy, float64) restatement of the reference EKF-SLAM hot path.

This is the ORACLE: the checker the CUDA path is compared against. It is not a product path and nothing
under ``slam-duckietown_b200/`` imports it. Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may call it.

Parity pinning: the reference has no tests or golden vectors of its own, so this restatement
is pinned against the UNMODIFIED reference functions imported in-process (``oracle/ref_loader.py``) by
``tests/test_oracle_vs_reference.py`` (runs where ``/root/reference`` exists) and against the golden vectors
``oracle/gen_golden.py`` dumped from the reference into ``tests/golden/`` (runs everywhere).

Reference lines followed (all in ``/root/reference/src``):
  delta_phi            replay_no_ros.py:250-266
  displacement         replay_no_ros.py:484-497
  gate + associate     replay_no_ros.py:280-301
  window average       replay_no_ros.py:311-337
  augment              replay_no_ros.py:341-360
  predict mean + G     replay_no_ros.py:368-417
  predict covariance   replay_no_ros.py:420-430
  update               replay_no_ros.py:435-480
  3-state predict      EKF-SLAM.py:29-56 (noise :12)
  3-state update       EKF-SLAM.py:59-84 (noise :13)

Two arithmetic forms are provided:
  form="rank"   O(n^2): the same algebra written as row/column gathers and a rank-2 (plain) or rank-4
                (Joseph) update — the shape the CUDA kernels use.
  form="dense"  O(n^3): dense G_F / H / (I-KH) matrices multiplied out, i.e. the reference's own cost
                profile (:430, :473, :480) — used as the "reference CPU path" timing baseline.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

TWO_PI = 2.0 * np.pi


@dataclass
class Flags:
    """Module constants the reference reads at call time (replay_no_ros.py:15-36)."""
    motion_var: float = 0.1          # MOTION_MODEL_VARIANCE (:15)
    meas_var: float = 0.7            # MEASUREMENT_MODEL_VARIANCE (:16)
    enable_measurement: bool = True  # ENABLE_MEASUREMENT_MODEL (:18)
    circular: bool = True            # ENABLE_CIRCULAR_INTERPOLATION (:19)
    disable_motion: bool = False     # DISABLE_MOTION_MODEL (:28)
    ignore_tags: tuple = ()          # IGNORE_TAGS (:36)
    joseph: bool = False             # False = plain (I-KH)P as the reference (:480)


@dataclass
class Detection:
    """Duck type of a dt_apriltags detection as consumed by replay_no_ros.py:128,:286-301."""
    tag_id: int
    pose_t: np.ndarray                      # (3,1) float64
    pose_err: float = 0.0
    pose_R: np.ndarray = field(default_factory=lambda: np.eye(3))
    center: np.ndarray = field(default_factory=lambda: np.zeros(2))
    corners: np.ndarray = field(default_factory=lambda: np.zeros((4, 2)))


def wrap_pi(a: float) -> float:
    """(a + pi) % (2 pi) - pi with Python/numpy float-modulo semantics (:397, :458)."""
    r = math.fmod(a + math.pi, TWO_PI)
    if r < 0.0:
        r += TWO_PI
    return r - math.pi


def delta_phi(ticks, prev_ticks, resolution):
    """replay_no_ros.py:262-266."""
    return (ticks - prev_ticks) * (2 * np.pi / resolution)


def displacement(R, baseline, dphi_left, dphi_right):
    """replay_no_ros.py:491-497 — returns (angular, linear)."""
    d_r = R * dphi_right
    d_l = R * dphi_left
    return (d_r - d_l) / baseline, (d_l + d_r) / 2


def associate_and_average(mean, pairs, tag_index, ignore_tags=()):
    """Gate, slot assignment, per-slot window average and global guess (replay_no_ros.py:280-337).

    Mutates ``tag_index`` (:295). Returns dict slot -> [gx, gy, err, tag_id, range, bearing] in
    first-seen-in-window order (dict insertion order, :313).
    """
    buckets = {}
    for _ts, dets in pairs:
        for det in dets:
            if det.tag_id in ignore_tags:                       # :286
                continue
            tx, tz = det.pose_t[0][0], det.pose_t[2][0]
            if tz * tz + tx * tx > 1.5 ** 2:                    # :289 (strict >)
                continue
            if det.tag_id not in tag_index:                     # :294-295
                tag_index[det.tag_id] = len(tag_index)
            buckets.setdefault(tag_index[det.tag_id], []).append(det)
    inv = {slot: tid for tid, slot in tag_index.items()}        # :311
    out = {}
    for slot, dets in buckets.items():
        k = len(dets)
        # np.mean over the window == pairwise/sequential sum / k (bit-exact for k <= 8)
        t = np.mean([d.pose_t for d in dets], axis=0)           # :315
        err = np.mean([d.pose_err for d in dets], axis=0)       # :317
        x_rel, y_rel = t[2][0], -t[0][0]                        # :321
        rng = np.sqrt(x_rel ** 2 + y_rel ** 2)                  # :329
        brg = np.arctan2(y_rel, x_rel)                          # :330
        out[slot] = [mean[0] + rng * np.cos(brg + mean[2]),     # :331 (PRE-predict pose)
                     mean[1] + rng * np.sin(brg + mean[2]),     # :332
                     err, inv[slot], rng, brg]
        del k
    return out


def augment(mean, cov, n_slots, tags_positions):
    """Grow (mean, cov) to 3+2*n_slots (replay_no_ros.py:341-360). KeyError like :359 if a new slot was not
    observed in this window."""
    n_old, n_new = len(mean), 3 + 2 * n_slots
    if n_old >= n_new:
        return mean, cov
    m2 = np.zeros(n_new)
    m2[:n_old] = mean
    c2 = np.zeros((n_new, n_new))
    c2[:n_old, :n_old] = cov
    for i in range(n_old, n_new, 2):
        slot = (i - 3) // 2
        c2[i, i] = 10000.0
        c2[i + 1, i + 1] = 10000.0
        m2[i] = tags_positions[slot][0]
        m2[i + 1] = tags_positions[slot][1]
    return m2, c2


def motion_step(theta, ang, lin, flags: Flags):
    """Pose increment (dx, dy), new theta and Jacobian entries g0=G[0,2], g1=G[1,2] (:368-417)."""
    if flags.disable_motion:                                    # :372-373
        return 0.0, 0.0, theta, 0.0, 0.0
    c, s = np.cos(theta), np.sin(theta)
    if flags.circular:
        if abs(ang) <= 1e-2:                                    # :376-389 straight; theta untouched
            return lin * c, lin * s, theta + 0, -lin * s, lin * c
        r = lin / ang                                           # :391
        s2, c2 = np.sin(theta + ang), np.cos(theta + ang)
        dx = -r * s + r * s2                                    # :393
        dy = r * c - r * c2                                     # :394
        th = wrap_pi(theta + ang)                               # :395-397
        return dx, dy, th, -r * c + r * c2, -r * s + r * s2     # :401-402
    # linear interpolation (:405-417): theta += ang, NOT wrapped
    return lin * c, lin * s, theta + ang, -lin * s, lin * c


def innovation(mean, j, rng, brg):
    """delta, q, wrapped innovation y and the 2x5 block of H over columns {0,1,2,j,j+1} (:443-469)."""
    dx = mean[j] - mean[0]
    dy = mean[j + 1] - mean[1]
    q = dx * dx + dy * dy                                       # :446
    sq = np.sqrt(q)
    y0 = rng - sq                                               # :455
    y1 = wrap_pi(brg - (np.arctan2(dy, dx) - mean[2]))          # :453,:458
    h5 = np.array([[-sq * dx, -sq * dy, 0.0, sq * dx, sq * dy],
                   [dy, -dx, -q, -dy, dx]]) / q                 # :466-469
    return np.array([y0, y1]), h5


def _inv2(s):
    det = s[0, 0] * s[1, 1] - s[0, 1] * s[1, 0]
    return np.array([[s[1, 1], -s[0, 1]], [-s[1, 0], s[0, 0]]]) / det


def predict_cov_rank(cov, g0, g1, mv):
    """In-place P <- G_F P G_F^T + F^T R F in O(n) (Appendix A.1; reference :420-430)."""
    cov[0, :] += g0 * cov[2, :]
    cov[1, :] += g1 * cov[2, :]
    cov[:, 0] += g0 * cov[:, 2]
    cov[:, 1] += g1 * cov[:, 2]
    cov[0, 0] += mv * mv
    cov[1, 1] += mv * mv
    cov[2, 2] += (mv / 2) * (mv / 2)
    return cov


def update_rank(mean, cov, slot, rng, brg, mm, joseph=False):
    """One sequential range-bearing update, in place, O(n^2) (Appendix A.2/A.4; reference :436-480).

    K comes from COLUMNS P[:, idx5]; HP from ROWS P[idx5, :] — never substitute one for the other
    (SURVEY.md §0.4).
    """
    j = 3 + 2 * slot
    idx = [0, 1, 2, j, j + 1]
    y, h5 = innovation(mean, j, rng, brg)
    pht = cov[:, idx] @ h5.T                                    # n x 2  (columns)
    hp = h5 @ cov[idx, :]                                       # 2 x n  (rows)
    q2 = np.array([[mm * mm, 0.0], [0.0, mm * mm]])             # :438
    s = hp[:, idx] @ h5.T + q2                                  # (H P) H^T + Q
    k = pht @ _inv2(s)
    mean += k @ y                                               # :476
    if not joseph:
        cov -= k @ hp                                           # :480
    else:
        aht = pht - k @ (s - q2)
        w = hp - q2 @ k.T
        cov -= k @ w
        cov -= aht @ k.T
    return mean, cov


def _predict_cov_dense(cov, g0, g1, mv):
    n = cov.shape[0]
    gf = np.eye(n)
    gf[0, 2], gf[1, 2] = g0, g1
    noise = np.zeros((n, n))
    noise[0, 0] = noise[1, 1] = mv ** 2
    noise[2, 2] = (mv / 2) ** 2
    return gf @ cov @ gf.T + noise


def _update_dense(mean, cov, slot, rng, brg, mm, joseph=False):
    n = cov.shape[0]
    j = 3 + 2 * slot
    y, h5 = innovation(mean, j, rng, brg)
    h = np.zeros((2, n))
    h[:, [0, 1, 2, j, j + 1]] = h5
    q2 = np.diag([mm ** 2, mm ** 2])
    k = cov @ h.T @ np.linalg.inv(h @ cov @ h.T + q2)
    mean = mean + k @ y
    ikh = np.eye(n) - k @ h
    if joseph:
        cov = ikh @ cov @ ikh.T + k @ q2 @ k.T
    else:
        cov = ikh @ cov
    return mean, cov


def ekf_pose_estimation(ang, lin, mean, cov, delta_t, pairs, tag_index, flags: Flags | None = None,
                        form: str = "rank"):
    """Restatement of EKF_pose_estimation (replay_no_ros.py:269-482). Returns NEW arrays
    (mean, cov, tags_positions); mutates ``tag_index`` only. ``delta_t`` is unused (as in the reference)."""
    flags = flags or Flags()
    tags = associate_and_average(mean, pairs, tag_index, flags.ignore_tags)
    mean = np.array(mean, dtype=np.float64, copy=True)
    cov = np.array(cov, dtype=np.float64, copy=True)
    mean, cov = augment(mean, cov, len(tag_index), tags)
    dx, dy, th, g0, g1 = motion_step(mean[2], ang, lin, flags)
    mean[0] += dx
    mean[1] += dy
    mean[2] = th
    if form == "dense":
        cov = _predict_cov_dense(cov, g0, g1, flags.motion_var)
    else:
        cov = predict_cov_rank(cov, g0, g1, flags.motion_var)
    if flags.enable_measurement:
        for slot, rec in tags.items():
            if form == "dense":
                mean, cov = _update_dense(mean, cov, slot, rec[4], rec[5], flags.meas_var, flags.joseph)
            else:
                mean, cov = update_rank(mean, cov, slot, rec[4], rec[5], flags.meas_var, flags.joseph)
    return mean, cov, tags


def step_from_observations(mean, cov, ang, lin, obs, flags: Flags | None = None, n_lm_known=None,
                           form: str = "rank"):
    """Steady-state / batched form of one filter-step on a FIXED-capacity state: ``obs`` is a list of
    (slot, range, bearing) already associated and averaged. Slots >= n_lm_known are augmented first from
    the pre-predict pose (Appendix A.3). In place on copies; returns (mean, cov, n_lm_known)."""
    flags = flags or Flags()
    mean = np.array(mean, dtype=np.float64, copy=True)
    cov = np.array(cov, dtype=np.float64, copy=True)
    if n_lm_known is not None:
        for slot, rng, brg in obs:
            if slot >= n_lm_known:
                assert slot == n_lm_known, "slots are handed out in first-seen order (:295)"
                j = 3 + 2 * slot
                mean[j] = mean[0] + rng * np.cos(brg + mean[2])
                mean[j + 1] = mean[1] + rng * np.sin(brg + mean[2])
                cov[j, j] = 10000.0
                cov[j + 1, j + 1] = 10000.0
                n_lm_known += 1
    dx, dy, th, g0, g1 = motion_step(mean[2], ang, lin, flags)
    mean[0] += dx
    mean[1] += dy
    mean[2] = th
    if form == "dense":      # the reference's own O(n^3) cost profile (:430, :473, :480)
        cov = _predict_cov_dense(cov, g0, g1, flags.motion_var)
        if flags.enable_measurement:
            for slot, rng, brg in obs:
                mean, cov = _update_dense(mean, cov, slot, rng, brg, flags.meas_var, flags.joseph)
        return mean, cov, n_lm_known
    predict_cov_rank(cov, g0, g1, flags.motion_var)
    if flags.enable_measurement:
        for slot, rng, brg in obs:
            update_rank(mean, cov, slot, rng, brg, flags.meas_var, flags.joseph)
    return mean, cov, n_lm_known


# ---------------------------------------------------------------------------------------------------------
# 3-state localisation EKF (src/EKF-SLAM.py)
# ---------------------------------------------------------------------------------------------------------
MOTION_NOISE3 = np.diag([0.1, 0.1, np.radians(5)])      # EKF-SLAM.py:12
OBS_NOISE3 = np.diag([0.5, 0.5])                        # EKF-SLAM.py:13


def predict3(state, cov, control, dt):
    """EKF-SLAM.py:29-56."""
    v, om = control
    th = state[2]
    if abs(om) > 1e-6:
        dx = -v / om * np.sin(th) + v / om * np.sin(th + om * dt)
        dy = v / om * np.cos(th) - v / om * np.cos(th + om * dt)
        dth = om * dt
    else:
        dx, dy, dth = v * np.cos(th) * dt, v * np.sin(th) * dt, 0.0
    out = np.array([state[0] + dx, state[1] + dy, state[2] + dth], dtype=np.float64)
    out[2] = np.arctan2(np.sin(out[2]), np.cos(out[2]))
    f = np.array([[1.0, 0.0, -v * dt * np.sin(th)], [0.0, 1.0, v * dt * np.cos(th)], [0.0, 0.0, 1.0]])
    return out, f @ cov @ f.T + MOTION_NOISE3


def update3(state, cov, observation, landmark_pos):
    """EKF-SLAM.py:59-84."""
    dx = landmark_pos[0] - state[0]
    dy = landmark_pos[1] - state[1]
    q = dx ** 2 + dy ** 2
    sq = np.sqrt(q)
    h = np.array([[-dx / sq, -dy / sq, 0.0], [dy / q, -dx / q, -1.0]])
    s = h @ cov @ h.T + OBS_NOISE3
    k = cov @ h.T @ np.linalg.inv(s)
    y = np.array([observation[0] - sq, observation[1] - (np.arctan2(dy, dx) - state[2])])
    y[1] = np.arctan2(np.sin(y[1]), np.cos(y[1]))
    return np.asarray(state, dtype=np.float64) + k @ y, (np.eye(3) - k @ h) @ cov


# ---------------------------------------------------------------------------------------------------------
# parity metrics
# ---------------------------------------------------------------------------------------------------------
def parity_errors(mu, p, mu_ref, p_ref):
    """(mu max-norm rel, P max-norm rel, P covariance-scaled) — the three numbers every parity test bounds."""
    mu, p, mu_ref, p_ref = (np.asarray(a, dtype=np.float64) for a in (mu, p, mu_ref, p_ref))
    e_mu = np.max(np.abs(mu - mu_ref)) / max(np.max(np.abs(mu_ref)), 1e-300)
    e_p = np.max(np.abs(p - p_ref)) / max(np.max(np.abs(p_ref)), 1e-300)
    d = np.sqrt(np.abs(np.diagonal(p_ref, axis1=-2, axis2=-1)))
    scale = d[..., :, None] * d[..., None, :]
    mask = scale > 0
    e_s = float(np.max(np.abs(p - p_ref)[mask] / scale[mask])) if mask.any() else 0.0
    return float(e_mu), float(e_p), e_s
