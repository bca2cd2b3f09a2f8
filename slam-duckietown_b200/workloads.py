"""Seeded synthetic worlds, trajectories and observations for the configurations of BASELINE.json
(SURVEY.md §8d). Host-side numpy only; inputs are generated ONCE on the host and fed to both the oracle and
the GPU path so association / branch decisions see identical bits.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class BatchedWorkload:
    """Monte Carlo replay inputs: T steps, B filters, up to M observations per filter-step."""
    n_landmarks: int
    landmarks: np.ndarray      # (L, 2) world
    truth: np.ndarray          # (T+1, 3) common true pose
    ang: np.ndarray            # (T, B) per-filter noisy angular displacement
    lin: np.ndarray            # (T, B)
    slot: np.ndarray           # (T, B, M) int32, -1 = no observation
    rng: np.ndarray            # (T, B, M)
    brg: np.ndarray            # (T, B, M)
    motion_var: np.ndarray     # (B,)
    meas_var: np.ndarray       # (B,)
    warm_steps: int            # first steps that initialise landmarks in slot order

    @property
    def steps(self):
        return self.ang.shape[0]

    @property
    def batch(self):
        return self.ang.shape[1]

    @property
    def m(self):
        return self.slot.shape[2]


def _wrap(a):
    return (a + np.pi) % (2 * np.pi) - np.pi


COMPACT_WORLD = (-0.55, 0.55, 0.05, 1.15)   # every landmark stays inside the reference's 1.5 m gate (:289)


def make_batched(batch: int, n_landmarks: int = 20, m: int = 3, steps: int = 16, seed: int = 1,
                 perturb_seed: int = 2, world=COMPACT_WORLD) -> BatchedWorkload:
    """Config B / M generator: landmarks uniform in `world` (default: a compact box around the
    robot's ~0.6 m turning circle so that the reference's 1.5 m gate passes every observation and the very
    same inputs can be replayed through the reference; SURVEY's [-1,3]x[-2,2] is `world=(-1,3,-2,2)`);
    common trajectory with
    ang ~ N(0.08, 0.05), lin ~ |N(0.05, 0.02)| per 0.7 s window; per filter motion_var ~ logU[0.05, 1.0],
    meas_var ~ logU[0.05, 1.5]; odometry + N(0, 0.01 rad / 3 mm); range/bearing + N(0, 0.02).

    The first ceil(L/m) steps observe slots in order (0..m-1, m..2m-1, ...) so that every landmark is
    initialised by augmentation (first-seen slot order, replay_no_ros.py:294-295); afterwards each step
    observes the m landmarks nearest to the true pose, nearest first.
    """
    wr = np.random.default_rng(seed)
    L = n_landmarks
    lms = np.stack([wr.uniform(world[0], world[1], L), wr.uniform(world[2], world[3], L)], axis=1)
    warm = -(-L // m)
    T = steps
    a_true = wr.normal(0.08, 0.05, T)
    l_true = np.abs(wr.normal(0.05, 0.02, T))
    truth = np.zeros((T + 1, 3))
    for t in range(T):
        x, y, th = truth[t]
        if abs(a_true[t]) <= 1e-2:
            truth[t + 1] = (x + l_true[t] * np.cos(th), y + l_true[t] * np.sin(th), th)
        else:
            r = l_true[t] / a_true[t]
            truth[t + 1] = (x - r * np.sin(th) + r * np.sin(th + a_true[t]),
                            y + r * np.cos(th) - r * np.cos(th + a_true[t]), _wrap(th + a_true[t]))
    pr = np.random.default_rng(perturb_seed)
    B = batch
    motion_var = np.exp(pr.uniform(np.log(0.05), np.log(1.0), B))
    meas_var = np.exp(pr.uniform(np.log(0.05), np.log(1.5), B))
    ang = a_true[:, None] + pr.normal(0.0, 0.01, (T, B))
    lin = l_true[:, None] + pr.normal(0.0, 0.003, (T, B))
    slot = np.full((T, B, m), -1, dtype=np.int32)
    rng = np.zeros((T, B, m))
    brg = np.zeros((T, B, m))
    for t in range(T):
        # detections of window t are taken along the way and averaged; model them at the pre-step pose
        x, y, th = truth[t]
        d = lms - np.array([x, y])
        dist = np.hypot(d[:, 0], d[:, 1])
        if t < warm:
            ids = np.arange(t * m, min((t + 1) * m, L))
        else:
            ids = np.argsort(dist, kind="stable")[:m]
        k = len(ids)
        slot[t, :, :k] = ids[None, :].astype(np.int32)
        r_true = dist[ids]
        b_true = _wrap(np.arctan2(d[ids, 1], d[ids, 0]) - th)
        rng[t, :, :k] = r_true[None, :] + pr.normal(0.0, 0.02, (B, k))
        brg[t, :, :k] = b_true[None, :] + pr.normal(0.0, 0.02, (B, k))
    return BatchedWorkload(L, lms, truth, ang, lin, slot, rng, brg, motion_var, meas_var, warm)


@dataclass
class LargeMapWorkload:
    n_landmarks: int
    landmarks: np.ndarray      # (L, 2)
    mu0: np.ndarray            # (n,)
    factor: np.ndarray         # (n, 64): P0 = factor @ factor.T / 64 + 0.1 I
    ops: list                  # sequence of ("predict", ang, lin) / ("update", slot, range, bearing)

    @property
    def n(self):
        return 3 + 2 * self.n_landmarks

    def p0(self) -> np.ndarray:
        p = self.factor @ self.factor.T
        p /= 64.0
        p[np.diag_indices(self.n)] += 0.1
        return p


def make_large_map(n_landmarks: int = 4096, updates: int = 200, grid=None, seed: int = 3, p_seed: int = 4,
                   rank: int = 64) -> LargeMapWorkload:
    """Config L / XL generator: jittered grid at 0.5 m, mu = truth + N(0, 0.05), dense SPD
    P0 = A A^T/64 + 0.1 I (every entry non-zero); `updates` single-landmark updates (landmark uniform among
    those within 1.5 m of the true pose) with an odometry predict after every 3rd update."""
    L = n_landmarks
    if grid is None:
        gx = int(np.ceil(np.sqrt(L)))
        gy = -(-L // gx)
    else:
        gx, gy = grid
    wr = np.random.default_rng(seed)
    ii, jj = np.meshgrid(np.arange(gx), np.arange(gy), indexing="ij")
    pts = np.stack([ii.ravel(), jj.ravel()], axis=1)[:L].astype(np.float64) * 0.5
    pts += wr.uniform(-0.1, 0.1, pts.shape)
    pose = np.array([0.5 * gx / 2, 0.5 * gy / 2, 0.3])
    truth = np.concatenate([pose, pts.ravel()])
    mu0 = truth + wr.normal(0.0, 0.05, truth.shape)
    a = np.random.default_rng(p_seed).normal(0.0, 1.0, (3 + 2 * L, rank))
    ops = []
    for u in range(updates):
        d = pts - pose[:2]
        dist = np.hypot(d[:, 0], d[:, 1])
        near = np.flatnonzero(dist <= 1.4)   # + noise stays inside the 1.5 m gate
        s = int(near[wr.integers(len(near))])
        r = dist[s] + wr.normal(0.0, 0.02)
        b = _wrap(np.arctan2(d[s, 1], d[s, 0]) - pose[2]) + wr.normal(0.0, 0.02)
        ops.append(("update", s, float(r), float(b)))
        if u % 3 == 2:
            ang, lin = float(wr.normal(0.08, 0.05)), float(abs(wr.normal(0.05, 0.02)))
            ops.append(("predict", ang, lin))
            if abs(ang) <= 1e-2:
                pose = pose + np.array([lin * np.cos(pose[2]), lin * np.sin(pose[2]), 0.0])
            else:
                rr = lin / ang
                pose = np.array([pose[0] - rr * np.sin(pose[2]) + rr * np.sin(pose[2] + ang),
                                 pose[1] + rr * np.cos(pose[2]) - rr * np.cos(pose[2] + ang),
                                 _wrap(pose[2] + ang)])
    return LargeMapWorkload(L, pts, mu0, a, ops)
