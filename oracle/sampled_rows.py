"""TEST INFRASTRUCTURE — the rank-2 restatement of the reference update / predict evaluated on a SUBSET of the
rows of P (leg (iii) of SURVEY.md §7 "Oracle cost at scale").

At n = 65 539 the full covariance is 34 GB: neither the verbatim reference nor the full rank-2 restatement can run
beside it. But the reference's update (replay_no_ros.py:436-480) in rank form touches row i of P only through

    P[i, :] -= K[i, :] @ (H5 @ P[idx5, :]),      K[i, :] = (P[i, idx5] @ H5^T) @ S^-1,
    S = H5 @ P[idx5, idx5] @ H5^T + Q,            idx5 = {0, 1, 2, j, j+1}

so a set T of rows can be carried forward EXACTLY as the full oracle would carry them, provided T contains the
rows {0, 1, 2} and {j_u, j_u + 1} of every update u that will be applied (they are the "5 gathered rows"; the 5
gathered columns of a tracked row are entries of that row). The predict (:420-430) is row-local too:
rows 0-1 += g * row 2, then every row's columns 0-1 += g * its column 2, then three diagonal adds.

`TrackedRows` holds mu (full, n doubles) and P[T, :]; its per-row arithmetic is `ekf_oracle.update_rank` /
`predict_cov_rank` restricted to T (checked against them in tests/test_oracle_golden.py).
"""
from __future__ import annotations

import numpy as np

from . import ekf_oracle as eo


def required_rows(slots) -> list:
    """Rows every update of the sequence gathers: {0,1,2} U {3+2s, 4+2s}."""
    out = [0, 1, 2]
    for s in slots:
        for r in (3 + 2 * int(s), 4 + 2 * int(s)):
            if r not in out:
                out.append(r)
    return out


class TrackedRows:
    def __init__(self, mu, rows, p_rows):
        """mu (n,), rows: global row indices (must cover `required_rows` of whatever will be applied),
        p_rows (len(rows), n): those rows of the initial covariance."""
        self.mu = np.array(mu, dtype=np.float64, copy=True)
        self.rows = [int(r) for r in rows]
        self.pos = {r: i for i, r in enumerate(self.rows)}
        self.P = np.array(p_rows, dtype=np.float64, copy=True)
        assert self.P.shape == (len(self.rows), len(self.mu))
        assert all(r in self.pos for r in (0, 1, 2)), "rows 0-2 must be tracked"

    def update(self, slot, rng, brg, mm):
        """replay_no_ros.py:436-480 (plain form) on the tracked rows."""
        j = 3 + 2 * int(slot)
        idx = [0, 1, 2, j, j + 1]
        if not all(r in self.pos for r in idx):
            raise KeyError(f"rows {j}, {j + 1} of slot {slot} are not tracked")
        y, h5 = eo.innovation(self.mu, j, rng, brg)
        five = self.P[[self.pos[r] for r in idx], :]               # the 5 gathered ROWS of the current P
        hp = h5 @ five                                             # 2 x n
        q2 = np.array([[mm * mm, 0.0], [0.0, mm * mm]])
        s = hp[:, idx] @ h5.T + q2
        pht = self.P[:, idx] @ h5.T                                # tracked rows' 5 gathered COLUMNS
        k = pht @ eo._inv2(s)                                      # K[T, :]
        self.mu[self.rows] += k @ y                                # only the tracked entries of mu are maintained
        self.P -= k @ hp

    def predict(self, ang, lin, flags: eo.Flags):
        dx, dy, th, g0, g1 = eo.motion_step(self.mu[2], ang, lin, flags)
        self.mu[0] += dx
        self.mu[1] += dy
        self.mu[2] = th
        p0, p1, p2 = self.pos[0], self.pos[1], self.pos[2]
        self.P[p0, :] += g0 * self.P[p2, :]
        self.P[p1, :] += g1 * self.P[p2, :]
        self.P[:, 0] += g0 * self.P[:, 2]
        self.P[:, 1] += g1 * self.P[:, 2]
        mv = flags.motion_var
        self.P[p0, 0] += mv * mv
        self.P[p1, 1] += mv * mv
        self.P[p2, 2] += (mv / 2) * (mv / 2)

    def errors(self, mu_rows, p_rows, diag, subset=None):
        """(mu max-norm rel over the compared entries, P max-norm rel, P covariance-scaled) of device values
        `mu_rows` (k,), `p_rows` (k, n) against this oracle's rows `subset` (default: all tracked rows, in order);
        `diag` (n,) = diagonal of the device P (the covariance scale of the columns; the compared rows' own diagonal
        entries come from the oracle)."""
        rows = self.rows if subset is None else [int(r) for r in subset]
        sel = [self.pos[r] for r in rows]
        p_ref = self.P[sel]
        mu_ref = self.mu[rows]
        e_mu = float(np.max(np.abs(mu_rows - mu_ref)) / max(np.max(np.abs(self.mu)), 1e-300))
        e_p = float(np.max(np.abs(p_rows - p_ref)) / max(np.max(np.abs(self.P)), 1e-300))
        d_rows = np.sqrt(np.abs(np.array([p_ref[i, r] for i, r in enumerate(rows)])))
        d_cols = np.sqrt(np.abs(np.asarray(diag, dtype=np.float64)))
        scale = d_rows[:, None] * d_cols[None, :]
        mask = scale > 0
        e_s = float(np.max(np.abs(p_rows - p_ref)[mask] / scale[mask])) if mask.any() else 0.0
        return e_mu, e_p, e_s
