"""Row-block sharding of ONE large-map EKF-SLAM filter across GPUs.

GPU g owns a contiguous block of rows of P whose boundaries fall on landmark pairs (rows j, j+1 share an
owner); rank 0 also owns the robot rows 0-2. Per range-bearing update there is exactly one exchange: the five
rows {0, 1, 2, j, j+1} of P (with their mu entries riding at column ``ld`` of each exported row) travel from
their owners to every rank — a broadcast of three rows from rank 0 and of two rows from owner(j), 40*n bytes —
after which every rank runs the gather + streaming rank-2 update on its own block
(``ekf_lm_export_rows`` / ``ekf_lm_update_from_rows``). ``K`` needs only columns of local rows; predict is
communication-free (reference semantics: replay_no_ros.py:420-430 touches rows/columns 0-1 only).

It is ROWS that travel: the reference keeps full, slightly asymmetric storage (replay_no_ros.py:480), where
using (P H^T)^T for H P is numerically unstable.

The partition helpers and ``exchange_rows`` are device-agnostic (they are covered by world_size-2 ``gloo`` tests
on CPU); everything that computes goes through the CUDA library and fails loudly without it.
"""
from __future__ import annotations

import os

from . import ekf_bindings as eb


def _landmark_counts(n_landmarks: int, world: int, rank0_fewer: int = 0):
    """Landmarks per rank: near-equal contiguous ranges; with ``rank0_fewer`` > 0 rank 0 takes that many landmarks
    less than the equal share (never fewer than one) and the others split the rest evenly."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if rank0_fewer < 0:
        raise ValueError("rank0_fewer must be >= 0")
    if world == 1 or rank0_fewer == 0:
        base, extra = divmod(n_landmarks, world)
        return [base + (1 if r < extra else 0) for r in range(world)]
    first = min(n_landmarks, max(1, n_landmarks // world - rank0_fewer))
    base, extra = divmod(n_landmarks - first, world - 1)
    return [first] + [base + (1 if r < extra else 0) for r in range(world - 1)]


def row_partition(n_landmarks: int, world: int, rank0_fewer: int = 0):
    """[(row_begin, row_end)] per rank: landmarks split into ``world`` contiguous ranges (``_landmark_counts``);
    rank 0 additionally owns rows 0-2. Every boundary is at a row 3 + 2k."""
    out, lm = [], 0
    for r, cnt in enumerate(_landmark_counts(n_landmarks, world, rank0_fewer)):
        begin = 0 if r == 0 else 3 + 2 * lm
        lm += cnt
        out.append((begin, 3 + 2 * lm))
    return out


def owner_of_slot(slot: int, n_landmarks: int, world: int, rank0_fewer: int = 0) -> int:
    """Rank that owns rows 3+2*slot and 4+2*slot under ``row_partition``."""
    if not 0 <= slot < n_landmarks:
        raise ValueError(f"slot {slot} outside [0, {n_landmarks})")
    if world > 1 and rank0_fewer > 0:
        counts = _landmark_counts(n_landmarks, world, rank0_fewer)
        if slot < counts[0]:
            return 0
        base, extra = divmod(n_landmarks - counts[0], world - 1)
        rest, split = slot - counts[0], extra * (base + 1)
        return 1 + (rest // (base + 1) if rest < split else extra + (rest - split) // base)
    base, extra = divmod(n_landmarks, world)
    split = extra * (base + 1)
    if slot < split:
        return slot // (base + 1)
    return extra + (slot - split) // base


def rank0_relief(n_landmarks: int, world: int) -> int:
    """Landmarks rank 0 gives up under the peer-memory exchange. Besides its rows, rank 0 stores the three robot rows
    of every update into all world - 1 peers (48 n (world - 1) bytes on the wire); next to an HBM-streaming pass that
    costs the time of ~33 rows of its block per peer, whatever n is (32 768 landmarks on 8 GPUs: rank 0's streaming
    kernel 1314 us against 1278 us on the others with equal shares; 1282 / 1285 us with the relief, 1327 -> 1318 us per
    update; profiles/r2v_p2p_timeline_relief_spread_8gpu.txt). Equal finish times: rows_0 + 33 (world - 1) = rows_other.
    Only for row blocks that stream from HBM: at 4096 landmarks on 8 GPUs (67 MB per rank, L2-resident) the update is
    bound by the arrival of the rows, not by rank 0's pass, and the relief measured 3 % SLOWER (same file)."""
    if world < 2:
        return 0
    n = 3 + 2 * n_landmarks
    if 8.0 * n * n / world <= 160e6:      # the threshold of the 16-row-tile selection (ekf_lm_bind_state)
        return 0
    fewer = round(33 * (world - 1) * (world - 1) / (2.0 * world))
    return int(min(fewer, max(0, n_landmarks // world // 2)))


def exchange_rows(rows5, owner: int, group=None):
    """The per-update collective: rows 0-2 from rank 0, rows 3-4 (= P rows j, j+1) from ``owner``.
    ``rows5`` is a (5, ld+4) tensor (CUDA -> NCCL over NVLink; CPU -> gloo in the tests)."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.broadcast(rows5[0:3], src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    dist.broadcast(rows5[3:5], src=dist.get_global_rank(group, owner) if group is not None else owner,
                   group=group)


def exchange_rows_multi(rows, owners, group=None):
    """Exchange of a fused step: rows 0-2 from rank 0, rows 3+2u, 4+2u (= P rows j_u, j_u+1) from owners[u]."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    g = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    dist.broadcast(rows[0:3], src=g(0), group=group)
    for u, owner in enumerate(owners):
        dist.broadcast(rows[3 + 2 * u:5 + 2 * u], src=g(owner), group=group)


class ShardedLargeMapEKF:
    """One process per GPU; every rank calls the same methods with the same arguments (SPMD)."""

    def __init__(self, capacity: int, rank: int, world: int, device: int | None = None, group=None,
                 exchange: str = "nccl", rank0_fewer: int | None = None):
        """exchange: "nccl" — export + two broadcasts per update through torch.distributed; "p2p" — the owners push
        their rows into every rank's exchange area over NVLink peer memory (cudaIpc) from INSIDE the update's own two
        kernels (gather: rows j, j+1; streaming: rows 0-2 for the next update), no collective and no extra launch."""
        self.capacity, self.rank, self.world, self.group = int(capacity), int(rank), int(world), group
        self.exchange = exchange
        # rank 0 also feeds the robot rows to every peer: under the peer-memory exchange it owns fewer landmarks
        # (rank0_relief; EKF_LM_RANK0_FEWER or the argument override it, 0 = equal shares)
        if rank0_fewer is None:
            env = os.environ.get("EKF_LM_RANK0_FEWER")
            rank0_fewer = int(env) if env not in (None, "") else \
                (rank0_relief(self.capacity, self.world) if exchange == "p2p" else 0)
        self.rank0_fewer = int(rank0_fewer)
        self.ranges = row_partition(self.capacity, self.world, self.rank0_fewer)
        if exchange == "p2p" and any(hi <= lo for lo, hi in self.ranges):
            # a rank without rows launches no streaming kernel, so it would never acknowledge an exchange buffer
            raise ValueError(f"peer-memory exchange needs at least one landmark per rank "
                             f"({self.capacity} landmarks on {self.world} ranks)")
        self.local = eb.LargeMapEKF(self.capacity, device=self.rank if device is None else device,
                                    row_range=self.ranges[self.rank])
        self.n, self.ld = self.local.n, self.local.ld
        torch = self.local.torch
        self.rows5 = torch.zeros((5, self.ld + 4), dtype=torch.float64, device=self.local.mu.device)
        self.local.rows5 = self.rows5
        if exchange == "p2p":
            self._connect_p2p()
        elif exchange != "nccl":
            raise ValueError(f"unknown exchange {exchange!r}")

    def _connect_p2p(self):
        """torch.distributed is the plumbing only: one all-gather of the 64-byte IPC handles at set-up."""
        import torch.distributed as dist
        handle = self.local.p2p_init(self.world, self.rank)
        if self.world == 1:
            self.local.p2p_connect(handles=[handle])
            return
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handle, group=self.group)
        self.local.p2p_connect(handles=gathered)
        dist.barrier(group=self.group)

    def close(self):
        self.local.close()

    def owner(self, slot: int) -> int:
        """Rank that owns the rows of landmark ``slot`` under this job's partition."""
        return owner_of_slot(slot, self.capacity, self.world, self.rank0_fewer)

    def set_noise(self, motion_var: float, meas_var: float):
        self.local.set_noise(motion_var, meas_var)

    def set_flags(self, circular=True, measurement=True, disable_motion=False, joseph=False):
        """joseph=True (one-pass Joseph form) needs K at every column on every rank: the peer-memory exchange
        all-gathers it inside the gather kernel; the broadcast exchange does not carry it."""
        if joseph and self.exchange != "p2p" and self.world > 1:
            raise ValueError("the Joseph form on row-block shards needs exchange='p2p'")
        self.local.set_flags(circular, measurement, disable_motion, joseph=joseph)

    def load(self, mean, cov_rows):
        """``mean`` (m,) replicated; ``cov_rows``: the FULL (m, m) covariance or a callable
        (row_begin, row_end) -> (rows, m) block, so that a 34 GB matrix never has to exist on one host."""
        import numpy as np
        lo, hi = self.ranges[self.rank]
        m = len(mean)
        hi = min(hi, m)
        blk = cov_rows(lo, hi) if callable(cov_rows) else np.asarray(cov_rows)[lo:hi, :]
        torch = self.local.torch
        mu = np.zeros(self.ld)
        mu[:m] = mean
        self.local.mu.copy_(torch.from_numpy(mu))
        self.local.P.zero_()
        if hi > lo:
            self.local.P[: hi - lo, :m] = torch.from_numpy(np.ascontiguousarray(blk)).to(self.local.P.device)
        self.local.p2p_invalidate()

    def augment(self, slot: int, gx: float, gy: float):
        self.local.augment(slot, gx, gy)       # mu entries everywhere, diagonal on the owner only

    def predict(self, ang: float, lin: float):
        self.local.predict(ang, lin)           # no communication

    def update(self, slot: int, rng: float, brg: float):
        if self.exchange == "p2p":
            self.local.p2p_update(slot, rng, brg, self.owner(slot))
            return
        self.local.export_rows(slot)
        exchange_rows(self.rows5, self.owner(slot), self.group)
        self.local.update_from_rows(slot, rng, brg, self.rows5)

    def update_multi(self, slots, rngs, brgs):
        """The step's sequential updates fused into ONE pass over every shard's row block and ONE exchange of the
        rows {0,1,2} U {j_u, j_u+1} (groups of at most 4 updates; a single update takes the plain path)."""
        slots, rngs, brgs = list(slots), list(rngs), list(brgs)
        for first in range(0, len(slots), 4):
            sl, rg, bg = slots[first:first + 4], rngs[first:first + 4], brgs[first:first + 4]
            if len(sl) == 1:
                self.update(sl[0], rg[0], bg[0])
                continue
            if self.exchange == "p2p":
                self.local.p2p_update_multi(sl, rg, bg, [self.owner(s) for s in sl])
                continue
            rows = self.local.export_rows_multi(sl)
            exchange_rows_multi(rows, [self.owner(s) for s in sl], self.group)
            self.local.update_multi_from_rows(sl, rg, bg, rows)

    def local_state(self):
        """(mu slice owned by this rank incl. the replicated pose on rank 0, P row block) as numpy."""
        lo, hi = self.ranges[self.rank]
        mu, blk = self.local.state()
        return mu[lo:hi], blk


class EmulatedShards:
    """All shards of a ``ShardedLargeMapEKF`` job inside ONE process on ONE GPU (the exchange is a device copy
    instead of a broadcast). Same kernels, same entry points, same per-element operation order — used by the
    single-GPU tests to check sharded == unsharded bitwise without needing a multi-GPU box."""

    def __init__(self, capacity: int, world: int, device: int = 0):
        self.capacity, self.world = int(capacity), int(world)
        self.ranges = row_partition(self.capacity, self.world)
        self.shards = [eb.LargeMapEKF(self.capacity, device=device, row_range=r) for r in self.ranges]
        self.n, self.ld = self.shards[0].n, self.shards[0].ld

    def close(self):
        for s in self.shards:
            s.close()

    def set_noise(self, mv, mm):
        for s in self.shards:
            s.set_noise(mv, mm)

    def load(self, mean, cov):
        for s in self.shards:
            s.load(mean, cov)

    def augment(self, slot, gx, gy):
        for s in self.shards:
            s.augment(slot, gx, gy)

    def predict(self, ang, lin):
        for s in self.shards:
            s.predict(ang, lin)

    def update(self, slot, rng, brg):
        owner = owner_of_slot(slot, self.capacity, self.world)
        bufs = [s.export_rows(slot) for s in self.shards]
        for r, s in enumerate(self.shards):           # "broadcast": copy the owners' rows to everyone
            if r != 0:
                bufs[r][0:3].copy_(bufs[0][0:3])
            if r != owner:
                bufs[r][3:5].copy_(bufs[owner][3:5])
        for s in self.shards:
            s.update_from_rows(slot, rng, brg)

    def update_multi(self, slots, rngs, brgs):
        slots = list(slots)
        owners = [owner_of_slot(s, self.capacity, self.world) for s in slots]
        bufs = [s.export_rows_multi(slots) for s in self.shards]
        for r in range(self.world):                   # "broadcast": copy the owners' rows to everyone
            if r != 0:
                bufs[r][0:3].copy_(bufs[0][0:3])
            for u, o in enumerate(owners):
                if r != o:
                    bufs[r][3 + 2 * u:5 + 2 * u].copy_(bufs[o][3 + 2 * u:5 + 2 * u])
        for s, b in zip(self.shards, bufs):
            s.update_multi_from_rows(slots, rngs, brgs, b)

    def state(self):
        import numpy as np
        mus, blks = [], []
        for (lo, hi), s in zip(self.ranges, self.shards):
            mu, blk = s.state()
            mus.append(mu[lo:hi])
            blks.append(blk)
        return np.concatenate(mus), np.concatenate(blks, axis=0)
