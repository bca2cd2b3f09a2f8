"""Filter-sharded Monte Carlo ensembles across GPUs (batched mode).

Filters never interact, so the batched mode shards by FILTER: rank r runs a contiguous range of the ensemble with
the common inputs replicated and there is NO collective in the step loop. The one optional exchange is at the end of
a replay: the ensemble mean and covariance of the state — what the reference's ablation studies show per run as
plots (README.md:421-492) — from per-rank sums that ``ekf_batch_moments`` forms on the device in a fixed order
(``BatchedEKF.moments()``: n + n^2 doubles) and one ``all_reduce`` of those sums plus the filter count.

``filter_partition`` and ``allreduce_moments`` / ``moments_to_mean_cov`` are device-agnostic (covered by the
world_size-2 ``gloo`` test on CPU); the sums themselves come from the CUDA library.
"""
from __future__ import annotations


def filter_partition(batch: int, world: int):
    """[(begin, end)] per rank: ``batch`` filters split into ``world`` contiguous, near-equal ranges (the first
    ``batch % world`` ranks hold one more). Ranges are empty when there are more ranks than filters."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if batch < 0:
        raise ValueError("batch must be >= 0")
    base, extra = divmod(batch, world)
    out, b = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append((b, b + cnt))
        b += cnt
    return out


def allreduce_moments(count: int, sums, group=None):
    """Sum the per-rank (filter count, moment sums) over the ranks; ``sums`` is the tensor of ``moments()`` (any
    device; NCCL for CUDA tensors, gloo for CPU). Returns (total count, summed tensor); without an initialised
    process group it is the identity (one rank)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return int(count), sums
    packed = torch.cat([sums.reshape(-1), torch.tensor([float(count)], dtype=sums.dtype, device=sums.device)])
    dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    return int(round(float(packed[-1]))), packed[:-1]


def moments_to_mean_cov(count: int, sums, n: int):
    """(mean (n,), covariance (n, n)) of the ensemble from its count and summed moments: E[x], E[x x^T] - E[x] E[x]^T
    (population form, as numpy's ``cov(..., bias=True)``)."""
    if count < 1:
        raise ValueError("empty ensemble")
    s = sums.reshape(-1)
    mean = s[:n] / count
    second = s[n:n + n * n].reshape(n, n) / count
    return mean, second - mean[:, None] * mean[None, :]


def ensemble_mean_cov(filt, group=None):
    """Ensemble mean / covariance of the state over ALL ranks' filters of a ``BatchedEKF`` (device tensors)."""
    count, sums = allreduce_moments(filt.batch, filt.moments(), group)
    return moments_to_mean_cov(count, sums, filt.n)
