"""TEST INFRASTRUCTURE — numpy twin of the device-side Monte Carlo perturbation generator
(slam-duckietown_b200/csrc/ekf_philox.cuh), bit for bit.

The batched Monte Carlo mode perturbs ONE common trajectory per ensemble member. Round 1 drew the perturbations on
the host (numpy) or with torch.randn and shipped them per filter per step (58.7 MB per step at configuration M).
Here every number is a pure function of (seed, global filter index, step, stream), so the device generates it in
place and this module reproduces it on the host for the per-member parity tests:

  * Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; the Random123 reference
    implementation's known-answer vectors are checked in tests/test_mc_noise.py): key = (seed lo, seed hi),
    counter = (filter lo, filter hi, step, stream) -> four 32-bit words;
  * two uniforms in (0, 1): u = (k + 0.5) 2^-52 with k = (w0 >> 6) 2^26 + (w1 >> 6) (and w2, w3) — exact in float64;
  * a PAIR of standard normals by Box-Muller, r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2), with the logarithm
    (frexp + atanh series) and the sine / cosine (octant reduction + Taylor polynomials on [0, pi/4]) built from IEEE
    add / multiply / divide / sqrt ONLY, each rounded separately and in one fixed order (the device uses __dadd_rn /
    __dmul_rn / __ddiv_rn / __dsqrt_rn: no FMA contraction), so host and device agree in every bit. Straight-line code:
    a warp never diverges (the first cut used the inverse CDF AS241, whose tail branch almost every warp took:
    89 us per 1 M filter-steps against 0.99 ms for the step itself, tools/mc_probe.py).
"""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10: counters / keys are uint32 arrays (broadcastable); returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint32) for x in (c0, c1, c2, c3))
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = np.asarray(k0, dtype=np.uint32)
    k1 = np.asarray(k1, dtype=np.uint32)
    with np.errstate(over="ignore"):
        for r in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & MASK32).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & MASK32).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = (k0 + W0).astype(np.uint32)
            k1 = (k1 + W1).astype(np.uint32)
    return c0, c1, c2, c3


def uniform52(w_hi, w_lo):
    """(k + 0.5) 2^-52, k = (w_hi >> 6) 2^26 + (w_lo >> 6): strictly inside (0, 1), exact in float64."""
    k = (w_hi.astype(np.uint64) >> np.uint64(6)) * np.uint64(1 << 26) + (w_lo.astype(np.uint64) >> np.uint64(6))
    return (k.astype(np.float64) + 0.5) * 2.0 ** -52


LN2 = 0.6931471805599453
SQRT_HALF = 0.7071067811865476
LOG_COEF = [1.0 / (2 * k + 1) for k in range(12)]       # atanh series: 1, 1/3, 1/5, ... (rounded once, as literals)


def det_log(x):
    """log(x) for x > 0 from frexp + the atanh series, IEEE add / mul / div only (~1e-16 relative)."""
    x = np.asarray(x, dtype=np.float64)
    m, e = np.frexp(x)                                   # x = m 2^e, m in [0.5, 1)
    small = m < SQRT_HALF
    m = np.where(small, m * 2.0, m)
    e = np.where(small, e - 1, e).astype(np.float64)
    s = (m - 1.0) / (m + 1.0)
    z = s * s
    p = np.full_like(z, LOG_COEF[11])
    for k in range(10, -1, -1):
        p = p * z + LOG_COEF[k]
    return e * LN2 + (2.0 * s) * p


PIO4 = 0.7853981633974483
SIN_COEF = [-1.0 / 6.0, 1.0 / 120.0, -1.0 / 5040.0, 1.0 / 362880.0, -1.0 / 39916800.0, 1.0 / 6227020800.0,
            -1.0 / 1307674368000.0, 1.0 / 355687428096000.0]
COS_COEF = [-1.0 / 2.0, 1.0 / 24.0, -1.0 / 720.0, 1.0 / 40320.0, -1.0 / 3628800.0, 1.0 / 479001600.0,
            -1.0 / 87178291200.0, 1.0 / 20922789888000.0]


def sincos_2pi(t):
    """(sin, cos) of 2 pi t for t in [0, 1): octant k = floor(8 t) (exact), Taylor polynomials on [0, pi/4] in a fixed
    IEEE add / mul order, octant symmetries by selection and sign."""
    t = np.asarray(t, dtype=np.float64)
    x = t * 8.0
    k = np.floor(x)
    f = x - k
    ki = k.astype(np.int64)
    g = np.where((ki & 1) == 1, 1.0 - f, f)
    a = g * PIO4
    a2 = a * a
    ps = np.full_like(a, SIN_COEF[7])
    pc = np.full_like(a, COS_COEF[7])
    for i in range(6, -1, -1):
        ps = ps * a2 + SIN_COEF[i]
        pc = pc * a2 + COS_COEF[i]
    sa = a + a * (a2 * ps)
    ca = 1.0 + a2 * pc
    swap = (((ki + 1) >> 1) & 1) == 1
    s = np.where(swap, ca, sa)
    c = np.where(swap, sa, ca)
    s = np.where(ki >= 4, -s, s)
    c = np.where((ki >= 2) & (ki <= 5), -c, c)
    return s, c


def box_muller(u1, u2):
    """Two independent standard normals from two uniforms in (0, 1): r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2)."""
    r = np.sqrt(-2.0 * det_log(u1))
    s, c = sincos_2pi(u2)
    return r * c, r * s


def normal_pair(seed, filt, step, stream):
    """The two standard normals of (seed, global filter index, step, stream); array arguments broadcast."""
    seed = int(seed)
    filt = np.asarray(filt, dtype=np.uint64)
    w = philox4x32_10((filt & MASK32).astype(np.uint32), (filt >> np.uint64(32)).astype(np.uint32),
                      np.asarray(step, dtype=np.uint32), np.asarray(stream, dtype=np.uint32),
                      np.uint32(seed & 0xFFFFFFFF), np.uint32((seed >> 32) & 0xFFFFFFFF))
    return box_muller(uniform52(w[0], w[1]), uniform52(w[2], w[3]))


# streams of one filter-step (ekf_philox.cuh: kStreamOdometry, kStreamObservation0, kStreamDetection0)
STREAM_ODOMETRY = 0
STREAM_OBSERVATION0 = 1
STREAM_DETECTION0 = 64


def perturb_step(seed, first_filter, batch, step, ang, lin, slots, rngs, brgs, sigma):
    """What ekf_batch_perturb writes for one step: (ang[B], lin[B], slot[B][m], rng[B][m], brg[B][m]).
    sigma = (sigma_ang, sigma_lin, sigma_rng, sigma_brg); empty slots (-1) stay unperturbed."""
    filt = np.arange(first_filter, first_filter + batch, dtype=np.uint64)
    z0, z1 = normal_pair(seed, filt, step, STREAM_ODOMETRY)
    a = ang + sigma[0] * z0
    l_ = lin + sigma[1] * z1
    m = len(slots)
    slot = np.tile(np.asarray(slots, dtype=np.int32)[None, :], (batch, 1))
    rr = np.zeros((batch, m))
    bb = np.zeros((batch, m))
    for k in range(m):
        z0, z1 = normal_pair(seed, filt, step, STREAM_OBSERVATION0 + k)
        rr[:, k] = rngs[k] + sigma[2] * z0
        bb[:, k] = brgs[k] + sigma[3] * z1
    return a, l_, slot, rr, bb


def detection_noise(seed, first_filter, batch, step, n_det, sigma):
    """What ekf_batch_detection_noise writes: (noise_tx[D][B], noise_tz[D][B]) = sigma * normal pairs of streams
    STREAM_DETECTION0 + d."""
    filt = np.arange(first_filter, first_filter + batch, dtype=np.uint64)
    tx = np.zeros((n_det, batch))
    tz = np.zeros((n_det, batch))
    for d in range(n_det):
        z0, z1 = normal_pair(seed, filt, step, STREAM_DETECTION0 + d)
        tx[d], tz[d] = sigma * z0, sigma * z1
    return tx, tz
