"""TEST INFRASTRUCTURE — numpy twin of the device-side Monte Carlo perturbation generator
(slam-duckietown_b200/csrc/ekf_philox.cuh), bit for bit.

The batched Monte Carlo mode perturbs ONE common trajectory per ensemble member. Round 1 drew the perturbations on
the host (numpy) or with torch.randn and shipped them per filter per step (58.7 MB per step at configuration M).
Here every number is a pure function of (seed, global filter index, step, stream), so the device generates it in
place and this module reproduces it on the host for the per-member parity tests:

  * Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; the Random123 reference
    implementation's known-answer vectors are checked in tests/test_mc_noise.py): key = (seed lo, seed hi),
    counter = (filter lo, filter hi, step, stream) -> four 32-bit words;
  * two uniforms strictly inside (0, 1) with 23 random bits each: u = (k + 0.5) 2^-23, k = w >> 9 (exact in float32);
    where the radial uniform draws k = 0 a second word refines it to (k' + 0.5) 2^-46 (tails to 8 sigma);
  * a PAIR of standard normals by Box-Muller, r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2), in SINGLE precision
    with the logarithm (frexp + atanh series) and the sine / cosine (octant reduction + Taylor polynomials on
    [0, pi/4]) built from IEEE add / multiply / divide / sqrt ONLY, each rounded separately and in one fixed order
    (the device uses __fadd_rn / __fmul_rn / __fdiv_rn / __fsqrt_rn: no FMA contraction; numpy float32 arithmetic
    rounds every operation the same way), so host and device agree in every bit; the pair is widened to float64
    (exact) before it is scaled. Straight-line code: a warp never diverges (the first cut used the inverse CDF AS241,
    whose tail branch almost every warp took). The first Box-Muller was in double precision with 52-bit uniforms
    (1e-16 accurate, ~270 instructions per draw, +12-15 % on the generated step); noise deviates do not need that.
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


F32 = np.float32


def _c(hexfloat):
    """A float32 constant from the hex literal the device code spells (ekf_philox.cuh)."""
    v = float.fromhex(hexfloat)
    assert float(F32(v)) == v
    return F32(v)


HALF, ONE, TWO, EIGHT = F32(0.5), F32(1.0), F32(2.0), F32(8.0)
P23, P46 = _c("0x1p-23"), _c("0x1p-46")


def uniform23(w):
    """(k + 0.5) 2^-23, k = w >> 9: strictly inside (0, 1), exact in float32."""
    k = (np.asarray(w, dtype=np.uint32) >> np.uint32(9)).astype(F32)
    return (k + HALF) * P23


def uniform23_tail(w, w2):
    """The radial uniform: uniform23(w), refined to (k' + 0.5) 2^-46 by a second word where k = 0."""
    k = np.asarray(w, dtype=np.uint32) >> np.uint32(9)
    fine = ((np.asarray(w2, dtype=np.uint32) >> np.uint32(9)).astype(F32) + HALF) * P46
    return np.where(k != 0, (k.astype(F32) + HALF) * P23, fine).astype(F32)


LN2 = _c("0x1.62e43p-1")
SQRT_HALF = _c("0x1.6a09e6p-1")
LOG_COEF = [ONE, _c("0x1.555556p-2"), _c("0x1.99999ap-3"), _c("0x1.24924ap-3"), _c("0x1.c71c72p-4")]  # 1/(2k+1)


def det_log(x):
    """log(x) for x > 0 from frexp + the atanh series, float32 add / mul / div only (~2e-7 relative)."""
    x = np.asarray(x, dtype=F32)
    m, e = np.frexp(x)                                   # x = m 2^e, m in [0.5, 1)
    small = m < SQRT_HALF
    m = np.where(small, m * TWO, m).astype(F32)
    e = np.where(small, e - 1, e).astype(F32)
    s = (m - ONE) / (m + ONE)
    z = s * s
    p = np.full_like(z, LOG_COEF[4])
    for k in range(3, -1, -1):
        p = p * z + LOG_COEF[k]
    out = e * LN2 + (TWO * s) * p
    assert out.dtype == F32
    return out


PIO4 = _c("0x1.921fb6p-1")
SIN_COEF = [-_c("0x1.555556p-3"), _c("0x1.111112p-7"), -_c("0x1.a01a02p-13"), _c("0x1.71de3ap-19")]   # -1/3! .. 1/9!
COS_COEF = [-_c("0x1p-1"), _c("0x1.555556p-5"), -_c("0x1.6c16c2p-10"), _c("0x1.a01a02p-16")]          # -1/2! .. 1/8!


def sincos_2pi(t):
    """(sin, cos) of 2 pi t for t in [0, 1): octant k = floor(8 t) (exact), Taylor polynomials on [0, pi/4] in a fixed
    float32 add / mul order, octant symmetries by selection and sign."""
    t = np.asarray(t, dtype=F32)
    x = t * EIGHT
    k = np.floor(x)
    f = x - k
    ki = k.astype(np.int64)
    g = np.where((ki & 1) == 1, ONE - f, f).astype(F32)
    a = g * PIO4
    a2 = a * a
    ps = np.full_like(a, SIN_COEF[3])
    pc = np.full_like(a, COS_COEF[3])
    for i in range(2, -1, -1):
        ps = ps * a2 + SIN_COEF[i]
        pc = pc * a2 + COS_COEF[i]
    sa = a + a * (a2 * ps)
    ca = ONE + a2 * pc
    swap = (((ki + 1) >> 1) & 1) == 1
    s = np.where(swap, ca, sa)
    c = np.where(swap, sa, ca)
    s = np.where(ki >= 4, -s, s)
    c = np.where((ki >= 2) & (ki <= 5), -c, c)
    assert s.dtype == F32 and c.dtype == F32
    return s, c


def box_muller(u1, u2):
    """Two independent standard normals (float32) from two uniforms in (0, 1): r = sqrt(-2 log u1),
    (r cos 2 pi u2, r sin 2 pi u2)."""
    r = np.sqrt(-TWO * det_log(u1))
    s, c = sincos_2pi(u2)
    z0, z1 = r * c, r * s
    assert z0.dtype == F32 and z1.dtype == F32
    return z0, z1


def normal_pair(seed, filt, step, stream):
    """The two standard normals of (seed, global filter index, step, stream), widened to float64 (exact); array
    arguments broadcast."""
    seed = int(seed)
    filt = np.asarray(filt, dtype=np.uint64)
    w = philox4x32_10((filt & MASK32).astype(np.uint32), (filt >> np.uint64(32)).astype(np.uint32),
                      np.asarray(step, dtype=np.uint32), np.asarray(stream, dtype=np.uint32),
                      np.uint32(seed & 0xFFFFFFFF), np.uint32((seed >> 32) & 0xFFFFFFFF))
    z0, z1 = box_muller(uniform23_tail(w[0], w[1]), uniform23(w[2]))
    return z0.astype(np.float64), z1.astype(np.float64)


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
