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
  * standard normals by the inverse CDF, Wichura's AS241 PPND16 (Appl. Statist. 37 (1988) 477-484), evaluated
    with IEEE add / multiply / divide / sqrt ONLY, each rounded separately and in one fixed order (the device uses
    __dadd_rn / __dmul_rn / __ddiv_rn / __dsqrt_rn: no FMA contraction), and a logarithm built from the same
    operations (frexp + atanh series), so host and device agree in every bit.
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


A = [3.3871328727963666080, 1.3314166789178437745e+2, 1.9715909503065514427e+3, 1.3731693765509461125e+4,
     4.5921953931549871457e+4, 6.7265770927008700853e+4, 3.3430575583588128105e+4, 2.5090809287301226727e+3]
B = [1.0, 4.2313330701600911252e+1, 6.8718700749205790830e+2, 5.3941960214247511077e+3, 2.1213794301586595867e+4,
     3.9307895800092710610e+4, 2.8729085735721942674e+4, 5.2264952788528545610e+3]
C = [1.42343711074968357734, 4.63033784615654529590, 5.76949722146069140550, 3.64784832476320460504,
     1.27045825245236838258, 2.41780725177450611770e-1, 2.27238449892691845833e-2, 7.74545014278341407640e-4]
D = [1.0, 2.05319162663775882187, 1.67638483018380384940, 6.89767334985100004550e-1, 1.48103976427480074590e-1,
     1.51986665636164571966e-2, 5.47593808499534494600e-4, 1.05075007164441684324e-9]
E = [6.65790464350110377720, 5.46378491116411436990, 1.78482653991729133580, 2.96560571828504891230e-1,
     2.65321895265761230930e-2, 1.24266094738807843860e-3, 2.71155556874348757815e-5, 2.01033439929228813265e-7]
F = [1.0, 5.99832206555887937690e-1, 1.36929880922735805310e-1, 1.48753612908506148525e-2,
     7.86869131145613259100e-4, 1.84631831751005468180e-5, 1.42151175831644588870e-7, 2.04426310338993978564e-15]


def _horner(coef, r):
    p = np.full_like(r, coef[7])
    for k in range(6, -1, -1):
        p = p * r + coef[k]
    return p


def normal_icdf(u):
    """Standard-normal quantile of u in (0, 1): AS241 PPND16, fixed operation order (see module docstring)."""
    u = np.asarray(u, dtype=np.float64)
    q = u - 0.5
    out = np.empty_like(u)
    central = np.abs(q) <= 0.425
    r = 0.180625 - q * q
    with np.errstate(invalid="ignore", divide="ignore"):
        out_c = q * _horner(A, r) / _horner(B, r)
        rt = np.where(q < 0.0, u, 1.0 - u)
        rt = np.sqrt(-det_log(np.where(central, 0.25, rt)))
        r1 = rt - 1.6
        v1 = _horner(C, r1) / _horner(D, r1)
        r2 = rt - 5.0
        v2 = _horner(E, r2) / _horner(F, r2)
    tail = np.where(rt <= 5.0, v1, v2)
    tail = np.where(q < 0.0, -tail, tail)
    out = np.where(central, out_c, tail)
    return out


def normal_pair(seed, filt, step, stream):
    """The two standard normals of (seed, global filter index, step, stream); array arguments broadcast."""
    seed = int(seed)
    filt = np.asarray(filt, dtype=np.uint64)
    w = philox4x32_10((filt & MASK32).astype(np.uint32), (filt >> np.uint64(32)).astype(np.uint32),
                      np.asarray(step, dtype=np.uint32), np.asarray(stream, dtype=np.uint32),
                      np.uint32(seed & 0xFFFFFFFF), np.uint32((seed >> 32) & 0xFFFFFFFF))
    return normal_icdf(uniform52(w[0], w[1])), normal_icdf(uniform52(w[2], w[3]))


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
