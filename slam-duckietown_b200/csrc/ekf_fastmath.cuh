// ekf_fastmath.cuh — branch-free fp64 reciprocal / division / reciprocal square root / atan2 / sincos for the motion
// model (a6, src/replay_no_ros.py:368-417), the innovation (a8, :443-469) and the 2x2 inverse (a9, :473).
//
// Why: the filter-step is one long dependent chain per filter (motion -> [innovation -> S^-1 -> K -> mu] x m) and
// throughput = resident filters / chain latency. CUDA's fp64 sqrt, division and atan2 each end in a predicated call
// to a slow path, which splits the innovation into basic blocks that cannot be interleaved, and the library atan2
// takes data-dependent branches (octant fix-ups) that DIVERGE when several filters share a warp
// (ekf_batched_group.cu: 170 warp-instructions per call against 100 for one filter per warp). The versions here are
// straight-line code: one MUFU seed + Newton steps, selects instead of branches, an Estrin-evaluated polynomial, so
// that sqrt, 1/q and atan2 of one innovation overlap.
//
// Accuracy (tests/test_fastmath.py, same source compiled for the host with an emulated 24-bit seed, 4 M points
// against long double libm): atan2 <= 2.5 ulp, 1/sqrt <= 2 ulp, sin / cos <= 1.5 ulp, division / reciprocal / sqrt
// <= 1 ulp. The reference's numpy calls are
// correctly rounded (sqrt, /) or < 1 ulp (arctan2); the parity bar for floating point is 1e-9 relative (SURVEY.md
// §8c) and the golden-vector distance did not move. Domain: finite operands, results in the normal
// range; a zero denominator gives NaN (the reference divides by zero there too: q = 0 at :461-469).
// Every kernel of the library goes through the same functions, so all variants still agree bitwise.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define EKF_FM __host__ __device__ __forceinline__
#else
#define EKF_FM static inline
#endif

namespace ekf {

// ~2^-20-accurate seeds: MUFU.RCP64H / MUFU.RSQ64H on the device, a float-precision stand-in on the host (tests)
EKF_FM double fm_rcp_seed(double d) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  return r;
#else
  return (double)(1.0f / (float)d);
#endif
}
EKF_FM double fm_rsqrt_seed(double d) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  return r;
#else
  return (double)(1.0f / sqrtf((float)d));
#endif
}

// 1 / d: seed, one cubic and one quadratic Newton step
EKF_FM double fm_rcp(double d) {
  double r = fm_rcp_seed(d);
  double e = fma(-d, r, 1.0);
  e = fma(e, e, e);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  return fma(r, e, r);
}

// n / d with one residual correction
EKF_FM double fm_div(double n, double d) {
  const double r = fm_rcp(d);
  const double t = n * r;
  const double rem = fma(-t, d, n);
  return fma(rem, r, t);
}

// sqrt(q) and 1 / sqrt(q) together (coupled Newton iteration on g ~ sqrt(q), h ~ 1 / (2 sqrt(q)))
EKF_FM void fm_sqrt_rsqrt(double q, double* root, double* inv_root) {
  const double y = fm_rsqrt_seed(q);
  double g = q * y, h = 0.5 * y;
  double r = fma(-h, g, 0.5);
  g = fma(g, r, g);
  h = fma(h, r, h);
  r = fma(-h, g, 0.5);
  g = fma(g, r, g);
  h = fma(h, r, h);
  const double d = fma(-g, g, q);
  g = fma(d, h, g);
  *root = g;
  *inv_root = h + h;
}

// atan2(y, x), branch-free. Reduction: with mn = min(|x|,|y|), mx = max(|x|,|y|) the argument of the polynomial is
// mn / mx when that is below tan(pi/8), else (mn - mx) / (mn + mx) = tan(atan(mn/mx) - pi/4) — chosen BEFORE the one
// division; atan(t) = t + t s P(s), s = t^2 <= 0.1717, P of degree 11 (Chebyshev fit, |error| < 1.2e-18 on P).
EKF_FM double fm_atan2(double y, double x) {
  const double ax = fabs(x), ay = fabs(y);
  const bool steep = ay > ax;
  const double mx = steep ? ay : ax, mn = steep ? ax : ay;  // selects: fmax / fmin carry NaN handling
  const bool big = mn > 0.41421356237309503 * mx;
  const double num = big ? mn - mx : mn;
  double den = big ? mn + mx : mx;
  den = (den == 0.0) ? 1.0 : den;  // atan2(+-0, +-0): angle 0 before the quadrant fix-ups, as libm
  const double t = fm_div(num, den);
  const double s = t * t;
  const double s2 = s * s, s4 = s2 * s2, s8 = s4 * s4;
  const double p01 = fma(0.19999999999999804, s, -0.3333333333333333);
  const double p23 = fma(0.11111111105131119, s, -0.1428571428565958);
  const double p45 = fma(0.07692296340047557, s, -0.09090908752290591);
  const double p67 = fma(0.05878921776625797, s, -0.06666424257540567);
  const double p89 = fma(0.04551350736874263, s, -0.05230401099058019);
  const double pab = fma(0.01627884915138577, s, -0.03456432733540151);
  const double q0 = fma(p23, s2, p01), q1 = fma(p67, s2, p45), q2 = fma(pab, s2, p89);
  const double p = fma(q2, s8, fma(q1, s4, q0));
  double a = fma(t * s, p, t);                                                 // atan(t), |a| <= pi/8
  a = big ? 0.7853981633974483 + (a + 3.061616997868383e-17) : a;              // + pi/4
  a = steep ? 1.5707963267948966 - (a - 6.123233995736766e-17) : a;            // pi/2 - a
#if defined(__CUDA_ARCH__)
  const bool xneg = __double2hiint(x) < 0;
#else
  const bool xneg = signbit(x);
#endif
  a = xneg ? 3.141592653589793 - (a - 1.2246467991473532e-16) : a;            // pi - a
  return copysign(a, y);
}

// sin(x), cos(x) together, branch-free for |x| <= 1e6 (three-term Cody-Waite reduction by pi/2 with FMAs, kernels
// sin r = r + r z S(z), cos r = 1 - z/2 + z^2 C(z), z = r^2 <= (pi/4)^2, S and C of degree 6 in z, quadrant by
// selects); beyond that the library call (a uniform, practically never taken branch: headings are wrapped or small).
EKF_FM void fm_sincos(double x, double* sn, double* cs) {
  if (!(fabs(x) <= 1.0e6)) {
#if defined(__CUDA_ARCH__)
    sincos(x, sn, cs);
#else
    *sn = sin(x);
    *cs = cos(x);
#endif
    return;
  }
  const double k = rint(x * 0.6366197723675814);
  double r = fma(-k, 1.5707963267948966, x);
  r = fma(-k, 6.123233995736766e-17, r);
  r = fma(-k, -1.4973849048591698e-33, r);
  const int q = (int)k;
  const double z = r * r, z2 = z * z, z4 = z2 * z2;
  const double s01 = fma(0.008333333333333331, z, -0.16666666666666666);
  const double s23 = fma(2.755731921932058e-06, z, -0.0001984126984126504);
  const double s45 = fma(1.6058530596645953e-10, z, -2.50521062259944e-08);
  const double sp = fma(fma(-7.586636882825159e-13, z2, s45), z4, fma(s23, z2, s01));
  const double c01 = fma(-0.0013888888888888887, z, 0.041666666666666664);
  const double c23 = fma(-2.755731922139249e-07, z, 2.4801587301584632e-05);
  const double c45 = fma(-1.1470460319209511e-11, z, 2.0876755787492604e-09);
  const double cp = fma(fma(4.745838412531023e-14, z2, c45), z4, fma(c23, z2, c01));
  const double sr = fma(r * z, sp, r);
  const double hz = 0.5 * z, w = 1.0 - hz;
  const double cr = w + (((1.0 - w) - hz) + z2 * cp);
  const bool swap = (q & 1) != 0;
  const double s0 = swap ? cr : sr, c0 = swap ? sr : cr;
  *sn = (q & 2) ? -s0 : s0;
  *cs = ((q + 1) & 2) ? -c0 : c0;
}

}  // namespace ekf
