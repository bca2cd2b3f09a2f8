// ekf_common.cuh — device helpers shared by the batched, large-map and 3-state kernels (sm_100a, fp64).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "ekf_fastmath.cuh"

namespace ekf {

constexpr double kPi = 3.14159265358979323846;
constexpr double kTwoPi = 6.28318530717958647692;
constexpr double kNewLandmarkVar = 10000.0;  // replay_no_ros.py:356-357

// flag bits shared by host and device
enum : int { kFlagCircular = 1, kFlagMeasurement = 2, kFlagDisableMotion = 4, kFlagJoseph = 8 };

// (a + pi) % (2 pi) - pi with Python float-modulo semantics (replay_no_ros.py:397,:458): the result of the
// modulo lies in [0, 2 pi). r - floor(r / 2pi) * 2pi is evaluated with one FMA (exact whenever the quotient is
// the true one, like fmod); a quotient that rounding pushed across an integer is repaired by the two
// range checks, which changes the value by at most one rounding (1e-16), far inside the 1e-9 parity bound.
__host__ __device__ __forceinline__ double wrap_pi(double a) {
  const double r = a + kPi;
  const double k = floor(r * (1.0 / kTwoPi));
  double v = fma(-k, kTwoPi, r);
  if (v < 0.0) v += kTwoPi;
  if (v >= kTwoPi) v -= kTwoPi;
  return v - kPi;
}

// a1/a2 (replay_no_ros.py:262-266, :491-497)
__host__ __device__ __forceinline__ double delta_phi(long long ticks, long long prev, long long resolution) {
  return (double)(ticks - prev) * (kTwoPi / (double)resolution);
}
__host__ __device__ __forceinline__ void displacement(double R, double baseline, double dl, double dr,
                                                      double* ang, double* lin) {
#ifdef __CUDA_ARCH__
  // no FMA contraction: numpy rounds R*dphi before the add, and tick-derived odometry must be bit-identical
  // to the host path (tests: ticks entry point == delta_phi/displacement on the host + plain step)
  const double wr = __dmul_rn(R, dr), wl = __dmul_rn(R, dl);
  *lin = __dadd_rn(wl, wr) / 2.0;
  *ang = __dsub_rn(wr, wl) / baseline;
#else
  const double wr = R * dr, wl = R * dl;
  *lin = (wl + wr) / 2.0;
  *ang = (wr - wl) / baseline;
#endif
}

// a6 (replay_no_ros.py:368-417): pose increment, new heading and Jacobian column g0 = G[0,2], g1 = G[1,2]
struct Motion {
  double dx, dy, theta, g0, g1;
};
// (s, c) = sincos(th), (s2, c2) = sincos(th + ang) are passed in so that a warp can evaluate both in one call
__device__ __forceinline__ Motion motion_from_sincos(double th, double ang, double lin, int flags, double s,
                                                     double c, double s2, double c2) {
  Motion m;
  if (flags & kFlagDisableMotion) {
    m.dx = 0.0; m.dy = 0.0; m.theta = th; m.g0 = 0.0; m.g1 = 0.0;
    return m;
  }
  if ((flags & kFlagCircular) && fabs(ang) > 1e-2) {
    // every product rounded before the add, as numpy evaluates :391-403 (and so that no kernel's FMA contraction
    // choice can differ from another's)
    const double r = fm_div(lin, ang);
    const double rs = __dmul_rn(r, s), rs2 = __dmul_rn(r, s2), rc = __dmul_rn(r, c), rc2 = __dmul_rn(r, c2);
    m.dx = __dadd_rn(-rs, rs2);
    m.dy = __dsub_rn(rc, rc2);
    m.theta = wrap_pi(th + ang);
    m.g0 = __dadd_rn(-rc, rc2);
    m.g1 = __dadd_rn(-rs, rs2);
  } else {
    m.dx = lin * c;
    m.dy = lin * s;
    // straight branch of the circular model leaves theta alone; the linear model adds ang unwrapped (:405-410)
    m.theta = (flags & kFlagCircular) ? th : th + ang;
    m.g0 = -lin * s;
    m.g1 = lin * c;
  }
  return m;
}
__device__ __forceinline__ Motion motion_step(double th, double ang, double lin, int flags) {
  double s, c, s2, c2;
  fm_sincos(th, &s, &c);
  fm_sincos(th + ang, &s2, &c2);
  return motion_from_sincos(th, ang, lin, flags, s, c, s2, c2);
}

// a8 (replay_no_ros.py:443-469): innovation y and the 2x5 block of H over columns {0,1,2,j,j+1}
struct Innov {
  double y0, y1;
  double h[2][5];
  double q;
};
__device__ __forceinline__ Innov innovation(double mx, double my, double mth, double lx, double ly, double rng,
                                            double brg) {
  Innov o;
  const double dx = lx - mx, dy = ly - my;
  const double q = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
  // sqrt, 1/sqrt and atan2 as straight-line code (ekf_fastmath.cuh): the three overlap instead of running as
  // three library calls with a slow-path branch each; H entries from 1/sqrt(q): sqrt(q) d / q = d / sqrt(q)
  double sq, isq;
  fm_sqrt_rsqrt(q, &sq, &isq);
  o.q = q;
  o.y0 = rng - sq;
  o.y1 = wrap_pi(brg - (fm_atan2(dy, dx) - mth));
  const double rq = isq * isq;
  const double ax = dx * isq, ay = dy * isq, bx = dx * rq, by = dy * rq;
  o.h[0][0] = -ax; o.h[0][1] = -ay; o.h[0][2] = 0.0; o.h[0][3] = ax;  o.h[0][4] = ay;
  o.h[1][0] = by;  o.h[1][1] = -bx; o.h[1][2] = -1.0; o.h[1][3] = -by; o.h[1][4] = bx;
  return o;
}

// closed-form inverse of the 2x2 innovation covariance (replaces np.linalg.inv, :473)
__device__ __forceinline__ void inv2(const double s[2][2], double si[2][2]) {
  const double det = __dsub_rn(__dmul_rn(s[0][0], s[1][1]), __dmul_rn(s[0][1], s[1][0]));
  const double rd = fm_rcp(det);
  si[0][0] = s[1][1] * rd;  si[0][1] = -s[0][1] * rd;
  si[1][0] = -s[1][0] * rd; si[1][1] = s[0][0] * rd;
}

// a0 b0 + a1 b1 with the rounding sequence spelled out (first product rounded, second fused into the add)
__device__ __forceinline__ double dot2(double a0, double b0, double a1, double b1) {
  return fma(a1, b1, __dmul_rn(a0, b0));
}

// a10 (replay_no_ros.py:476) for one entry: mu + (K[i,0] y0 + K[i,1] y1), with the rounding sequence spelled out so
// that every large-map kernel that applies it (sequential, fused multi-update, shard pose refresh) agrees bitwise
__device__ __forceinline__ double mu_inc(double mu, double k0, double k1, double y0, double y1) {
  return __dadd_rn(mu, fma(k1, y1, __dmul_rn(k0, y0)));
}

// Joseph form as a one-pass rank-4 update: W = H P - Q K^T and AHt = P H^T - K (S - Q), entry by
// entry, with the rounding sequence spelled out so that every kernel that forms them (staged / W-in-registers /
// sub-warp group batched kernels, large-map gathers) agrees bitwise. sq = S - Q (the innovation covariance before the
// measurement noise is added).
__device__ __forceinline__ double joseph_w(double hp, double qn, double k) { return fma(-qn, k, hp); }
__device__ __forceinline__ double joseph_a(double pht, double k0, double k1, double sq0, double sq1) {
  return __dsub_rn(pht, fma(k1, sq1, __dmul_rn(k0, sq0)));
}

// 256-bit global accesses (sm_100a: LDG.E.256 / STG.E.256)
struct __align__(32) double4x {
  double a, b, c, d;
};
__device__ __forceinline__ double4x ldg256_stream(const double* p) {
  double4x v;
  asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.a), "=d"(v.b), "=d"(v.c), "=d"(v.d) : "l"(p));
  return v;
}
__device__ __forceinline__ double4x ldg256(const double* p) {
  double4x v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.a), "=d"(v.b), "=d"(v.c), "=d"(v.d) : "l"(p));
  return v;
}
__device__ __forceinline__ void stg256(double* p, const double4x& v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.a), "d"(v.b), "d"(v.c), "d"(v.d)
               : "memory");
}

}  // namespace ekf
