// ekf_philox.cuh — counter-based Monte Carlo perturbations on the device (sm_100a), reproducible bit for bit on the
// host (oracle/mc_noise.py is the numpy twin; tests/test_mc_noise.py pins both to the Random123 known answers).
//
// Every perturbation of the batched Monte Carlo mode is a pure function of (seed, GLOBAL filter index, step, stream):
//   Philox4x32-10 (Salmon et al., SC'11), key = (seed lo, seed hi), counter = (filter lo, filter hi, step, stream)
//   -> two uniforms u = (k + 0.5) 2^-52, k = (w0 >> 6) 2^26 + (w1 >> 6)   (exact in fp64, strictly inside (0, 1))
//   -> a pair of standard normals by Box-Muller: r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2).
// The logarithm and the sine / cosine use ONLY IEEE add / multiply / divide / sqrt, each rounded separately in a fixed
// order (__dadd_rn / __dmul_rn / __ddiv_rn / __dsqrt_rn: the compiler may not contract them into FMAs), which is
// what numpy does on the host — so an ensemble member's inputs can be regenerated off the device and replayed through
// the oracle, whatever the sharding (the counter holds the global filter index, not the rank-local one).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

enum : unsigned { kStreamOdometry = 0, kStreamObservation0 = 1, kStreamDetection0 = 64 };

struct Philox4 {
  uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                          uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}

#ifdef __CUDACC__
__device__ __forceinline__ double pn_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double pn_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double pn_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double pn_div(double a, double b) { return __ddiv_rn(a, b); }

__device__ __forceinline__ double uniform52(uint32_t hi, uint32_t lo) {
  const uint64_t k = ((uint64_t)(hi >> 6) << 26) + (uint64_t)(lo >> 6);
  return pn_mul(pn_add((double)k, 0.5), 2.220446049250313e-16 /* 2^-52 */);
}

// log(x), x > 0: x = m 2^e, m in [sqrt(1/2), sqrt 2); log m = 2 s (1 + z/3 + z^2/5 + ...), s = (m-1)/(m+1), z = s^2
__device__ __forceinline__ double det_log(double x) {
  int e;
  double m = frexp(x, &e);  // exact
  if (m < 0.7071067811865476) {
    m = pn_mul(m, 2.0);
    e -= 1;
  }
  const double s = pn_div(pn_sub(m, 1.0), pn_add(m, 1.0));
  const double z = pn_mul(s, s);
  double p = 1.0 / 23.0;
  p = pn_add(pn_mul(p, z), 1.0 / 21.0);
  p = pn_add(pn_mul(p, z), 1.0 / 19.0);
  p = pn_add(pn_mul(p, z), 1.0 / 17.0);
  p = pn_add(pn_mul(p, z), 1.0 / 15.0);
  p = pn_add(pn_mul(p, z), 1.0 / 13.0);
  p = pn_add(pn_mul(p, z), 1.0 / 11.0);
  p = pn_add(pn_mul(p, z), 1.0 / 9.0);
  p = pn_add(pn_mul(p, z), 1.0 / 7.0);
  p = pn_add(pn_mul(p, z), 1.0 / 5.0);
  p = pn_add(pn_mul(p, z), 1.0 / 3.0);
  p = pn_add(pn_mul(p, z), 1.0);
  return pn_add(pn_mul((double)e, 0.6931471805599453), pn_mul(pn_mul(2.0, s), p));
}

// (sin, cos) of 2 pi t, t in [0, 1): octant k = floor(8 t) (exact), Taylor polynomials on [0, pi/4], octant symmetries
// by selection and sign — operation order == oracle/mc_noise.py:sincos_2pi. Straight-line.
__device__ __forceinline__ void sincos_2pi(double t, double* s, double* c) {
  const double x = pn_mul(t, 8.0);
  const double kf = floor(x);
  const double f = pn_sub(x, kf);
  const int k = (int)kf;
  const double g = (k & 1) ? pn_sub(1.0, f) : f;
  const double a = pn_mul(g, 0.7853981633974483);
  const double a2 = pn_mul(a, a);
  double ps = 1.0 / 355687428096000.0, pc = 1.0 / 20922789888000.0;
  ps = pn_add(pn_mul(ps, a2), -1.0 / 1307674368000.0); pc = pn_add(pn_mul(pc, a2), -1.0 / 87178291200.0);
  ps = pn_add(pn_mul(ps, a2), 1.0 / 6227020800.0);     pc = pn_add(pn_mul(pc, a2), 1.0 / 479001600.0);
  ps = pn_add(pn_mul(ps, a2), -1.0 / 39916800.0);      pc = pn_add(pn_mul(pc, a2), -1.0 / 3628800.0);
  ps = pn_add(pn_mul(ps, a2), 1.0 / 362880.0);         pc = pn_add(pn_mul(pc, a2), 1.0 / 40320.0);
  ps = pn_add(pn_mul(ps, a2), -1.0 / 5040.0);          pc = pn_add(pn_mul(pc, a2), -1.0 / 720.0);
  ps = pn_add(pn_mul(ps, a2), 1.0 / 120.0);            pc = pn_add(pn_mul(pc, a2), 1.0 / 24.0);
  ps = pn_add(pn_mul(ps, a2), -1.0 / 6.0);             pc = pn_add(pn_mul(pc, a2), -1.0 / 2.0);
  const double sa = pn_add(a, pn_mul(a, pn_mul(a2, ps)));
  const double ca = pn_add(1.0, pn_mul(a2, pc));
  const bool swap = ((k + 1) >> 1) & 1;
  double ss = swap ? ca : sa, cc = swap ? sa : ca;
  if (k >= 4) ss = -ss;
  if (k >= 2 && k <= 5) cc = -cc;
  *s = ss;
  *c = cc;
}

// the two standard normals of (seed, global filter index, step, stream)
__device__ __forceinline__ void normal_pair(uint64_t seed, uint64_t filter, uint32_t step, uint32_t stream, double* z0,
                                            double* z1) {
  const Philox4 w = philox4x32_10((uint32_t)filter, (uint32_t)(filter >> 32), step, stream, (uint32_t)seed,
                                  (uint32_t)(seed >> 32));
  // Box-Muller: r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2)
  const double r = __dsqrt_rn(pn_mul(-2.0, det_log(uniform52(w.x, w.y))));
  double sn, cs;
  sincos_2pi(uniform52(w.z, w.w), &sn, &cs);
  *z0 = pn_mul(r, cs);
  *z1 = pn_mul(r, sn);
}
#endif  // __CUDACC__

}  // namespace ekf
