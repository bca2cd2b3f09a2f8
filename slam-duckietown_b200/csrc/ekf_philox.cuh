// ekf_philox.cuh — counter-based Monte Carlo perturbations on the device (sm_100a), reproducible bit for bit on the
// host (oracle/mc_noise.py is the numpy twin; tests/test_mc_noise.py pins both to the Random123 known answers).
//
// Every perturbation of the batched Monte Carlo mode is a pure function of (seed, GLOBAL filter index, step, stream):
//   Philox4x32-10 (Salmon et al., SC'11), key = (seed lo, seed hi), counter = (filter lo, filter hi, step, stream)
//   -> two uniforms u = (k + 0.5) 2^-52, k = (w0 >> 6) 2^26 + (w1 >> 6)   (exact in fp64, strictly inside (0, 1))
//   -> two standard normals by the inverse CDF (Wichura's AS241 PPND16).
// The inverse CDF and its logarithm use ONLY IEEE add / multiply / divide / sqrt, each rounded separately in a fixed
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

#define EKF_PN_HORNER(r, c0, c1, c2, c3, c4, c5, c6, c7)                                                        \
  pn_add(pn_mul(pn_add(pn_mul(pn_add(pn_mul(pn_add(pn_mul(pn_add(pn_mul(pn_add(pn_mul(pn_add(pn_mul((c7), (r)), \
  (c6)), (r)), (c5)), (r)), (c4)), (r)), (c3)), (r)), (c2)), (r)), (c1)), (r)), (c0))

// standard-normal quantile of u in (0, 1): AS241 PPND16, operation order == oracle/mc_noise.py:normal_icdf
__device__ __forceinline__ double normal_icdf(double u) {
  const double q = pn_sub(u, 0.5);
  if (fabs(q) <= 0.425) {
    const double r = pn_sub(0.180625, pn_mul(q, q));
    const double num = EKF_PN_HORNER(r, 3.3871328727963666080, 1.3314166789178437745e+2, 1.9715909503065514427e+3,
                                     1.3731693765509461125e+4, 4.5921953931549871457e+4, 6.7265770927008700853e+4,
                                     3.3430575583588128105e+4, 2.5090809287301226727e+3);
    const double den = EKF_PN_HORNER(r, 1.0, 4.2313330701600911252e+1, 6.8718700749205790830e+2,
                                     5.3941960214247511077e+3, 2.1213794301586595867e+4, 3.9307895800092710610e+4,
                                     2.8729085735721942674e+4, 5.2264952788528545610e+3);
    return pn_div(pn_mul(q, num), den);
  }
  double r = (q < 0.0) ? u : pn_sub(1.0, u);
  r = __dsqrt_rn(-det_log(r));
  double v;
  if (r <= 5.0) {
    r = pn_sub(r, 1.6);
    const double num = EKF_PN_HORNER(r, 1.42343711074968357734, 4.63033784615654529590, 5.76949722146069140550,
                                     3.64784832476320460504, 1.27045825245236838258, 2.41780725177450611770e-1,
                                     2.27238449892691845833e-2, 7.74545014278341407640e-4);
    const double den = EKF_PN_HORNER(r, 1.0, 2.05319162663775882187, 1.67638483018380384940,
                                     6.89767334985100004550e-1, 1.48103976427480074590e-1, 1.51986665636164571966e-2,
                                     5.47593808499534494600e-4, 1.05075007164441684324e-9);
    v = pn_div(num, den);
  } else {
    r = pn_sub(r, 5.0);
    const double num = EKF_PN_HORNER(r, 6.65790464350110377720, 5.46378491116411436990, 1.78482653991729133580,
                                     2.96560571828504891230e-1, 2.65321895265761230930e-2, 1.24266094738807843860e-3,
                                     2.71155556874348757815e-5, 2.01033439929228813265e-7);
    const double den = EKF_PN_HORNER(r, 1.0, 5.99832206555887937690e-1, 1.36929880922735805310e-1,
                                     1.48753612908506148525e-2, 7.86869131145613259100e-4, 1.84631831751005468180e-5,
                                     1.42151175831644588870e-7, 2.04426310338993978564e-15);
    v = pn_div(num, den);
  }
  return (q < 0.0) ? -v : v;
}

// the two standard normals of (seed, global filter index, step, stream)
__device__ __forceinline__ void normal_pair(uint64_t seed, uint64_t filter, uint32_t step, uint32_t stream, double* z0,
                                            double* z1) {
  const Philox4 w = philox4x32_10((uint32_t)filter, (uint32_t)(filter >> 32), step, stream, (uint32_t)seed,
                                  (uint32_t)(seed >> 32));
  *z0 = normal_icdf(uniform52(w.x, w.y));
  *z1 = normal_icdf(uniform52(w.z, w.w));
}
#endif  // __CUDACC__

}  // namespace ekf
