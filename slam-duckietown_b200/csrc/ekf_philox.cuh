// ekf_philox.cuh — counter-based Monte Carlo perturbations on the device (sm_100a), reproducible bit for bit on the
// host (oracle/mc_noise.py is the numpy twin; tests/test_mc_noise.py pins both to the Random123 known answers).
//
// Every perturbation of the batched Monte Carlo mode is a pure function of (seed, GLOBAL filter index, step, stream):
//   Philox4x32-10 (Salmon et al., SC'11), key = (seed lo, seed hi), counter = (filter lo, filter hi, step, stream)
//   -> two uniforms strictly inside (0, 1) with 23 random bits each, u = (k + 0.5) 2^-23, k = w >> 9 (exact in fp32);
//      where the radial one draws k = 0 (once in 2^23) a second word refines it to (k' + 0.5) 2^-46, so the tails reach
//      8 sigma instead of stopping at 5.8
//   -> a pair of standard normals by Box-Muller: r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2).
// The arithmetic is SINGLE precision (the deviates are sensor / odometry noise, scaled by sigmas of 1e-3..1e-2: 2e-7
// relative is far below anything the filter resolves) and uses ONLY IEEE add / multiply / divide / sqrt, each rounded
// separately in a fixed order (__fadd_rn / __fmul_rn / __fdiv_rn / __fsqrt_rn: the compiler may not contract them into
// FMAs), which is what numpy float32 does on the host — so an ensemble member's inputs can be regenerated off the
// device and replayed through the oracle, whatever the sharding (the counter holds the global filter index, not the
// rank-local one). Round 2 first did this in double precision with 52-bit uniforms: ~270 instructions per draw, half
// of them on the FP64 pipe the step kernels live on, +12-15 % on the generated step (profiles/r2y_mc_e2e_probe.txt);
// the single-precision draw is ~140 instructions on the otherwise idle FP32 pipe.
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
__device__ __forceinline__ float pn_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float pn_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float pn_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float pn_div(float a, float b) { return __fdiv_rn(a, b); }

// (k + 0.5) 2^-23, k = w >> 9: 24 significant bits, exact, strictly inside (0, 1)
__device__ __forceinline__ float uniform23(uint32_t w) {
  return pn_mul(pn_add((float)(w >> 9), 0.5f), 0x1p-23f);
}
// the radial uniform: as uniform23(w), refined by a second word where k = 0 (exact as well)
__device__ __forceinline__ float uniform23_tail(uint32_t w, uint32_t w2) {
  const uint32_t k = w >> 9;
  const float coarse = pn_mul(pn_add((float)k, 0.5f), 0x1p-23f);
  const float fine = pn_mul(pn_add((float)(w2 >> 9), 0.5f), 0x1p-46f);
  return k ? coarse : fine;
}

// log(x), x > 0: x = m 2^e, m in [sqrt(1/2), sqrt 2); log m = 2 s (1 + z/3 + z^2/5 + z^3/7 + z^4/9), s = (m-1)/(m+1)
__device__ __forceinline__ float det_log(float x) {
  int e;
  float m = frexpf(x, &e);  // exact
  if (m < 0x1.6a09e6p-1f) {
    m = pn_mul(m, 2.0f);
    e -= 1;
  }
  const float s = pn_div(pn_sub(m, 1.0f), pn_add(m, 1.0f));
  const float z = pn_mul(s, s);
  float p = 0x1.c71c72p-4f;                       // 1/9
  p = pn_add(pn_mul(p, z), 0x1.24924ap-3f);       // 1/7
  p = pn_add(pn_mul(p, z), 0x1.99999ap-3f);       // 1/5
  p = pn_add(pn_mul(p, z), 0x1.555556p-2f);       // 1/3
  p = pn_add(pn_mul(p, z), 1.0f);
  return pn_add(pn_mul((float)e, 0x1.62e43p-1f), pn_mul(pn_mul(2.0f, s), p));
}

// (sin, cos) of 2 pi t, t in [0, 1): octant k = floor(8 t) (exact), Taylor polynomials on [0, pi/4], octant symmetries
// by selection and sign — operation order == oracle/mc_noise.py:sincos_2pi. Straight-line.
__device__ __forceinline__ void sincos_2pi(float t, float* s, float* c) {
  const float x = pn_mul(t, 8.0f);
  const float kf = floorf(x);
  const float f = pn_sub(x, kf);
  const int k = (int)kf;
  const float g = (k & 1) ? pn_sub(1.0f, f) : f;
  const float a = pn_mul(g, 0x1.921fb6p-1f);      // pi/4
  const float a2 = pn_mul(a, a);
  float ps = 0x1.71de3ap-19f, pc = 0x1.a01a02p-16f;                                         // 1/9!, 1/8!
  ps = pn_add(pn_mul(ps, a2), -0x1.a01a02p-13f); pc = pn_add(pn_mul(pc, a2), -0x1.6c16c2p-10f);  // -1/7!, -1/6!
  ps = pn_add(pn_mul(ps, a2), 0x1.111112p-7f);   pc = pn_add(pn_mul(pc, a2), 0x1.555556p-5f);    // 1/5!, 1/4!
  ps = pn_add(pn_mul(ps, a2), -0x1.555556p-3f);  pc = pn_add(pn_mul(pc, a2), -0x1p-1f);          // -1/3!, -1/2!
  const float sa = pn_add(a, pn_mul(a, pn_mul(a2, ps)));
  const float ca = pn_add(1.0f, pn_mul(a2, pc));
  const bool swap = ((k + 1) >> 1) & 1;
  float ss = swap ? ca : sa, cc = swap ? sa : ca;
  if (k >= 4) ss = -ss;
  if (k >= 2 && k <= 5) cc = -cc;
  *s = ss;
  *c = cc;
}

// the two standard normals of (seed, global filter index, step, stream), widened to double (exact)
__device__ __forceinline__ void normal_pair(uint64_t seed, uint64_t filter, uint32_t step, uint32_t stream, double* z0,
                                            double* z1) {
  const Philox4 w = philox4x32_10((uint32_t)filter, (uint32_t)(filter >> 32), step, stream, (uint32_t)seed,
                                  (uint32_t)(seed >> 32));
  // Box-Muller: r = sqrt(-2 log u1), (r cos 2 pi u2, r sin 2 pi u2)
  const float r = __fsqrt_rn(pn_mul(-2.0f, det_log(uniform23_tail(w.x, w.y))));
  float sn, cs;
  sincos_2pi(uniform23(w.z), &sn, &cs);
  *z0 = (double)pn_mul(r, cs);
  *z1 = (double)pn_mul(r, sn);
}
#endif  // __CUDACC__

}  // namespace ekf
