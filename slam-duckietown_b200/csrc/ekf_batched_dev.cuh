// ekf_batched_dev.cuh — device helpers shared by the batched step kernels: mbarrier + 1-D TMA bulk copies,
// 128/256-bit shared/global accesses and the rank-2 element update (sm_100a).
#pragma once
#include "ekf_common.cuh"
#include "ekf_philox.cuh"

namespace ekf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on the mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// L2 prefetch of the data of the filter that will occupy this CTA's place on the SM one "generation" later. Measured
// (tools/phase_probe.py, config M): under load a CTA waits ~30 % of its life for its inputs, mu and P to arrive from
// a saturated HBM; prefetched into L2 about 3/8 of a generation ahead they arrive at L2 latency instead
// (789 M -> 1 049 M filter-steps/s at config M, the plateau is 1/4 .. 1/2 of a generation).
__device__ __forceinline__ void l2_prefetch_bulk(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void l2_prefetch(const void* src) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(src));
}

// one rank-2 (plain) step on a single element; the SAME sequence is used wherever an entry of an
// intermediate P_u is needed, so phase 2's rebuilt rows/columns equal what phase 3 finally writes.
__device__ __forceinline__ double rank2(double v, double k0, double k1, double w0, double w1) {
  return fma(-k1, w1, fma(-k0, w0, v));
}
__device__ __forceinline__ void rank2x4(double4x& p, double k0, double k1, const double4x& w0,
                                        const double4x& w1) {
  p.a = rank2(p.a, k0, k1, w0.a, w1.a);
  p.b = rank2(p.b, k0, k1, w0.b, w1.b);
  p.c = rank2(p.c, k0, k1, w0.c, w1.c);
  p.d = rank2(p.d, k0, k1, w0.d, w1.d);
}

constexpr int kTailDoubles = 4;  // mbarrier
constexpr unsigned kFull = 0xffffffffu;

struct __align__(16) d2 {
  double x, y;
};
__device__ __forceinline__ double4x lds256(const double* p) { return *reinterpret_cast<const double4x*>(p); }
__device__ __forceinline__ d2 lds128(const double* p) { return *reinterpret_cast<const d2*>(p); }
__device__ __forceinline__ void stg128(double* p, const d2& v) {
  asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

}  // namespace ekf
