// ekf_largemap_p2p.cuh — row exchange of the row-block sharded large-map update over NVLink PEER MEMORY, as device
// helpers that the update's own kernels call (sm_100a). See LmExchange in ekf_largemap.h for the protocol.
//
// Per range-bearing update every rank needs the five rows {0, 1, 2, j, j+1} of P (and their mu entries); rows 0-2
// live on rank 0, rows j, j+1 on owner(j). Round 1 first ran two NCCL broadcasts per update (~60 us
// of launch + protocol latency), then a separate push kernel per row part (14.8 us per update under ncu). Round 2
// moved the stores into the update's own kernels, first with a flag protocol (data stores, __threadfence_system(),
// CTA count, second fence, flag store; consumer: poll flag, fence): correct, but the globaltimer timeline
// (tools/lm_p2p_timeline.py, profiles/r2t_p2p_timeline_flags_8gpu.txt) showed that the two system-scope fences and the
// flag round trip cost 8.6-11.8 us per push — 65 KB of data — and stretched rank 0's streaming kernel from 13.3 to
// 19.5 us. Now the data carries its own validity (the LL idea of NCCL's low-latency protocol):
//  * a row of the exchange buffer is an array of 16-byte CELLS {value.lo, tag, value.hi, tag}, tag = the low 32 bits
//    of the update's sequence number; a cell is written with ONE 16-byte volatile store of two 8-byte elements, each
//    carrying its own tag, so nothing beyond the atomicity of an aligned 8-byte element is assumed;
//  * the producer just stores cells into every peer's buffer through peer-mapped pointers (cudaIpc, one process per
//    GPU): no fence, no flag, no CTA counter; the consumer thread that needs an entry polls that cell with 16-byte
//    volatile loads until both tags match. One-way store latency is all that is left on the critical path;
//  * buffer of update s is s & 1; before overwriting it the producer checks (one lane per rank, local memory) that
//    every rank has acknowledged the gather of update s - 2 (landmark part) resp. s - 1 (pre-push into s + 1);
//  * every spin has a cycle budget: on expiry the kernel sets an error word and goes on instead of hanging the GPU.
#pragma once
#include "ekf_largemap.h"

#ifndef EKF_TL  // diagnostic stamps exist only in the timeline build of ekf_largemap.cu
#define EKF_TL(seq, k) ((void)0)
#define EKF_TLMAX(seq, k) ((void)0)
#endif

namespace ekf {

constexpr long long kSpinBudget = 4000000000LL;  // ~2 s of SM clocks

// cells of row slot k of the buffer of update `seq` inside one rank's exchange area (x.ldx cells per row)
__device__ __forceinline__ double* x_row(double* area, const LmExchange& x, unsigned long long seq, int k) {
  return area + ((int64_t)(seq & 1) * P2P_ROWS + k) * x.ldx * 2;
}

// a cell is two 8-byte words {tag << 32 | value.lo, tag << 32 | value.hi}: every word is one scalar element of the
// vector access, so a reader can never see half a word
__device__ __forceinline__ void ll_store(double* row_cells, int64_t c, double v, unsigned long long seq) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  const unsigned long long tag = (unsigned long long)(unsigned)seq << 32;
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(reinterpret_cast<uint4*>(row_cells) + c),
               "l"((b & 0xffffffffULL) | tag), "l"((b >> 32) | tag)
               : "memory");
}

// two adjacent cells (c even: 32-byte aligned) in one 32-byte store
__device__ __forceinline__ void ll_store2(double* row_cells, int64_t c, double va, double vb, unsigned long long seq) {
  const unsigned long long a = (unsigned long long)__double_as_longlong(va);
  const unsigned long long b = (unsigned long long)__double_as_longlong(vb);
  const unsigned long long tag = (unsigned long long)(unsigned)seq << 32;
  asm volatile("st.volatile.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(reinterpret_cast<uint4*>(row_cells) + c),
               "l"((a & 0xffffffffULL) | tag), "l"((a >> 32) | tag), "l"((b & 0xffffffffULL) | tag), "l"((b >> 32) | tag)
               : "memory");
}

// one look at a cell (no waiting): several of these can be in flight before the first ll_resolve
struct LLRaw {
  unsigned long long w0, w1;  // tag << 32 | lo, tag << 32 | hi
};
__device__ __forceinline__ LLRaw ll_peek(const double* row_cells, int64_t c) {
  LLRaw r;
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];"
               : "=l"(r.w0), "=l"(r.w1)
               : "l"(reinterpret_cast<const uint4*>(row_cells) + c)
               : "memory");
  return r;
}
__device__ __forceinline__ bool ll_valid(const LLRaw& r, unsigned long long seq) {
  return (unsigned)(r.w0 >> 32) == (unsigned)seq && (unsigned)(r.w1 >> 32) == (unsigned)seq;
}
__device__ __forceinline__ double ll_value(const LLRaw& r) {
  return __longlong_as_double((long long)((r.w1 << 32) | (r.w0 & 0xffffffffULL)));
}

// on expiry: err = max(err, 3 << 32 | tag)
__device__ __forceinline__ double ll_load(const double* row_cells, int64_t c, unsigned long long seq,
                                          unsigned long long* err) {
  LLRaw r;
  long long start = 0;
  for (int it = 0;; ++it) {
    r = ll_peek(row_cells, c);
    if (ll_valid(r, seq)) break;
    if (it == 0) {
      start = clock64();
    } else if (clock64() - start > kSpinBudget) {
      atomicMax(err, (3ULL << 32) | (seq & 0xffffffffULL));
      break;
    }
  }
  return ll_value(r);
}

__device__ __forceinline__ double ll_resolve(const LLRaw& r, const double* row_cells, int64_t c, unsigned long long seq,
                                             unsigned long long* err) {
  return ll_valid(r, seq) ? ll_value(r) : ll_load(row_cells, c, seq, err);
}

// entry c of an exchanged row: a cell of the exchange buffer (ll) or a plain double (own P / replicated rows buffer)
__device__ __forceinline__ double x_read(const double* row, int64_t c, bool ll, const LmExchange& x) {
  return ll ? ll_load(row, c, x.seq, x.err) : row[c];
}

// warp 0 of a CTA (all 32 lanes): every rank has acknowledged the gather of update `need`, so the buffer it read may
// be overwritten. Lane r polls rank r's ack word (local memory). on expiry: err = max(err, code << 32 | need) —
// code 1: before a gather-side push, 2: before the streaming kernel's pre-push
__device__ __forceinline__ void x_wait_acks(const LmExchange& x, unsigned long long need, unsigned long long code = 1) {
  const int lane = threadIdx.x & 31;
  if (need != 0 && lane < x.world) {
    const volatile unsigned long long* p = x.acks + lane;
    const long long t0 = clock64();
    while (*p < need) {
      if (clock64() - t0 > kSpinBudget) {
        atomicMax(x.err, (code << 32) | (need & 0xffffffffULL));
        break;
      }
    }
  }
  __syncwarp();
}

// the streaming kernels' first CTA: this rank's gather of update x.seq is complete (the launch follows it on the stream)
__device__ __forceinline__ void x_ack(const LmExchange& x) {
  const int d = threadIdx.x;
  if (d < x.world) {
    volatile unsigned long long* ack = reinterpret_cast<unsigned long long*>(x.peers[d] + x.flag_off) + x.rank;
    *ack = x.seq;
  }
}

}  // namespace ekf
