// ekf_largemap_p2p.cuh — row exchange of the row-block sharded large-map update over NVLink PEER MEMORY, as device
// helpers that the update's own kernels call (sm_100a). See LmExchange in ekf_largemap.h for the protocol.
//
// Per range-bearing update every rank needs the five rows {0, 1, 2, j, j+1} of P (and their mu entries); rows 0-2
// live on rank 0, rows j, j+1 on owner(j). Round 1 first ran two NCCL broadcasts per update (~60 us
// of launch + protocol latency), then a separate push kernel per row part (14.8 us per update under ncu, and 14 % of
// pure protocol cost even at world = 1). Now there is NO extra launch: the owners' gather CTAs store their chunk of the
// rows straight into every peer's exchange buffer through peer-mapped pointers (cudaIpc, one process per GPU) and the
// last CTA publishes a sequence number; rank 0's streaming kernel does the same for the robot rows of the NEXT update
// from the registers that hold them. Consumers wait on flags in their OWN memory.
//  * ordering: data stores, __threadfence_system(), CTA barrier, then an atomic CTA count; the last CTA fences again and
//    stores the flags (the threadfence-reduction pattern); a waiter polls with volatile loads and fences before the
//    CTA reads the buffer;
//  * every spin has a cycle budget: on expiry the kernel sets an error word and goes on instead of hanging the GPU.
#pragma once
#include "ekf_largemap.h"

namespace ekf {

constexpr long long kSpinBudget = 4000000000LL;  // ~2 s of SM clocks
constexpr int kPrepushCounter = P2P_PARTS;       // index of the streaming kernel's CTA counter in LmExchange::count

// on expiry: err = max(err, code << 32 | the number waited for) — code 1: acks before a gather-side push, 2: acks
// before the streaming kernel's pre-push, 3: a row part's flag
__device__ __forceinline__ bool spin_until_ge(const volatile unsigned long long* p, unsigned long long want,
                                              unsigned long long* err, unsigned long long code = 1) {
  const long long t0 = clock64();
  while (*p < want) {
    if (clock64() - t0 > kSpinBudget) {
      atomicMax(err, (code << 32) | (want & 0xffffffffULL));
      return false;
    }
    __nanosleep(32);
  }
  return true;
}

// row slot k / flag word `part` of the buffer of update `seq` inside one rank's exchange area
__device__ __forceinline__ double* x_row(double* area, const LmExchange& x, unsigned long long seq, int k) {
  return area + ((int64_t)(seq & 1) * P2P_ROWS + k) * x.ldx;
}
__device__ __forceinline__ volatile unsigned long long* x_flag(double* area, const LmExchange& x, unsigned long long seq,
                                                               int part) {
  return reinterpret_cast<unsigned long long*>(area + x.flag_off) + (seq & 1) * P2P_PARTS + part;
}

// thread 0 of a CTA: every rank has acknowledged the gather of update `need` (so its buffer may be overwritten)
__device__ __forceinline__ void x_wait_acks(const LmExchange& x, unsigned long long need, unsigned long long code = 1) {
  if (need == 0) return;
  for (int r = 0; r < x.world; ++r)
    if (!spin_until_ge(x.acks + r, need, x.err, code)) return;
}

// After a CTA's stores into the peers' buffers: count the CTA in; the last of `total` CTAs publishes `seq` in the flag
// words of `part_mask` (bit p = part p) of the buffer of update `seq` at every peer. Call with all threads of the CTA.
__device__ __forceinline__ void x_publish(const LmExchange& x, unsigned* counter, unsigned total, unsigned long long seq,
                                          unsigned part_mask) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(counter, 1u) == total - 1) {  // every other CTA's stores are fenced and counted in
      *counter = 0;
      __threadfence_system();
      for (int d = 0; d < x.world; ++d) {
        if (d == x.rank) continue;
        for (int p = 0; p < P2P_PARTS; ++p)
          if ((part_mask >> p) & 1u) *x_flag(x.peers[d], x, seq, p) = seq;
      }
    }
  }
}

// thread 0 of a CTA: the parts of `part_mask` of this update's buffer have arrived in THIS rank's area
__device__ __forceinline__ void x_wait_parts(const LmExchange& x, const unsigned long long* flags, unsigned part_mask) {
  for (int p = 0; p < P2P_PARTS; ++p)
    if ((part_mask >> p) & 1u)
      if (!spin_until_ge(flags + p, x.seq, x.err, 3)) break;
  __threadfence_system();
}

// the streaming kernels' first CTA: this rank's gather of update x.seq is complete (the launch follows it on the stream)
__device__ __forceinline__ void x_ack(const LmExchange& x) {
  const int d = threadIdx.x;
  if (d < x.world) {
    volatile unsigned long long* ack =
        reinterpret_cast<unsigned long long*>(x.peers[d] + x.flag_off) + 2 * P2P_PARTS + x.rank;
    *ack = x.seq;
  }
}

}  // namespace ekf
