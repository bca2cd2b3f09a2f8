// ekf_largemap_p2p.cu — row exchange of the row-block sharded large-map update over NVLink PEER MEMORY (sm_100a).
//
// Per range-bearing update every rank needs the five rows {0, 1, 2, j, j+1} of P (and their mu entries); rows 0-2
// live on rank 0, rows j, j+1 on owner(j). The first implementation exported the rows and ran two
// NCCL broadcasts per update: ~0.06 ms of launch + protocol latency, which is half of a 4096-landmark update on 2
// GPUs and 85 % of it on 8. Here the owners PUSH their rows straight into every rank's exchange buffer with plain
// stores through peer-mapped pointers (cudaIpc handles, one process per GPU) and publish a sequence number; the
// consumers wait on a flag in their OWN memory. No collective, no host synchronisation, everything stream-ordered:
//
//     push (owners only) -> gather (waits on its own flags, then reads the exchange buffer) -> streaming update
//     (its first CTA acknowledges the buffer to every rank)
//
//  * the exchange buffer is double-buffered on the update's sequence number; a push into buffer s & 1 first waits
//    until every rank has acknowledged the gather of update s - 2 (acks are tiny peer stores too);
//  * ordering: data stores, __threadfence_system(), then the flag store by the last CTA of the push (the
//    threadfence-reduction pattern); the waiter polls with volatile loads and fences before returning, and the
//    gather kernel is a later launch on the same stream;
//  * every spin has a cycle budget: on expiry the kernel sets an error word and returns instead of hanging the GPU.
#include "ekf_common.cuh"
#include "ekf_largemap.h"

namespace ekf {

namespace {

constexpr long long kSpinBudget = 4000000000LL;  // ~2 s of SM clocks

__device__ __forceinline__ bool spin_until_ge(const volatile unsigned long long* p, unsigned long long want,
                                              unsigned long long* err) {
  const long long t0 = clock64();
  while (*p < want) {
    if (clock64() - t0 > kSpinBudget) {
      atomicExch(err, 1ULL);
      return false;
    }
    __nanosleep(64);
  }
  return true;
}

}  // namespace

// Owner side: copy `nrows` local rows (P rows first_row..first_row+nrows-1 -> exchange slots first_k..) and their mu
// entries into the exchange buffer `buf` of EVERY rank (blockIdx.z = destination), then publish `seq` in part `part`
// of that rank's flags.
__global__ void __launch_bounds__(256) lm_p2p_push_kernel(P2PArgs a, int first_row, int first_k, int nrows, int part) {
  const int dest = blockIdx.z;
  const int k = blockIdx.y;
  __shared__ bool ok;
  if (threadIdx.x == 0) {  // back-pressure: everyone is done reading this buffer (update seq - 2)
    ok = true;
    if (a.seq > 2)
      for (int r = 0; r < a.world && ok; ++r) ok = spin_until_ge(a.acks + r, a.seq - 2, a.err);
  }
  __syncthreads();
  if (!ok) return;
  const double* src = a.P + (int64_t)(first_row + k - a.row_begin) * a.ld;
  double* area = a.peers[dest];
  double* dst = area + ((int64_t)(a.seq & 1) * P2P_ROWS + first_k + k) * a.ldx;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < a.ld; c += (int64_t)gridDim.x * blockDim.x)
    dst[c] = src[c];
  if (blockIdx.x == 0 && threadIdx.x == 0) dst[a.ld] = a.mu[first_row + k];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y;
    if (atomicAdd(a.push_count + dest, 1u) == total - 1) {  // last CTA for this destination
      a.push_count[dest] = 0;
      __threadfence_system();
      volatile unsigned long long* flag =
          reinterpret_cast<unsigned long long*>(area + a.flag_off) + (a.seq & 1) * P2P_PARTS + part;
      *flag = a.seq;
    }
  }
}

cudaError_t lm_launch_p2p_push(const P2PArgs& a, int first_row, int first_k, int nrows, int part, cudaStream_t st) {
  dim3 grid(8, nrows, a.world);
  lm_p2p_push_kernel<<<grid, 256, 0, st>>>(a, first_row, first_k, nrows, part);
  return cudaGetLastError();
}
}  // namespace ekf
