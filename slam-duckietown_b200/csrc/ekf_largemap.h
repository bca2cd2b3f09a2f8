// ekf_largemap.h — argument blocks and launchers of the large-map kernels (internal, not the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

namespace ekf {

// Programmatic dependent launch (sm_90+): the large-map kernels of one update form a chain on one stream
// (gather -> streaming update -> next gather ...). Launched with this attribute a kernel's CTAs may be scheduled
// while its predecessor drains; every such kernel begins with pdl_wait() (griddepcontrol.wait: predecessor complete
// and its memory visible) before it reads or writes anything, so only launch latency and CTA ramp-up overlap.
// EKF_LM_PDL=0 switches it off.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  static const bool on = [] { const char* v = getenv("EKF_LM_PDL"); return !(v && v[0] == '0'); }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

constexpr int LM_THREADS = 128;  // threads per CTA of the streaming update = column groups per tile

// slots of the small scalar workspace (device doubles)
enum : int { kScalG0 = 0, kScalG1 = 1, kScalY0 = 2, kScalY1 = 3, kScalArrive = 6 /* u64 arrival counter of the predict */, kScalCount = 8 };

struct LmGatherArgs {
  const double* P;        // local row block
  const double* row[5];   // rows {0,1,2,j,j+1}: into P itself (single GPU) or the exchanged buffer (shard)
  const double* mu5[5];   // mu entries {0,1,2,j,j+1}
  double *K0, *K1, *A0, *A1, *W0, *W1, *scal;
  int64_t ld;
  int n, row_begin, row_end, j, flags;
  int shard;              // 1: rows/mu5 come from the exchange buffer of a row-block shard
  double* mu_out;         // shard only: replicated pose entries are refreshed here
  double rng, brg, mm;
  // peer-memory exchange: wait until both flags of the exchange buffer reached wait_seq before reading `row`
  const unsigned long long* wait_flags;  // null = no wait; else wait_parts consecutive flag words
  unsigned long long wait_seq;
  unsigned long long* wait_err;
  int wait_parts;
};

struct LmUpdateArgs {
  double* P;              // local row block
  double* mu;
  const double *K0, *K1, *A0, *A1, *W0, *W1, *scal;
  int64_t ld;
  int n, row_begin, row_end;
  int full_tiles, tail_groups;  // filled by the launcher
  // peer-memory exchange: the first CTA tells every rank that this rank's gather of update ack_seq is done
  double* const* ack_peers;     // null = no ack
  int64_t ack_flag_off;
  unsigned long long ack_seq;
  int ack_world, ack_rank;
};

constexpr int LM_MAX_FUSED = 4;  // updates fused into one pass over P (ekf_largemap_multi.cu)
constexpr int LM_MAX_ROWS = 3 + 2 * LM_MAX_FUSED;  // rows a fused step exchanges

// fused multi-update: m sequential updates of one step, ONE read + write of P (plain form, whole matrix on one GPU)
struct LmMultiArgs {
  double* P;
  double* mu;
  double* K;    // [m][2][ldw]: K_u[:,0] at K + 2u ldw, K_u[:,1] at K + (2u+1) ldw
  double* W;    // [m][2][ldw]: (H P)_u rows, same indexing
  double* y;    // [m][2] innovations
  double* KR;   // [m][LM_MAX_ROWS][2] K_u at the exchanged rows R = {0,1,2} U {j_u, j_u+1} (replicated on every rank)
  const double* xrows;  // row-block shard: exchanged rows [3 + 2m][ldx] (+ mu entry at column ld); null = read P
  int64_t ld, ldw, ldx;
  int n, m, row_begin, row_end;
  // peer-memory exchange (see LmGatherArgs / LmUpdateArgs): the first gather waits, the streaming kernel acks
  const unsigned long long* wait_flags;
  unsigned long long wait_seq;
  unsigned long long* wait_err;
  double* const* ack_peers;
  int64_t ack_flag_off;
  int ack_world, ack_rank;
  int j[LM_MAX_FUSED];
  double rng[LM_MAX_FUSED], brg[LM_MAX_FUSED];
  double mm;
  int full_tiles, tail_groups;  // filled by the launcher
};
cudaError_t lm_launch_update_multi(const LmMultiArgs& a, cudaStream_t st);
cudaError_t lm_launch_export_rows_multi(const LmMultiArgs& a, double* rows, cudaStream_t st);

// peer-memory row exchange of the row-block shards (ekf_largemap_p2p.cu). One exchange AREA per rank (doubles):
//   [2][P2P_ROWS][ldx] double-buffered rows {0,1,2} U {j_u, j_u+1 : u < m} (+ mu entry at column ld), then at
//   flag_off: u64 flags[2][P2P_PARTS] (buffer, part: 0 = robot rows, 1 + u = landmark rows of update u),
//   u64 acks[P2P_MAX_WORLD]
constexpr int P2P_MAX_WORLD = 16;
constexpr int P2P_ROWS = 11;   // == LM_MAX_ROWS: rows of a fused step of up to 4 updates
constexpr int P2P_PARTS = 8;   // flag words per buffer (1 + LM_MAX_FUSED used)
struct P2PArgs {
  const double* P;              // local row block
  const double* mu;
  double* const* peers;         // device array [world]: every rank's area (own entry = local pointer)
  unsigned long long* acks;     // local area's ack words
  unsigned long long* err;      // set to 1 when a spin ran out of budget
  unsigned* push_count;         // device [world] CTA counters of the push kernel
  int64_t ld, ldx, flag_off;
  unsigned long long seq;       // update sequence number, starts at 1
  int world, rank, row_begin;
};
cudaError_t lm_launch_p2p_push(const P2PArgs& a, int first_row, int first_k, int nrows, int part, cudaStream_t st);

cudaError_t lm_launch_augment(double* mu, double* P, int64_t ld, int row_begin, int row_end, int j, double gx,
                              double gy, cudaStream_t st);
cudaError_t lm_launch_predict(double* mu, double* P, int64_t ld, int n, int row_begin, int row_end,
                              double* scal, double ang, double lin, double mv, int flags, cudaStream_t st);
cudaError_t lm_launch_gather(const LmGatherArgs& a, cudaStream_t st);
cudaError_t lm_launch_update(const LmUpdateArgs& a, bool joseph, int variant, cudaStream_t st);
cudaError_t lm_launch_export_rows(const double* P, const double* mu, int64_t ld, int row_begin, int row_end,
                                  int j, double* rows5, int64_t ldx, cudaStream_t st);

}  // namespace ekf
