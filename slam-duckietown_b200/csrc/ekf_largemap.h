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
enum : int { kScalG0 = 0, kScalG1 = 1, kScalY0 = 2, kScalY1 = 3, kScalArrive = 6 /* u64 arrival counter of the predict */, kScalReady = 7 /* u64 gather-done counter of the single-launch update */, kScalCount = 8 };

// Peer-memory row exchange, folded into the update's own kernels (round 2; round 1 launched a separate push kernel
// per row part: 14.8 us per update against a 17 us streaming pass at 8 GPUs, profiles/r1o_lm_launches.summary.txt).
//   * landmark rows {j, j+1}: the OWNER's gather kernel stores them (+ their mu entries) into every peer's buffer at
//     its head (CTA b = column chunk b) and itself reads the rows from its own P;
//   * robot rows {0, 1, 2}: rank 0's STREAMING kernel of update s stores the rows it has just brought up to date
//     (row tile 0, from registers) into the buffer of update s + 1 — the next gather finds three of its five rows
//     already there. A predict (or an external write to P) in between invalidates the pre-push; rank 0's next gather
//     then pushes the robot rows itself, like a landmark part.
// Rows travel as 16-byte cells that carry the update's sequence number next to each half of the value; consumers
// poll the cells they need (ekf_largemap_p2p.cuh): no flags, no fences. The buffer of update s is s & 1; a push into
// it first checks (warp 0 of the pushing CTA) that every rank has acknowledged the gather of update s - 2 (landmark
// part) resp. s - 1 (pre-push into the buffer of s + 1).
struct LmExchange {
  double* const* peers;          // device array [world]: every rank's exchange area; null = no exchange
  unsigned long long* acks;      // this rank's ack words [world] (local memory)
  unsigned long long* err;       // code << 32 | sequence number of a spin that ran out of budget
  int64_t ldx, flag_off;         // cells per exchanged row; offset (doubles) of the ack words in an area
  unsigned long long seq;        // sequence number of THIS update (starts at 1)
  int world, rank;
};

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
  // peer-memory exchange (x.peers != null): bit k of push_mask = this rank pushes part k (0: rows 0-2, 1: rows j, j+1)
  // at the head of the kernel; bit k of wait_mask = part k arrives from a peer: row[] / mu5[] of that part point at
  // CELLS of the exchange buffer, which are polled for this update's tag
  LmExchange x;
  int push_mask, wait_mask;
  int share_k;  // Joseph form on a row-block shard: all-gather K through the exchange buffer (rows P2P_KROW, +1)
};

struct LmUpdateArgs {
  double* P;              // local row block
  double* mu;
  const double *K0, *K1, *A0, *A1, *W0, *W1, *scal;
  int64_t ld;
  int n, row_begin, row_end;
  int full_tiles, tail_groups;  // filled by the launcher
  // peer-memory exchange (x.peers != null): the first CTA tells every rank that this rank's gather of update x.seq is
  // done; with prepush set (rank 0), row tile 0 pushes the updated rows 0-2 (+ mu) into the buffer of update seq + 1
  LmExchange x;
  int prepush;
};

constexpr int LM_MAX_FUSED = 4;  // updates fused into one pass over P (ekf_largemap_multi.cu)
constexpr int LM_MAX_ROWS = 3 + 2 * LM_MAX_FUSED;  // rows a fused step exchanges

// fused multi-update: m sequential updates of one step, ONE read + write of P (plain form, whole matrix on one GPU)
struct LmMultiArgs {
  double* P;
  double* mu;
  double* K;    // [m][2][ldw]: K_u[:,0] at K + 2u ldw, K_u[:,1] at K + (2u+1) ldw
  double* A;    // [m][2][ldw]: AHt_u, same indexing (Joseph form only)
  int joseph;   // 1: one-pass Joseph form (whole matrix on one GPU only)
  int variant;  // tiling of the streaming kernel: 0 = 32-row tiles, 1 = 16-row tiles
  double* W;    // [m][2][ldw]: (H P)_u rows, same indexing
  double* y;    // [m][2] innovations
  double* KR;   // [m][LM_MAX_ROWS][2] K_u at the exchanged rows R = {0,1,2} U {j_u, j_u+1} (replicated on every rank)
  const double* xrows;  // row-block shard: exchanged rows [3 + 2m][ldx] (+ mu entry at column ld); null = read P
  int64_t ld, ldw, ldx;
  int n, m, row_begin, row_end;
  // peer-memory exchange (see LmExchange): part 0 = rows 0-2, part 1 + u = rows j_u, j_u + 1. The fused gather pushes
  // the parts in push_mask at its head and waits for those in wait_mask; the streaming kernel acks and (prepush, rank
  // 0) pushes the updated rows 0-2 for the next step. rowp / mup: where row R[rho] of the step's initial covariance and
  // its mu entry are read from — this rank's own P / mu when it owns the row, the exchange buffer otherwise.
  LmExchange x;
  int push_mask, wait_mask, prepush;
  const double* rowp[LM_MAX_ROWS];
  const double* mup[LM_MAX_ROWS];
  int j[LM_MAX_FUSED];
  double rng[LM_MAX_FUSED], brg[LM_MAX_FUSED];
  double mm;
  int full_tiles, tail_groups;  // filled by the launcher
};
cudaError_t lm_launch_update_multi(const LmMultiArgs& a, cudaStream_t st);
cudaError_t lm_launch_export_rows_multi(const LmMultiArgs& a, double* rows, cudaStream_t st);

// peer-memory row exchange of the row-block shards (ekf_largemap_p2p.cuh). One exchange AREA per rank:
//   [2][P2P_ROWS][ldx] double-buffered rows {0,1,2} U {j_u, j_u+1 : u < m} of 16-byte cells (the mu entry rides in
//   cell ld), then at flag_off (doubles): u64 acks[P2P_MAX_WORLD]
constexpr int P2P_MAX_WORLD = 16;
constexpr int P2P_KROW = 11;   // == LM_MAX_ROWS: rows of a fused step of up to 4 updates come first ...
constexpr int P2P_ROWS = 13;   // ... then two rows of cells for K[:, 0] and K[:, 1] (Joseph form on shards: every rank
                               // needs K at every column; each rank stores its rows' K into every peer's buffer)
// in-kernel push helpers (ekf_largemap_p2p.cuh) are shared by the gather and streaming kernels

cudaError_t lm_launch_augment(double* mu, double* P, int64_t ld, int row_begin, int row_end, int j, double gx,
                              double gy, cudaStream_t st);
cudaError_t lm_launch_predict(double* mu, double* P, int64_t ld, int n, int row_begin, int row_end,
                              double* scal, double ang, double lin, double mv, int flags, cudaStream_t st);
cudaError_t lm_launch_gather(const LmGatherArgs& a, cudaStream_t st);
cudaError_t lm_launch_update(const LmUpdateArgs& a, bool joseph, int variant, cudaStream_t st);
// gather + streaming update in ONE launch (plain form); ready: device u64 counter, calls: host count of such launches
cudaError_t lm_launch_update_single(const LmGatherArgs& g, const LmUpdateArgs& u, int variant, unsigned long long* ready,
                                    unsigned long long* calls, cudaStream_t st);
cudaError_t lm_launch_export_rows(const double* P, const double* mu, int64_t ld, int row_begin, int row_end,
                                  int j, double* rows5, int64_t ldx, cudaStream_t st);

}  // namespace ekf
