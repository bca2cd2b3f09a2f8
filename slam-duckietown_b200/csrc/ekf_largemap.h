// ekf_largemap.h — argument blocks and launchers of the large-map kernels (internal, not the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

constexpr int LM_THREADS = 128;  // threads per CTA of the streaming update = column groups per tile

// slots of the small scalar workspace (device doubles)
enum : int { kScalG0 = 0, kScalG1 = 1, kScalY0 = 2, kScalY1 = 3, kScalCount = 8 };

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
};

struct LmUpdateArgs {
  double* P;              // local row block
  double* mu;
  const double *K0, *K1, *A0, *A1, *W0, *W1, *scal;
  int64_t ld;
  int n, row_begin, row_end;
  int full_tiles, tail_groups;  // filled by the launcher
};

constexpr int LM_MAX_FUSED = 4;  // updates fused into one pass over P (ekf_largemap_multi.cu)

// fused multi-update: m sequential updates of one step, ONE read + write of P (plain form, whole matrix on one GPU)
struct LmMultiArgs {
  double* P;
  double* mu;
  double* K;    // [m][2][ldw]: K_u[:,0] at K + 2u ldw, K_u[:,1] at K + (2u+1) ldw
  double* W;    // [m][2][ldw]: (H P)_u rows, same indexing
  double* y;    // [m][2] innovations
  int64_t ld, ldw;
  int n, m;
  int j[LM_MAX_FUSED];
  double rng[LM_MAX_FUSED], brg[LM_MAX_FUSED];
  double mm;
  int full_tiles, tail_groups;  // filled by the launcher
};
cudaError_t lm_launch_update_multi(const LmMultiArgs& a, cudaStream_t st);

cudaError_t lm_launch_augment(double* mu, double* P, int64_t ld, int row_begin, int row_end, int j, double gx,
                              double gy, cudaStream_t st);
cudaError_t lm_launch_predict(double* mu, double* P, int64_t ld, int n, int row_begin, int row_end,
                              double* scal, double ang, double lin, double mv, int flags, bool update_mean,
                              cudaStream_t st);
cudaError_t lm_launch_gather(const LmGatherArgs& a, cudaStream_t st);
cudaError_t lm_launch_update(const LmUpdateArgs& a, bool joseph, int variant, cudaStream_t st);
cudaError_t lm_launch_export_rows(const double* P, const double* mu, int64_t ld, int row_begin, int row_end,
                                  int j, double* rows5, int64_t ldx, cudaStream_t st);

}  // namespace ekf
