// ekf_batched.h — argument block and launchers of the batched fused-step kernels (internal).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

constexpr int BATCH_MAX_M = 8;  // observations per filter-step (the surrogate bag peaks at 6)

struct BatchStepArgs {
  double* mu;                    // [B][LD]
  double* P;                     // [B][N*LD]
  int* n_known;                  // [B]
  const double* motion_var;      // [B]
  const double* meas_var;        // [B]
  const double* ang;             // [B] (or null when ticks are given)
  const double* lin;             // [B]
  const int32_t *ticks_l, *ticks_r, *prev_l, *prev_r;  // [B] each, or null
  int ticks_common;              // 1: every filter sees the SAME encoder ticks c_tl / c_tr / c_pl / c_pr (one bag)
  int32_t c_tl, c_tr, c_pl, c_pr;
  double wheel_radius, baseline;
  long long resolution;
  const int32_t* slot;           // [B][m]
  const double* rng;             // [B][m]
  const double* brg;             // [B][m]
  // Monte Carlo inputs generated IN the kernel (gen_record != null; ang / lin / slot / rng / brg are then unused): the
  // step's COMMON record (device: doubles [ang, lin, rng[0..m), brg[0..m)] then int32 slot[0..m)) plus this filter's
  // perturbation, a pure function of (gen_seed, gen_first + b, gen_step, stream) — ekf_philox.cuh. Lanes 0..m-1 draw
  // the pair of observation u = lane, lane m the odometry pair, while the TMA load of the filter's state is in flight.
  const double* gen_record;
  unsigned long long gen_seed, gen_first;
  unsigned gen_step;
  double gen_s_ang, gen_s_lin, gen_s_rng, gen_s_brg;
  double* gen_pose_out;          // GEN kernels only: compact copy [B][3] of the updated pose (null = none), so that the
                                 // per-step result (ATE, pose moments) can be formed on another stream while the next step runs
  unsigned long long* faults;    // device counter
  int m;
  int flags;
};

bool batch_capacity_supported(int n_landmarks);
size_t batch_step_smem_bytes(int n_landmarks, int m, bool joseph);
cudaError_t batch_launch_step(int n_landmarks, const BatchStepArgs& a, int64_t batch, bool joseph,
                              cudaStream_t st);
// W-in-registers variant (ekf_batched_regw.cu): plain form with 1-6 observation slots, Joseph form with 1-4
bool batch_regw_supported(int m, bool joseph, bool generated);
size_t batch_regw_smem_bytes(int n_landmarks, int m, bool joseph);
cudaError_t batch_launch_step_regw(int n_landmarks, const BatchStepArgs& a, int64_t batch, bool joseph,
                                   cudaStream_t st);
// sub-warp groups for small filters (ekf_batched_group.cu): 8 or 16 lanes per filter, 1 <= m <= 4; both forms
bool batch_group_supported(int n_landmarks, int m);
cudaError_t batch_launch_step_group(int n_landmarks, int group_width, const BatchStepArgs& a, int64_t batch,
                                    bool joseph, cudaStream_t st);
cudaError_t batch_launch_init(double* mu, double* P, int* n_known, int64_t batch, int n, int ld, double p0,
                              cudaStream_t st);
cudaError_t batch_launch_pack(double* mu, double* P, const double* mean_ref, const double* cov_ref,
                              int64_t first, int64_t count, int n, int ld, cudaStream_t st);
cudaError_t batch_launch_unpack(const double* mu, const double* P, double* mean_ref, double* cov_ref,
                                int64_t first, int64_t count, int n, int ld, cudaStream_t st);
cudaError_t batch_launch_pose(const double* mu, double* pose, int64_t batch, int ld, cudaStream_t st);

}  // namespace ekf
