// ekf_ensemble.h — device-side association and ensemble error metrics of the batched mode (internal).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

constexpr int ASSOC_MAX_M = 8;  // == BATCH_MAX_M: observation slots of one filter-step

struct AssocArgs {
  int64_t batch;
  int n_det, n_tags, m;
  const int32_t* det_tag;   // [n_det] compact tag index (host maps tag ids; < 0 = ignored tag)
  const double* det_tx;     // [n_det] pose_t[0] of each detection of the window, in window order
  const double* det_tz;     // [n_det] pose_t[2]
  const double* noise_tx;   // [n_det][B] per-filter perturbation added to pose_t[0] (null = none)
  const double* noise_tz;   // [n_det][B]
  // gen != 0: no noise arrays — detection d of filter b is perturbed by gen_sigma x the normal pair of
  // (gen_seed, gen_first + b, gen_step, kStreamDetection0 + d), generated in the kernel (ekf_philox.cuh)
  int gen;
  uint64_t gen_seed, gen_first;
  uint32_t gen_step;
  double gen_sigma;
  int32_t* tag_slot;        // [n_tags][B] slot of a tag in this filter's TAG_INDEX, -1 = never seen
  int32_t* n_seen;          // [B] len(TAG_INDEX)
  int32_t* slot;            // out [B][m], -1 = unused
  double* rng;              // out [B][m]
  double* brg;              // out [B][m]
  unsigned long long* overflow;  // windows that held more than m distinct landmarks
};

cudaError_t batch_launch_associate(const AssocArgs& a, cudaStream_t st);

// Monte Carlo inputs generated on the device (ekf_philox.cuh): every filter gets the step's COMMON odometry /
// observations plus its own perturbation, a pure function of (seed, first_filter + b, step, stream).
// record (device): doubles [ang, lin, rng[0..m), brg[0..m)] followed by int32 slot[0..m)
struct PerturbArgs {
  int64_t batch;
  uint64_t seed, first_filter;
  uint32_t step;
  int m;
  const double* record;
  double s_ang, s_lin, s_rng, s_brg;  // standard deviations of the four perturbations
  double *ang, *lin;                  // out [B]
  int32_t* slot;                      // out [B][m]
  double *rng, *brg;                  // out [B][m]
};
cudaError_t batch_launch_perturb(const PerturbArgs& a, cudaStream_t st);
// noise_tx / noise_tz [n_det][B] = sigma * the normal pair of stream kStreamDetection0 + d (per-filter detection noise)
cudaError_t batch_launch_detection_noise(int64_t batch, uint64_t seed, uint64_t first_filter, uint32_t step, int n_det,
                                         double sigma, double* noise_tx, double* noise_tz, cudaStream_t st);
// ensemble moments of the robot pose only: sums[0..2] = sum_b (x, y, theta), sums[3 + 3 i + k] = sum_b pose_i pose_k;
// two passes, fixed summation order; partial: scratch [POSE_MOM_CTAS][12]
constexpr int POSE_MOM_CTAS = 592;  // 4 x 148 SMs
cudaError_t batch_launch_pose_moments(const double* mu, int64_t mu_stride, int64_t batch, double* partial, double* sums,
                                      cudaStream_t st);
// the same result with the partials in a file-scope device array (no allocation anywhere on the path); at most
// POSE_MOM_STATIC_MAX_FILTERS filters per call; calls on one device share the array and must not overlap
constexpr int64_t POSE_MOM_STATIC_MAX_FILTERS = int64_t(1) << 22;
cudaError_t batch_launch_pose_moments_static(const double* mu, int64_t mu_stride, int64_t batch, double* sums,
                                             int device, cudaStream_t st);
constexpr int MOM_CTAS = 592;  // 4 x 148 SMs: partial sums of the ensemble moments
// sums[n + n*n] = (sum_b mu_b, sum_b mu_b mu_b^T) in a fixed order; partial: scratch [MOM_CTAS][n + n*n]
cudaError_t batch_launch_moments(const double* mu, int64_t mu_stride, int64_t batch, int n, double* partial,
                                 double* sums, cudaStream_t st);
cudaError_t batch_launch_nonfinite(const double* mu, const double* P, int64_t mu_stride, int64_t p_stride,
                                   int64_t batch, unsigned long long* count, cudaStream_t st);
cudaError_t batch_launch_ate(const double* mu, int64_t mu_stride, int64_t batch, double gx, double gy, double* sse,
                             cudaStream_t st);

}  // namespace ekf
