// ekf_capi.cu — the C-ABI of libekf_b200.so (see include/ekf_b200.h for the contract and the reference
// interfaces each entry point replaces). Plain pointers and sizes only; no torch types; errors are returned
// as negative codes with a thread-local message.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "../../include/ekf_b200.h"
#include "ekf3.h"
#include "ekf_batched.h"
#include "ekf_common.cuh"
#include "ekf_largemap.h"
#include "ekf_ensemble.h"

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) return fail(EKF_ECUDA, "%s: %s", #call, cudaGetErrorString(e_));   \
  } while (0)

int make_flags(int circular, int measurement, int disable_motion, int cov_form) {
  return (circular ? ekf::kFlagCircular : 0) | (measurement ? ekf::kFlagMeasurement : 0) |
         (disable_motion ? ekf::kFlagDisableMotion : 0) | (cov_form == EKF_COV_JOSEPH ? ekf::kFlagJoseph : 0);
}

}  // namespace

struct ekf_batch {
  int nlm, n, ld, device;
  int64_t batch;
  double *mu = nullptr, *P = nullptr;
  int32_t* n_known = nullptr;
  const double *mv = nullptr, *mm = nullptr;
  int flags = ekf::kFlagCircular | ekf::kFlagMeasurement;
  // 0 = auto (W-in-registers kernel where it applies: plain form, 1-6 observation slots; else fully staged),
  // 2 = fully staged, 3 = warp per filter with W in registers even for small filters, 8 / 16 = sub-warp groups of
  // that width (capacity <= 8 landmarks; auto picks 8 there); EKF_BATCH_KERNEL=auto|staged|regw|group8|group16
  int kernel = 0;
  unsigned long long* faults = nullptr;  // [0] fault count, [1..2] work / done counters of the persistent kernel, [3] association overflows, [4] scratch of ekf_batch_nonfinite_count
  // staging for the host-buffer entry point
  cudaStream_t stream = nullptr;
  double *d_ang = nullptr, *d_lin = nullptr, *d_rng = nullptr, *d_brg = nullptr, *d_pose = nullptr;
  int32_t* d_slot = nullptr;
  // pipelined host entry point: copy-in / compute / copy-out streams over two staging sets
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
  double* p_stage[2] = {nullptr, nullptr};  // per set: ang | lin | rng[M] | brg[M] | pose[3] doubles, slot[M] ints
  int64_t pipe_step = 0;
  cudaEvent_t ev_caller = nullptr;  // orders the handle's streams after the caller's stream
  double* mom_partial = nullptr;  // scratch of ekf_batch_moments
  double* pose_partial = nullptr;  // scratch of ekf_batch_pose_moments
  double* mc_moments[2] = {nullptr, nullptr};  // pose moments of the step in flight on each staging set
  double* mc_record[2] = {nullptr, nullptr};  // device copies of the common step record (one per staging set)
};

struct ekf_lm {
  int cap, n, device;
  int64_t ld = 0, ldw = 0;
  double *mu = nullptr, *P = nullptr;
  int row_begin = 0, row_end = 0;
  double mv = 0.1, mm = 0.7;
  int flags = ekf::kFlagCircular | ekf::kFlagMeasurement;
  double *K0 = nullptr, *K1 = nullptr, *A0 = nullptr, *A1 = nullptr, *W0 = nullptr, *W1 = nullptr,
         *scal = nullptr;
  double *mK = nullptr, *mW = nullptr, *mA = nullptr, *my = nullptr;  // fused multi-update records [LM_MAX_FUSED][2][ldw], y [..][2]
  double* mKR = nullptr;                               // K_u at the exchanged rows [LM_MAX_FUSED][LM_MAX_ROWS][2]
  // peer-memory row exchange (ekf_lm_p2p_*)
  int p2p_world = 0, p2p_rank = 0;
  double* p2p_area = nullptr;                 // this rank's exchange area (cudaMalloc, exported by IPC handle)
  double* p2p_peer_host[ekf::P2P_MAX_WORLD] = {};
  bool p2p_opened[ekf::P2P_MAX_WORLD] = {};
  double** p2p_peers_dev = nullptr;
  unsigned long long* p2p_err = nullptr;
  unsigned long long p2p_seq = 0;
  unsigned long long p2p_robot_seq = 0;  // update number whose buffer already holds rows 0-2 (pre-pushed by rank 0)
  int variant = 0;          // tiling of the streaming kernel (EKF_LM_VARIANT; default chosen at bind time, see there)
  bool variant_auto = true;
  bool single_launch = true;             // gather + streaming update of one update in one launch (EKF_LM_SINGLE=0: two)
  unsigned long long single_calls = 0;   // single-launch updates issued so far (the ready counter's target grows with it)
};

extern "C" {

int ekf_version(void) { return 100; }
const char* ekf_last_error(void) { return g_err.c_str(); }

int ekf_device_count(int* count) {
  if (!count) return fail(EKF_EINVAL, "count is null");
  CU(cudaGetDeviceCount(count));
  return EKF_OK;
}

double ekf_delta_phi(int64_t ticks, int64_t prev_ticks, int64_t resolution) {
  return ekf::delta_phi(ticks, prev_ticks, resolution);
}

int ekf_displacement(double wheel_radius, double baseline, double dphi_left, double dphi_right, double* angular,
                     double* linear) {
  if (!angular || !linear) return fail(EKF_EINVAL, "output pointer is null");
  ekf::displacement(wheel_radius, baseline, dphi_left, dphi_right, angular, linear);
  return EKF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// batched
// ---------------------------------------------------------------------------------------------------------
int ekf_batch_layout(int n_landmarks, int* n, int* ld, int64_t* p_stride, int64_t* mu_stride) {
  if (n_landmarks < 0) return fail(EKF_EINVAL, "n_landmarks < 0");
  const int nn = 3 + 2 * n_landmarks, l = (nn + 3) & ~3;
  if (n) *n = nn;
  if (ld) *ld = l;
  if (p_stride) *p_stride = (int64_t)nn * l;
  if (mu_stride) *mu_stride = l;
  return EKF_OK;
}

int ekf_batch_create(int n_landmarks, int64_t batch, int device, ekf_batch_t** out) {
  if (!out) return fail(EKF_EINVAL, "out is null");
  if (!ekf::batch_capacity_supported(n_landmarks))
    return fail(EKF_EINVAL, "landmark capacity %d not built (use 4, 8, ..., 28, 32, 40, 48 or 56)", n_landmarks);
  if (batch <= 0 || batch > 0x7fffffffLL) return fail(EKF_EINVAL, "batch out of range");
  CU(cudaSetDevice(device));
  ekf_batch* h = new (std::nothrow) ekf_batch();
  if (!h) return fail(EKF_EINVAL, "out of host memory");
  h->nlm = n_landmarks;
  h->n = 3 + 2 * n_landmarks;
  h->ld = (h->n + 3) & ~3;
  h->device = device;
  h->batch = batch;
  if (const char* k = getenv("EKF_BATCH_KERNEL"))
    h->kernel = (strcmp(k, "staged") == 0) ? 2 : (strcmp(k, "regw") == 0) ? 3 :
                (strcmp(k, "group8") == 0) ? 8 : (strcmp(k, "group16") == 0) ? 16 : 0;
  cudaError_t e = cudaMalloc(&h->faults, 8 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(h->faults, 0, 8 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    ekf_batch_destroy(h);
    return fail(EKF_ECUDA, "ekf_batch_create: %s", cudaGetErrorString(e));
  }
  *out = h;
  return EKF_OK;
}

int ekf_batch_destroy(ekf_batch_t* h) {
  if (!h) return EKF_OK;
  cudaSetDevice(h->device);
  cudaFree(h->faults);
  cudaFree(h->mom_partial);
  cudaFree(h->pose_partial);
  cudaFree(h->mc_moments[0]);
  cudaFree(h->mc_moments[1]);
  cudaFree(h->mc_record[0]);
  cudaFree(h->mc_record[1]);
  cudaFree(h->d_ang);
  cudaFree(h->d_lin);
  cudaFree(h->d_rng);
  cudaFree(h->d_brg);
  cudaFree(h->d_slot);
  cudaFree(h->d_pose);
  for (int i = 0; i < 2; ++i) {
    cudaFree(h->p_stage[i]);
    if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
    if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]);
    if (h->ev_out[i]) cudaEventDestroy(h->ev_out[i]);
  }
  if (h->ev_caller) cudaEventDestroy(h->ev_caller);
  if (h->s_in) cudaStreamDestroy(h->s_in);
  if (h->s_out) cudaStreamDestroy(h->s_out);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return EKF_OK;
}

int ekf_batch_bind_state(ekf_batch_t* h, double* mu_dev, double* p_dev, int32_t* n_known_dev) {
  if (!h || !mu_dev || !p_dev || !n_known_dev) return fail(EKF_EINVAL, "null argument");
  if (((uintptr_t)mu_dev & 31) || ((uintptr_t)p_dev & 31))
    return fail(EKF_EINVAL, "mu/P must be 32-byte aligned (256-bit accesses, TMA bulk copies)");
  h->mu = mu_dev;
  h->P = p_dev;
  h->n_known = n_known_dev;
  return EKF_OK;
}

int ekf_batch_set_noise(ekf_batch_t* h, const double* motion_var_dev, const double* meas_var_dev) {
  if (!h || !motion_var_dev || !meas_var_dev) return fail(EKF_EINVAL, "null argument");
  h->mv = motion_var_dev;
  h->mm = meas_var_dev;
  return EKF_OK;
}

int ekf_batch_set_flags(ekf_batch_t* h, int circular, int measurement, int disable_motion, int cov_form) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (cov_form != EKF_COV_PLAIN && cov_form != EKF_COV_JOSEPH) return fail(EKF_EINVAL, "unknown cov_form");
  h->flags = make_flags(circular, measurement, disable_motion, cov_form);
  return EKF_OK;
}

int ekf_batch_init(ekf_batch_t* h, double p0_diag, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  CU(ekf::batch_launch_init(h->mu, h->P, h->n_known, h->batch, h->n, h->ld, p0_diag, (cudaStream_t)stream));
  CU(cudaMemsetAsync(h->faults, 0, sizeof(unsigned long long), (cudaStream_t)stream));
  return EKF_OK;
}

static int batch_step_common(ekf_batch_t* h, ekf::BatchStepArgs& a, int m, void* stream) {
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  if (!h->mv) return fail(EKF_ESTATE, "noise not set");
  if (m < 0 || m > ekf::BATCH_MAX_M) return fail(EKF_EINVAL, "m must be in [0, %d]", ekf::BATCH_MAX_M);
  if (m > 0 && !a.gen_record && (!a.slot || !a.rng || !a.brg)) return fail(EKF_EINVAL, "observation arrays are null");
  a.mu = h->mu;
  a.P = h->P;
  a.n_known = h->n_known;
  a.motion_var = h->mv;
  a.meas_var = h->mm;
  a.faults = h->faults;
  a.m = m;
  a.flags = h->flags;
  CU(cudaSetDevice(h->device));
  const bool joseph = (h->flags & ekf::kFlagJoseph) != 0;
  // dispatch (both covariance forms): capacities <= 8 landmarks -> sub-warp group kernel; otherwise one warp per filter
  // with W in registers; the fully staged kernel takes what is left (Joseph with 5+ slots, 7-8 slots, EKF_BATCH_KERNEL)
  const bool gen = a.gen_record != nullptr;  // in-kernel inputs: plain form on the fast kernels, else the staged kernel
  if ((h->kernel == 0 || h->kernel >= 8) && ekf::batch_group_supported(h->nlm, m) && !(gen && joseph))
    CU(ekf::batch_launch_step_group(h->nlm, h->kernel >= 8 ? h->kernel : 8, a, h->batch, joseph, (cudaStream_t)stream));
  else if ((h->kernel == 0 || h->kernel == 3 || h->kernel >= 8) && ekf::batch_regw_supported(m, joseph, gen) && h->nlm <= 28)
    CU(ekf::batch_launch_step_regw(h->nlm, a, h->batch, joseph, (cudaStream_t)stream));
  else
    CU(ekf::batch_launch_step(h->nlm, a, h->batch, joseph, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_step(ekf_batch_t* h, const double* ang_dev, const double* lin_dev, const int32_t* slot_dev,
                   const double* rng_dev, const double* brg_dev, int m, void* stream) {
  if (!h || !ang_dev || !lin_dev) return fail(EKF_EINVAL, "null argument");
  ekf::BatchStepArgs a{};
  a.ang = ang_dev;
  a.lin = lin_dev;
  a.slot = slot_dev;
  a.rng = rng_dev;
  a.brg = brg_dev;
  return batch_step_common(h, a, m, stream);
}

int ekf_batch_step_ticks(ekf_batch_t* h, const int32_t* ticks_left_dev, const int32_t* ticks_right_dev,
                         const int32_t* prev_left_dev, const int32_t* prev_right_dev, double wheel_radius,
                         double baseline, int resolution, const int32_t* slot_dev, const double* rng_dev,
                         const double* brg_dev, int m, void* stream) {
  if (!h || !ticks_left_dev || !ticks_right_dev || !prev_left_dev || !prev_right_dev)
    return fail(EKF_EINVAL, "null argument");
  if (resolution == 0 || baseline == 0.0) return fail(EKF_EINVAL, "resolution / baseline must be non-zero");
  ekf::BatchStepArgs a{};
  a.ticks_l = ticks_left_dev;
  a.ticks_r = ticks_right_dev;
  a.prev_l = prev_left_dev;
  a.prev_r = prev_right_dev;
  a.wheel_radius = wheel_radius;
  a.baseline = baseline;
  a.resolution = resolution;
  a.slot = slot_dev;
  a.rng = rng_dev;
  a.brg = brg_dev;
  return batch_step_common(h, a, m, stream);
}

int ekf_batch_step_ticks_common(ekf_batch_t* h, int32_t ticks_left, int32_t ticks_right, int32_t prev_left,
                                int32_t prev_right, double wheel_radius, double baseline, int resolution,
                                const int32_t* slot_dev, const double* rng_dev, const double* brg_dev, int m,
                                void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (resolution == 0 || baseline == 0.0) return fail(EKF_EINVAL, "resolution / baseline must be non-zero");
  ekf::BatchStepArgs a{};
  a.ticks_common = 1;
  a.c_tl = ticks_left; a.c_tr = ticks_right; a.c_pl = prev_left; a.c_pr = prev_right;
  a.wheel_radius = wheel_radius;
  a.baseline = baseline;
  a.resolution = resolution;
  a.slot = slot_dev;
  a.rng = rng_dev;
  a.brg = brg_dev;
  return batch_step_common(h, a, m, stream);
}

static int batch_step_generated(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                                const double* record_dev, const double* sigma4, double* pose_out, void* stream) {
  if (!record_dev || !sigma4) return fail(EKF_EINVAL, "null argument");
  if (step > 0xffffffffull) return fail(EKF_EINVAL, "step must fit 32 bits");
  ekf::BatchStepArgs a{};
  a.gen_record = record_dev;
  a.gen_seed = seed; a.gen_first = first_filter; a.gen_step = (unsigned)step;
  a.gen_s_ang = sigma4[0]; a.gen_s_lin = sigma4[1]; a.gen_s_rng = sigma4[2]; a.gen_s_brg = sigma4[3];
  a.gen_pose_out = pose_out;
  return batch_step_common(h, a, m, stream);
}

int ekf_batch_step_generated(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                             const double* record_dev, const double* sigma4, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  return batch_step_generated(h, seed, step, first_filter, m, record_dev, sigma4, nullptr, stream);
}

int ekf_batch_pack(ekf_batch_t* h, const double* mean_ref_dev, const double* cov_ref_dev, int64_t first,
                   int64_t count, void* stream) {
  if (!h || !mean_ref_dev || !cov_ref_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  if (first < 0 || count < 0 || first + count > h->batch) return fail(EKF_EINVAL, "filter range out of bounds");
  CU(cudaSetDevice(h->device));
  if (count)
    CU(ekf::batch_launch_pack(h->mu, h->P, mean_ref_dev, cov_ref_dev, first, count, h->n, h->ld,
                              (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_unpack(ekf_batch_t* h, double* mean_ref_dev, double* cov_ref_dev, int64_t first, int64_t count,
                     void* stream) {
  if (!h || !mean_ref_dev || !cov_ref_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  if (first < 0 || count < 0 || first + count > h->batch) return fail(EKF_EINVAL, "filter range out of bounds");
  CU(cudaSetDevice(h->device));
  if (count)
    CU(ekf::batch_launch_unpack(h->mu, h->P, mean_ref_dev, cov_ref_dev, first, count, h->n, h->ld,
                                (cudaStream_t)stream));
  return EKF_OK;
}

// Stream contract of the host entry points: the fused step runs on the handle's own non-blocking streams (copy-in /
// compute / copy-out). `stream` is the caller's stream that last touched the bound state (init, step, pack, ...):
// on entry the handle's streams wait for everything enqueued on it so far, on exit it is made to wait for the
// step's kernel, so that a later ekf_batch_step / unpack / moments on `stream` sees the state this call produced.
static int host_order_in(ekf_batch_t* h, cudaStream_t caller, cudaStream_t first, cudaStream_t second) {
  if (!h->ev_caller) CU(cudaEventCreateWithFlags(&h->ev_caller, cudaEventDisableTiming));
  CU(cudaEventRecord(h->ev_caller, caller));
  CU(cudaStreamWaitEvent(first, h->ev_caller, 0));
  if (second) CU(cudaStreamWaitEvent(second, h->ev_caller, 0));
  return EKF_OK;
}

int ekf_batch_step_host(ekf_batch_t* h, const double* ang_host, const double* lin_host,
                        const int32_t* slot_host, const double* rng_host, const double* brg_host, int m,
                        double* pose_host, void* stream) {
  if (!h || !ang_host || !lin_host || !pose_host) return fail(EKF_EINVAL, "null argument");
  if (m < 0 || m > ekf::BATCH_MAX_M) return fail(EKF_EINVAL, "m must be in [0, %d]", ekf::BATCH_MAX_M);
  if (m > 0 && (!slot_host || !rng_host || !brg_host)) return fail(EKF_EINVAL, "observation arrays are null");
  CU(cudaSetDevice(h->device));
  const int64_t B = h->batch;
  if (!h->d_ang) {
    CU(cudaMalloc(&h->d_ang, B * sizeof(double)));
    CU(cudaMalloc(&h->d_lin, B * sizeof(double)));
    CU(cudaMalloc(&h->d_rng, B * ekf::BATCH_MAX_M * sizeof(double)));
    CU(cudaMalloc(&h->d_brg, B * ekf::BATCH_MAX_M * sizeof(double)));
    CU(cudaMalloc(&h->d_slot, B * ekf::BATCH_MAX_M * sizeof(int32_t)));
    CU(cudaMalloc(&h->d_pose, B * 3 * sizeof(double)));
  }
  cudaStream_t st = h->stream;
  if (h->s_in) {  // an earlier pipelined call may still be in flight on the other two streams
    CU(cudaStreamSynchronize(h->s_in));
    CU(cudaStreamSynchronize(h->s_out));
  }
  int rc = host_order_in(h, (cudaStream_t)stream, st, nullptr);
  if (rc != EKF_OK) return rc;
  CU(cudaMemcpyAsync(h->d_ang, ang_host, B * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(h->d_lin, lin_host, B * sizeof(double), cudaMemcpyHostToDevice, st));
  if (m > 0) {
    CU(cudaMemcpyAsync(h->d_slot, slot_host, B * m * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->d_rng, rng_host, B * m * sizeof(double), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->d_brg, brg_host, B * m * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  rc = ekf_batch_step(h, h->d_ang, h->d_lin, h->d_slot, h->d_rng, h->d_brg, m, st);
  if (rc != EKF_OK) return rc;
  CU(ekf::batch_launch_pose(h->mu, h->d_pose, B, h->ld, st));
  CU(cudaMemcpyAsync(pose_host, h->d_pose, B * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));  // synchronous: the caller's stream needs no further ordering
  return EKF_OK;
}

// The pipeline behind both pipelined host entry points. Step i: `fill` puts the step's per-filter inputs into staging
// set i & 1 on the copy-in stream (H2D copies, or one small copy + the on-device perturbation generator); the fused
// step then runs on the compute stream; `post` turns the new state into the step's result in the set's result area
// (every filter's pose, or the ensemble's pose moments) on the compute stream; the result drains to `out_host` on the
// copy-out stream. Two staging sets ordered by events: fill(i) overlaps step(i-1) and drain(i-2).
}  // extern "C" (templates cannot have C linkage)

struct PipeSet {
  double *ang, *lin, *rng, *brg, *result;
  int32_t* slot;
};

// `drain` runs on the copy-out stream before the device-to-host copy (work on the step's compact result that may
// overlap the NEXT step's kernel); out_dev: where the bytes copied to out_host live (null = the set's result area)
template <typename Fill, typename Step, typename Post, typename Drain>
static int pipe_step(ekf_batch_t* h, void* caller, void* out_host, size_t out_bytes, const void* out_dev, Fill fill,
                     Step step, Post post, Drain drain) {
  CU(cudaSetDevice(h->device));
  const int64_t B = h->batch;
  const int M = ekf::BATCH_MAX_M;
  const size_t set_doubles = (size_t)B * (2 + 2 * M + 3) + (size_t)(B * M + 1) / 2 + 4;
  if (!h->s_in) {
    CU(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CU(cudaMalloc(&h->p_stage[i], set_doubles * sizeof(double)));
      CU(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
    }
  }
  const int set = (int)(h->pipe_step & 1);
  PipeSet ps;
  ps.ang = h->p_stage[set];
  ps.lin = ps.ang + B;
  ps.rng = ps.lin + B;
  ps.brg = ps.rng + B * M;
  ps.result = ps.brg + B * M;  // [B][3] doubles
  ps.slot = reinterpret_cast<int32_t*>(ps.result + B * 3);
  // the state may have been touched on the caller's stream (init / step / pack) since the last host step
  int rc = host_order_in(h, (cudaStream_t)caller, h->stream, nullptr);
  if (rc != EKF_OK) return rc;
  // copy-in: the set is free once the kernel of step i-2 (its last reader) is done
  if (h->pipe_step >= 2) CU(cudaStreamWaitEvent(h->s_in, h->ev_done[set], 0));
  rc = fill(ps, h->s_in);
  if (rc != EKF_OK) return rc;
  CU(cudaEventRecord(h->ev_in[set], h->s_in));
  // compute: after this step's inputs have landed and the result area of step i-2 has drained
  CU(cudaStreamWaitEvent(h->stream, h->ev_in[set], 0));
  if (h->pipe_step >= 2) CU(cudaStreamWaitEvent(h->stream, h->ev_out[set], 0));
  rc = step(ps, h->stream);
  if (rc != EKF_OK) return rc;
  rc = post(ps, h->stream);
  if (rc != EKF_OK) return rc;
  CU(cudaEventRecord(h->ev_done[set], h->stream));
  // the caller's stream sees the state of this step from here on
  CU(cudaStreamWaitEvent((cudaStream_t)caller, h->ev_done[set], 0));
  // copy-out
  CU(cudaStreamWaitEvent(h->s_out, h->ev_done[set], 0));
  rc = drain(ps, h->s_out);
  if (rc != EKF_OK) return rc;
  if (out_host && out_bytes)
    CU(cudaMemcpyAsync(out_host, out_dev ? out_dev : ps.result, out_bytes, cudaMemcpyDeviceToHost, h->s_out));
  CU(cudaEventRecord(h->ev_out[set], h->s_out));
  ++h->pipe_step;
  return EKF_OK;
}

extern "C" {

// Pipelined variant of ekf_batch_step_host: returns as soon as the work is enqueued. Host buffers must stay valid,
// and pose_host must not be read, until ekf_batch_host_sync returns.
int ekf_batch_step_host_async(ekf_batch_t* h, const double* ang_host, const double* lin_host,
                              const int32_t* slot_host, const double* rng_host, const double* brg_host, int m,
                              double* pose_host, void* stream) {
  if (!h || !ang_host || !lin_host || !pose_host) return fail(EKF_EINVAL, "null argument");
  if (m < 0 || m > ekf::BATCH_MAX_M) return fail(EKF_EINVAL, "m must be in [0, %d]", ekf::BATCH_MAX_M);
  if (m > 0 && (!slot_host || !rng_host || !brg_host)) return fail(EKF_EINVAL, "observation arrays are null");
  const int64_t B = h->batch;
  return pipe_step(
      h, stream, pose_host, (size_t)B * 3 * sizeof(double), nullptr,
      [&](const PipeSet& ps, cudaStream_t st) -> int {
        CU(cudaMemcpyAsync(ps.ang, ang_host, B * sizeof(double), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ps.lin, lin_host, B * sizeof(double), cudaMemcpyHostToDevice, st));
        if (m > 0) {
          CU(cudaMemcpyAsync(ps.slot, slot_host, B * m * sizeof(int32_t), cudaMemcpyHostToDevice, st));
          CU(cudaMemcpyAsync(ps.rng, rng_host, B * m * sizeof(double), cudaMemcpyHostToDevice, st));
          CU(cudaMemcpyAsync(ps.brg, brg_host, B * m * sizeof(double), cudaMemcpyHostToDevice, st));
        }
        return EKF_OK;
      },
      [&](const PipeSet& ps, cudaStream_t st) -> int {
        return ekf_batch_step(h, ps.ang, ps.lin, ps.slot, ps.rng, ps.brg, m, st);
      },
      [&](const PipeSet& ps, cudaStream_t st) -> int {
        CU(ekf::batch_launch_pose(h->mu, ps.result, B, h->ld, st));
        return EKF_OK;
      },
      [&](const PipeSet&, cudaStream_t) -> int { return EKF_OK; });
}

// ---- Monte Carlo replay with the perturbations generated on the device --------------------------------------
int ekf_batch_mc_record_doubles(int m, int64_t* doubles) {
  if (m < 0 || m > ekf::BATCH_MAX_M || !doubles) return fail(EKF_EINVAL, "m out of range or null argument");
  *doubles = 2 + 2 * m + (m + 1) / 2;
  return EKF_OK;
}

static int mc_fill_args(ekf_batch_t* h, ekf::PerturbArgs& a, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                        const double* sigma4) {
  if (m < 0 || m > ekf::BATCH_MAX_M) return fail(EKF_EINVAL, "m must be in [0, %d]", ekf::BATCH_MAX_M);
  if (!sigma4) return fail(EKF_EINVAL, "sigma4 is null");
  if (step > 0xffffffffull) return fail(EKF_EINVAL, "step must fit 32 bits");
  a.batch = h->batch; a.seed = seed; a.first_filter = first_filter; a.step = (uint32_t)step; a.m = m;
  a.s_ang = sigma4[0]; a.s_lin = sigma4[1]; a.s_rng = sigma4[2]; a.s_brg = sigma4[3];
  return EKF_OK;
}

int ekf_batch_perturb(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                      const double* record_dev, const double* sigma4, double* ang_dev, double* lin_dev,
                      int32_t* slot_dev, double* rng_dev, double* brg_dev, void* stream) {
  if (!h || !record_dev || !ang_dev || !lin_dev) return fail(EKF_EINVAL, "null argument");
  if (m > 0 && (!slot_dev || !rng_dev || !brg_dev)) return fail(EKF_EINVAL, "observation arrays are null");
  ekf::PerturbArgs a{};
  const int rc = mc_fill_args(h, a, seed, step, first_filter, m, sigma4);
  if (rc != EKF_OK) return rc;
  a.record = record_dev; a.ang = ang_dev; a.lin = lin_dev; a.slot = slot_dev; a.rng = rng_dev; a.brg = brg_dev;
  CU(cudaSetDevice(h->device));
  CU(ekf::batch_launch_perturb(a, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_detection_noise(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int n_det,
                              double sigma, double* noise_tx_dev, double* noise_tz_dev, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (n_det < 0 || step > 0xffffffffull) return fail(EKF_EINVAL, "n_det < 0 or step beyond 32 bits");
  if (n_det > 0 && (!noise_tx_dev || !noise_tz_dev)) return fail(EKF_EINVAL, "null argument");
  CU(cudaSetDevice(h->device));
  CU(ekf::batch_launch_detection_noise(h->batch, seed, first_filter, (uint32_t)step, n_det, sigma, noise_tx_dev,
                                       noise_tz_dev, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_pose_moments(ekf_batch_t* h, double* sums_dev, void* stream) {
  if (!h || !sums_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  if (!h->pose_partial) CU(cudaMalloc(&h->pose_partial, sizeof(double) * ekf::POSE_MOM_CTAS * 9));
  CU(ekf::batch_launch_pose_moments(h->mu, h->ld, h->batch, h->pose_partial, sums_dev, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_pose_moments_noalloc(const double* mu_dev, int64_t mu_stride, int64_t batch, double* sums_dev, int device,
                             void* stream) {
  if (!mu_dev || !sums_dev) return fail(EKF_EINVAL, "null argument");
  if (mu_stride < 3) return fail(EKF_EINVAL, "mu_stride < 3 (a record starts with x, y, theta)");
  if (batch < 1) return fail(EKF_EINVAL, "batch < 1");
  if (batch > ekf::POSE_MOM_STATIC_MAX_FILTERS)
    return fail(EKF_EINVAL, "ekf_pose_moments_noalloc: batch exceeds EKF_POSE_MOMENTS_NOALLOC_MAX_FILTERS (4194304); "
                            "use ekf_batch_pose_moments or split the call");
  CU(ekf::batch_launch_pose_moments_static(mu_dev, mu_stride, batch, sums_dev, device, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_mc_step_host_async(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                                 const double* record_host, const double* sigma4, int score, double gt_x, double gt_y,
                                 double* sse_dev, double* pose_moments_host, void* stream) {
  if (!h || !record_host) return fail(EKF_EINVAL, "null argument");
  if (score && !sse_dev) return fail(EKF_EINVAL, "score requested without an sse buffer");
  ekf::PerturbArgs a{};
  const int rc0 = mc_fill_args(h, a, seed, step, first_filter, m, sigma4);
  if (rc0 != EKF_OK) return rc0;
  CU(cudaSetDevice(h->device));
  if (!h->pose_partial) CU(cudaMalloc(&h->pose_partial, sizeof(double) * ekf::POSE_MOM_CTAS * 9));
  if (!h->mc_record[0])
    for (int i = 0; i < 2; ++i) CU(cudaMalloc(&h->mc_record[i], sizeof(double) * 32));
  const int set = (int)(h->pipe_step & 1);
  const size_t rec_bytes = sizeof(double) * (size_t)(2 + 2 * m + (m + 1) / 2);
  // the per-member inputs are drawn INSIDE the fused step (BatchStepArgs::gen_record): the copy-in stream only
  // carries the < 200-byte record; no perturbation kernel, no per-filter staging arrays. The step kernel leaves a
  // compact copy of every member's pose in the set's result area; ATE and the pose moments are formed from it on the
  // copy-out stream, i.e. while the NEXT step already runs.
  if (!h->mc_moments[0])
    for (int i = 0; i < 2; ++i) CU(cudaMalloc(&h->mc_moments[i], sizeof(double) * 12));
  return pipe_step(
      h, stream, pose_moments_host, pose_moments_host ? 12 * sizeof(double) : 0, h->mc_moments[set],
      [&](const PipeSet&, cudaStream_t st) -> int {
        CU(cudaMemcpyAsync(h->mc_record[set], record_host, rec_bytes, cudaMemcpyHostToDevice, st));
        return EKF_OK;
      },
      [&](const PipeSet& ps, cudaStream_t st) -> int {
        return batch_step_generated(h, seed, step, first_filter, m, h->mc_record[set], sigma4, ps.result, st);
      },
      [&](const PipeSet&, cudaStream_t) -> int { return EKF_OK; },
      [&](const PipeSet& ps, cudaStream_t st) -> int {
        if (score) CU(ekf::batch_launch_ate(ps.result, 3, h->batch, gt_x, gt_y, sse_dev, st));
        if (pose_moments_host)
          CU(ekf::batch_launch_pose_moments(ps.result, 3, h->batch, h->pose_partial, h->mc_moments[set], st));
        return EKF_OK;
      });
}

int ekf_batch_host_sync(ekf_batch_t* h) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  CU(cudaSetDevice(h->device));
  if (h->s_in) {
    CU(cudaStreamSynchronize(h->s_in));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaStreamSynchronize(h->s_out));
  } else if (h->stream) {
    CU(cudaStreamSynchronize(h->stream));
  }
  return EKF_OK;
}

int ekf_batch_fault_count(ekf_batch_t* h, int64_t* count) {
  if (!h || !count) return fail(EKF_EINVAL, "null argument");
  CU(cudaSetDevice(h->device));
  unsigned long long v = 0;
  CU(cudaMemcpy(&v, h->faults, sizeof(v), cudaMemcpyDeviceToHost));
  *count = (int64_t)v;
  return EKF_OK;
}

int ekf_batch_nonfinite_count(ekf_batch_t* h, int64_t* count, void* stream) {
  if (!h || !count) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaMemsetAsync(h->faults + 4, 0, sizeof(unsigned long long), st));
  CU(ekf::batch_launch_nonfinite(h->mu, h->P, h->ld, (int64_t)h->n * h->ld, h->batch, h->faults + 4, st));
  unsigned long long v = 0;
  CU(cudaMemcpyAsync(&v, h->faults + 4, sizeof(v), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  *count = (int64_t)v;
  return EKF_OK;
}

int ekf_batch_associate(ekf_batch_t* h, int n_det, const int32_t* det_tag_dev, const double* det_tx_dev,
                        const double* det_tz_dev, const double* noise_tx_dev, const double* noise_tz_dev,
                        int n_tags, int32_t* tag_slot_dev, int32_t* n_seen_dev, int32_t* slot_dev, double* rng_dev,
                        double* brg_dev, int m, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (n_det < 0 || n_tags <= 0) return fail(EKF_EINVAL, "n_det < 0 or n_tags <= 0");
  if (m < 1 || m > ekf::ASSOC_MAX_M) return fail(EKF_EINVAL, "m must be in [1, %d]", ekf::ASSOC_MAX_M);
  if ((n_det > 0 && (!det_tag_dev || !det_tx_dev || !det_tz_dev)) || !tag_slot_dev || !n_seen_dev || !slot_dev ||
      !rng_dev || !brg_dev)
    return fail(EKF_EINVAL, "null argument");
  if ((noise_tx_dev == nullptr) != (noise_tz_dev == nullptr)) return fail(EKF_EINVAL, "give both noise arrays or none");
  CU(cudaSetDevice(h->device));
  ekf::AssocArgs a{};
  a.batch = h->batch; a.n_det = n_det; a.n_tags = n_tags; a.m = m;
  a.det_tag = det_tag_dev; a.det_tx = det_tx_dev; a.det_tz = det_tz_dev;
  a.noise_tx = noise_tx_dev; a.noise_tz = noise_tz_dev;
  a.tag_slot = tag_slot_dev; a.n_seen = n_seen_dev;
  a.slot = slot_dev; a.rng = rng_dev; a.brg = brg_dev;
  a.overflow = h->faults + 3;
  CU(ekf::batch_launch_associate(a, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_associate_generated(ekf_batch_t* h, int n_det, const int32_t* det_tag_dev, const double* det_tx_dev,
                                  const double* det_tz_dev, uint64_t seed, uint64_t step, uint64_t first_filter,
                                  double sigma, int n_tags, int32_t* tag_slot_dev, int32_t* n_seen_dev,
                                  int32_t* slot_dev, double* rng_dev, double* brg_dev, int m, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (n_det < 0 || n_tags <= 0) return fail(EKF_EINVAL, "n_det < 0 or n_tags <= 0");
  if (m < 1 || m > ekf::ASSOC_MAX_M) return fail(EKF_EINVAL, "m must be in [1, %d]", ekf::ASSOC_MAX_M);
  if ((n_det > 0 && (!det_tag_dev || !det_tx_dev || !det_tz_dev)) || !tag_slot_dev || !n_seen_dev || !slot_dev ||
      !rng_dev || !brg_dev)
    return fail(EKF_EINVAL, "null argument");
  if (step > 0xffffffffull) return fail(EKF_EINVAL, "step must fit 32 bits");
  CU(cudaSetDevice(h->device));
  ekf::AssocArgs a{};
  a.batch = h->batch; a.n_det = n_det; a.n_tags = n_tags; a.m = m;
  a.det_tag = det_tag_dev; a.det_tx = det_tx_dev; a.det_tz = det_tz_dev;
  a.gen = 1; a.gen_seed = seed; a.gen_step = (uint32_t)step; a.gen_first = first_filter; a.gen_sigma = sigma;
  a.tag_slot = tag_slot_dev; a.n_seen = n_seen_dev;
  a.slot = slot_dev; a.rng = rng_dev; a.brg = brg_dev;
  a.overflow = h->faults + 3;
  CU(ekf::batch_launch_associate(a, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_associate_overflow_count(ekf_batch_t* h, int64_t* count) {
  if (!h || !count) return fail(EKF_EINVAL, "null argument");
  CU(cudaSetDevice(h->device));
  unsigned long long v = 0;
  CU(cudaMemcpy(&v, h->faults + 3, sizeof(v), cudaMemcpyDeviceToHost));
  *count = (int64_t)v;
  return EKF_OK;
}

int ekf_batch_accumulate_ate(ekf_batch_t* h, double gt_x, double gt_y, double* sse_dev, void* stream) {
  if (!h || !sse_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  CU(ekf::batch_launch_ate(h->mu, h->ld, h->batch, gt_x, gt_y, sse_dev, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_batch_moments(ekf_batch_t* h, double* sums_dev, void* stream) {
  if (!h || !sums_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->mu) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  if (!h->mom_partial)
    CU(cudaMalloc(&h->mom_partial, sizeof(double) * ekf::MOM_CTAS * (size_t)(h->n + h->n * h->n)));
  CU(ekf::batch_launch_moments(h->mu, h->ld, h->batch, h->n, h->mom_partial, sums_dev, (cudaStream_t)stream));
  return EKF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// large map
// ---------------------------------------------------------------------------------------------------------
int ekf_lm_layout(int capacity_landmarks, int* n, int64_t* ld) {
  if (capacity_landmarks < 0) return fail(EKF_EINVAL, "capacity < 0");
  const int nn = 3 + 2 * capacity_landmarks;
  if (n) *n = nn;
  if (ld) *ld = ((int64_t)nn + 15) & ~15LL;
  return EKF_OK;
}

int ekf_lm_create(int capacity_landmarks, int device, ekf_lm_t** out) {
  if (!out) return fail(EKF_EINVAL, "out is null");
  if (capacity_landmarks < 0 || capacity_landmarks > 500000000) return fail(EKF_EINVAL, "capacity out of range");
  CU(cudaSetDevice(device));
  ekf_lm* h = new (std::nothrow) ekf_lm();
  if (!h) return fail(EKF_EINVAL, "out of host memory");
  h->cap = capacity_landmarks;
  h->n = 3 + 2 * capacity_landmarks;
  h->device = device;
  h->ldw = (((int64_t)h->n + 15) & ~15LL) + 16;
  const char* v = getenv("EKF_LM_VARIANT");
  h->variant = v ? atoi(v) : 0;
  h->variant_auto = (v == nullptr);
  if (const char* sl = getenv("EKF_LM_SINGLE")) h->single_launch = atoi(sl) != 0;
  double* ws = nullptr;
  const size_t ws_doubles = (size_t)(6 + 6 * ekf::LM_MAX_FUSED) * h->ldw + ekf::kScalCount + 2 * ekf::LM_MAX_FUSED +
                            2 * ekf::LM_MAX_FUSED * ekf::LM_MAX_ROWS;
  cudaError_t e = cudaMalloc(&ws, ws_doubles * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(ws, 0, ws_doubles * sizeof(double));
  if (e != cudaSuccess) {
    cudaFree(ws);
    delete h;
    return fail(EKF_ECUDA, "ekf_lm_create: %s", cudaGetErrorString(e));
  }
  h->K0 = ws;
  h->K1 = ws + h->ldw;
  h->A0 = ws + 2 * h->ldw;
  h->A1 = ws + 3 * h->ldw;
  h->W0 = ws + 4 * h->ldw;
  h->W1 = ws + 5 * h->ldw;
  h->scal = ws + 6 * h->ldw;
  h->my = h->scal + ekf::kScalCount;
  h->mK = h->my + 2 * ekf::LM_MAX_FUSED;
  h->mW = h->mK + 2 * ekf::LM_MAX_FUSED * h->ldw;
  h->mA = h->mW + 2 * ekf::LM_MAX_FUSED * h->ldw;
  h->mKR = h->mA + 2 * ekf::LM_MAX_FUSED * h->ldw;
  *out = h;
  return EKF_OK;
}

int ekf_lm_destroy(ekf_lm_t* h) {
  if (!h) return EKF_OK;
  cudaSetDevice(h->device);
  for (int r = 0; r < h->p2p_world; ++r)
    if (h->p2p_opened[r]) cudaIpcCloseMemHandle(h->p2p_peer_host[r]);
  cudaFree(h->p2p_area);
  cudaFree(h->p2p_peers_dev);
  cudaFree(h->p2p_err);
  cudaFree(h->K0);
  delete h;
  return EKF_OK;
}

int ekf_lm_bind_state(ekf_lm_t* h, double* mu_dev, double* p_dev, int64_t ld, int row_begin, int row_end) {
  if (!h || !mu_dev || !p_dev) return fail(EKF_EINVAL, "null argument");
  if (ld < h->n || (ld & 3)) return fail(EKF_EINVAL, "ld must be >= n and a multiple of 4");
  if (((uintptr_t)p_dev & 31)) return fail(EKF_EINVAL, "P must be 32-byte aligned");
  if (row_begin < 0 || row_end > h->n || row_begin > row_end) return fail(EKF_EINVAL, "bad row range");
  if (row_begin != 0 && (row_begin < 3 || ((row_begin - 3) & 1)))
    return fail(EKF_EINVAL, "row blocks must start on a landmark pair (row 3 + 2k)");
  h->mu = mu_dev;
  h->P = p_dev;
  h->ld = ld;
  h->row_begin = row_begin;
  h->row_end = row_end;
  // tiling of the streaming kernel: a row block that fits L2 (a 1/8 or 1/4 shard of the 4096-landmark map: 67 / 134 MB)
  // runs 6-12 % faster with 16-row tiles (twice the CTAs for the launch ramp and the tail; tools/lm_shard_probe.py:
  // 20.7 -> 18.3 us at 1027 rows, 49.9 -> 47.1 us at 2051), blocks that stream from HBM keep the 32-row tiles
  if (h->variant_auto) h->variant = ((int64_t)(row_end - row_begin) * ld * 8 <= (160LL << 20)) ? 1 : 0;
  return EKF_OK;
}

// Reference-layout state <-> device layout in ONE call each (the scalar drop-in does this around every
// EKF_pose_estimation; through torch slicing it was ~60 us of host overhead per direction at n <= 29).
int ekf_lm_upload(ekf_lm_t* h, int m, const double* mean_host, const double* cov_host, void* stream) {
  if (!h || !mean_host || !cov_host) return fail(EKF_EINVAL, "null argument");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (m < 0 || m > h->n) return fail(EKF_EINVAL, "state size %d outside [0, %d]", m, h->n);
  CU(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int rows = h->row_end - h->row_begin;
  CU(cudaMemsetAsync(h->mu, 0, sizeof(double) * h->ld, st));
  CU(cudaMemsetAsync(h->P, 0, sizeof(double) * (size_t)rows * h->ld, st));
  if (m > 0) CU(cudaMemcpyAsync(h->mu, mean_host, sizeof(double) * m, cudaMemcpyHostToDevice, st));
  const int lo = h->row_begin, hi = h->row_end < m ? h->row_end : m;
  if (hi > lo)
    CU(cudaMemcpy2DAsync(h->P, sizeof(double) * h->ld, cov_host + (size_t)lo * m, sizeof(double) * m,
                         sizeof(double) * m, hi - lo, cudaMemcpyHostToDevice, st));
  h->p2p_robot_seq = 0;
  return EKF_OK;
}

// mean_host [m]; cov_host [(min(row_end, m) - row_begin) x m], C order: this block's live rows. Synchronous.
int ekf_lm_download(ekf_lm_t* h, int m, double* mean_host, double* cov_host, void* stream) {
  if (!h || !mean_host || !cov_host) return fail(EKF_EINVAL, "null argument");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (m < 0 || m > h->n) return fail(EKF_EINVAL, "state size %d outside [0, %d]", m, h->n);
  CU(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  if (m > 0) CU(cudaMemcpyAsync(mean_host, h->mu, sizeof(double) * m, cudaMemcpyDeviceToHost, st));
  const int lo = h->row_begin, hi = h->row_end < m ? h->row_end : m;
  if (hi > lo)
    CU(cudaMemcpy2DAsync(cov_host, sizeof(double) * m, h->P, sizeof(double) * h->ld, sizeof(double) * m, hi - lo,
                         cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return EKF_OK;
}

int ekf_lm_set_noise(ekf_lm_t* h, double motion_var, double meas_var) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  h->mv = motion_var;
  h->mm = meas_var;
  return EKF_OK;
}

int ekf_lm_set_flags(ekf_lm_t* h, int circular, int measurement, int disable_motion, int cov_form) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (cov_form != EKF_COV_PLAIN && cov_form != EKF_COV_JOSEPH) return fail(EKF_EINVAL, "unknown cov_form");
  h->flags = make_flags(circular, measurement, disable_motion, cov_form);
  return EKF_OK;
}

int ekf_lm_augment(ekf_lm_t* h, int slot, double gx, double gy, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (slot < 0 || slot >= h->cap) return fail(EKF_EINVAL, "slot %d outside capacity %d", slot, h->cap);
  CU(cudaSetDevice(h->device));
  CU(ekf::lm_launch_augment(h->mu, h->P, h->ld, h->row_begin, h->row_end, 3 + 2 * slot, gx, gy,
                            (cudaStream_t)stream));
  return EKF_OK;
}

// The robot rows pre-pushed for update p2p_seq + 1 have gone stale (predict, external write to P). Their cells already
// carry the tag p2p_seq + 1, so that number is BURNT: the next update takes p2p_seq + 2 and rank 0's gather pushes
// fresh rows under it — a consumer can never mistake the stale push for the fresh one (tags are matched exactly).
// Every rank makes the same calls in the same order, so the numbering stays in step; acks are compared with >=, so a
// skipped number is fine.
static void p2p_drop_prepush(ekf_lm_t* h) {
  if (h->p2p_robot_seq != 0 && h->p2p_robot_seq == h->p2p_seq + 1) ++h->p2p_seq;
  h->p2p_robot_seq = 0;
}

int ekf_lm_predict(ekf_lm_t* h, double ang, double lin, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  CU(cudaSetDevice(h->device));
  CU(ekf::lm_launch_predict(h->mu, h->P, h->ld, h->n, h->row_begin, h->row_end, h->scal, ang, lin, h->mv,
                            h->flags, (cudaStream_t)stream));
  p2p_drop_prepush(h);  // rows 0-1 and the pose changed: the robot rows pre-pushed by the last update are stale
  return EKF_OK;
}

// exchange descriptor of update / step number `seq` of this handle's peer-memory exchange
// area = [2][P2P_ROWS][ld + 4] 16-byte cells (two doubles each), then the ack words (ekf_largemap.h)
static int64_t p2p_flag_off(const ekf_lm* h) { return 2 * ekf::P2P_ROWS * (h->ld + 4) * 2; }
static int64_t p2p_area_doubles(const ekf_lm* h) { return p2p_flag_off(h) + ekf::P2P_MAX_WORLD; }

static ekf::LmExchange p2p_exchange(ekf_lm_t* h, unsigned long long seq) {
  ekf::LmExchange x{};
  x.peers = h->p2p_peers_dev;
  x.ldx = h->ld + 4;
  x.flag_off = p2p_flag_off(h);
  x.acks = reinterpret_cast<unsigned long long*>(h->p2p_area + x.flag_off);
  x.err = h->p2p_err;
  x.seq = seq;
  x.world = h->p2p_world;
  x.rank = h->p2p_rank;
  return x;
}

// One range-bearing update. rows5: replicated 5-row buffer of a row-block shard whose rows were exchanged by the host
// (NCCL path / emulated shards); owner >= 0: peer-memory exchange inside the kernels (ekf_lm_p2p_update).
static int lm_update_impl_split(ekf_lm_t* h, int slot, double rng, double brg, const double* rows5, void* stream,
                                int owner) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (slot < 0 || slot >= h->cap) return fail(EKF_EINVAL, "slot %d outside capacity %d", slot, h->cap);
  if (!(h->flags & ekf::kFlagMeasurement)) return EKF_OK;  // ENABLE_MEASUREMENT_MODEL = False (:435)
  const bool whole = (h->row_begin == 0 && h->row_end == h->n);
  const bool p2p = owner >= 0;
  if (!rows5 && !whole && !p2p) return fail(EKF_ESTATE, "row-block shard: use ekf_lm_update_from_rows");
  const bool joseph = (h->flags & ekf::kFlagJoseph) != 0;
  if (joseph && !whole && !p2p)
    return fail(EKF_EINVAL, "Joseph form on a row-block shard needs the peer-memory exchange (ekf_lm_p2p_update)");
  CU(cudaSetDevice(h->device));
  const int j = 3 + 2 * slot;
  const int idx5[5] = {0, 1, 2, j, j + 1};
  ekf::LmGatherArgs g{};
  ekf::LmUpdateArgs u{};
  g.P = h->P;
  int64_t ldx = h->ld + 4, mu_off = h->ld;  // replicated rows buffer: plain doubles, mu entry at column ld
  if (p2p) {
    const unsigned long long seq = ++h->p2p_seq;
    g.x = u.x = p2p_exchange(h, seq);
    ldx = 2 * (h->ld + 4);  // exchange buffer: 16-byte cells
    mu_off = 2 * h->ld;
    rows5 = h->p2p_area + (int64_t)(seq & 1) * ekf::P2P_ROWS * ldx;
    const bool robot_there = (h->p2p_robot_seq == seq);  // pre-pushed by rank 0's previous streaming kernel
    if (h->p2p_rank == 0) {
      if (!robot_there) g.push_mask |= 1;
    } else {
      g.wait_mask |= 1;
    }
    if (h->p2p_rank == owner) g.push_mask |= 2; else g.wait_mask |= 2;
    // every rank keeps the same book: rank 0's streaming kernel of this update pre-pushes the robot rows of the
    // next one (plain form; the Joseph streaming kernel does not, so rank 0's next gather pushes them)
    u.prepush = (h->p2p_rank == 0) && !joseph;
    h->p2p_robot_seq = joseph ? 0 : seq + 1;
    g.share_k = joseph ? 1 : 0;
  }
  for (int k = 0; k < 5; ++k) {
    const bool own = idx5[k] >= h->row_begin && idx5[k] < h->row_end;
    if (rows5 && !(p2p && own)) {
      g.row[k] = rows5 + k * ldx;
      g.mu5[k] = rows5 + k * ldx + mu_off;
    } else {
      g.row[k] = h->P + (int64_t)(idx5[k] - h->row_begin) * h->ld;
      g.mu5[k] = h->mu + idx5[k];
    }
  }
  g.K0 = h->K0; g.K1 = h->K1; g.A0 = h->A0; g.A1 = h->A1; g.W0 = h->W0; g.W1 = h->W1; g.scal = h->scal;
  g.ld = h->ld;
  g.n = h->n;
  g.row_begin = h->row_begin;
  g.row_end = h->row_end;
  g.j = j;
  g.flags = h->flags;
  g.rng = rng;
  g.brg = brg;
  g.mm = h->mm;
  g.shard = rows5 ? 1 : 0;
  g.mu_out = h->mu;
  u.P = h->P;
  u.mu = h->mu;
  u.K0 = h->K0; u.K1 = h->K1; u.A0 = h->A0; u.A1 = h->A1; u.W0 = h->W0; u.W1 = h->W1; u.scal = h->scal;
  u.ld = h->ld;
  u.n = h->n;
  u.row_begin = h->row_begin;
  u.row_end = h->row_end;
  if (!joseph && h->variant <= 1 && h->single_launch && !p2p) {
    // ONE launch: the leading CTAs run the gather, the streaming tiles pre-load their first rows and wait for it.
    // Not with the peer-memory exchange: measured no gain there with the flag protocol of the time (4096 landmarks on
    // 8 GPUs: 0.0419 vs 0.0403 ms) and, at 32 768 landmarks, one rank's wait of the very first update ran into its
    // 2 s budget (results intact; cause not found) — the sharded update keeps its two PDL-chained launches.
    CU(ekf::lm_launch_update_single(g, u, h->variant, reinterpret_cast<unsigned long long*>(h->scal + ekf::kScalReady),
                                    &h->single_calls, (cudaStream_t)stream));
    return EKF_OK;
  }
  CU(ekf::lm_launch_gather(g, (cudaStream_t)stream));
  CU(ekf::lm_launch_update(u, joseph, h->variant, (cudaStream_t)stream));
  return EKF_OK;
}

static int lm_update_impl(ekf_lm_t* h, int slot, double rng, double brg, const double* rows5, void* stream) {
  return lm_update_impl_split(h, slot, rng, brg, rows5, stream, -1);
}

int ekf_lm_update(ekf_lm_t* h, int slot, double rng, double brg, void* stream) {
  return lm_update_impl(h, slot, rng, brg, nullptr, stream);
}

static int lm_multi_check(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                          const double* brg_host, bool need_obs = true) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (m < 0) return fail(EKF_EINVAL, "m < 0");
  if (m > 0 && (!slots_host || (need_obs && (!rng_host || !brg_host))))
    return fail(EKF_EINVAL, "observation arrays are null");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  for (int u = 0; u < m; ++u)
    if (slots_host[u] < 0 || slots_host[u] >= h->cap)
      return fail(EKF_EINVAL, "slot %d outside capacity %d", slots_host[u], h->cap);
  return EKF_OK;
}

// xrows: exchanged rows of a row-block shard ([3 + 2g][ldx]); own_local: rows this rank owns are read from its own P
// (peer-memory exchange: owners do not push to themselves), otherwise every row comes from xrows (replicated buffer)
// (those xrows are then the 16-byte cells of the exchange buffer, not plain doubles)
static void lm_multi_fill(ekf_lm_t* h, ekf::LmMultiArgs& a, int g, const int32_t* slots, const double* rng,
                          const double* brg, const double* xrows, bool own_local = false) {
  const int64_t xscale = own_local ? 2 : 1;
  a.P = h->P; a.mu = h->mu; a.K = h->mK; a.W = h->mW; a.A = h->mA; a.y = h->my; a.KR = h->mKR;
  a.xrows = xrows;
  a.ld = h->ld; a.ldw = h->ldw; a.ldx = h->ld + 4;
  a.n = h->n; a.m = g; a.row_begin = h->row_begin; a.row_end = h->row_end; a.mm = h->mm;
  a.variant = h->variant == 1 ? 1 : 0;
  for (int u = 0; u < g; ++u) {
    a.j[u] = 3 + 2 * slots[u];
    if (rng) a.rng[u] = rng[u];
    if (brg) a.brg[u] = brg[u];
  }
  for (int rho = 0; rho < 3 + 2 * g; ++rho) {
    const int row = rho < 3 ? rho : a.j[(rho - 3) >> 1] + ((rho - 3) & 1);
    const bool own = row >= h->row_begin && row < h->row_end;
    if (xrows && !(own_local && own)) {
      a.rowp[rho] = xrows + (int64_t)rho * a.ldx * xscale;
      a.mup[rho] = a.rowp[rho] + h->ld * xscale;
    } else {
      a.rowp[rho] = h->P + (int64_t)(row - h->row_begin) * h->ld;
      a.mup[rho] = h->mu + row;
    }
  }
}

int ekf_lm_update_multi(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                        const double* brg_host, void* stream) {
  const int rc0 = lm_multi_check(h, m, slots_host, rng_host, brg_host);
  if (rc0 != EKF_OK) return rc0;
  const bool whole = (h->row_begin == 0 && h->row_end == h->n);
  const bool joseph = (h->flags & ekf::kFlagJoseph) != 0;
  if (!(h->flags & ekf::kFlagMeasurement)) return EKF_OK;
  if (!whole) {  // a shard goes through ekf_lm_update_multi_from_rows / ekf_lm_p2p_update_multi (fails like ekf_lm_update)
    for (int u = 0; u < m; ++u) {
      const int rc = lm_update_impl(h, slots_host[u], rng_host[u], brg_host[u], nullptr, stream);
      if (rc != EKF_OK) return rc;
    }
    return EKF_OK;
  }
  CU(cudaSetDevice(h->device));
  for (int first = 0; first < m; first += ekf::LM_MAX_FUSED) {  // groups of <= LM_MAX_FUSED, each ONE pass over P
    const int g = (m - first < ekf::LM_MAX_FUSED) ? m - first : ekf::LM_MAX_FUSED;
    if (g == 1) {
      const int rc = lm_update_impl(h, slots_host[first], rng_host[first], brg_host[first], nullptr, stream);
      if (rc != EKF_OK) return rc;
      continue;
    }
    ekf::LmMultiArgs a{};
    lm_multi_fill(h, a, g, slots_host + first, rng_host + first, brg_host + first, nullptr);
    a.joseph = joseph ? 1 : 0;
    CU(ekf::lm_launch_update_multi(a, (cudaStream_t)stream));
  }
  return EKF_OK;
}

int ekf_lm_export_rows_multi(ekf_lm_t* h, int m, const int32_t* slots_host, double* rows_dev, void* stream) {
  const int rc0 = lm_multi_check(h, m, slots_host, nullptr, nullptr, false);
  if (rc0 != EKF_OK) return rc0;
  if (!rows_dev) return fail(EKF_EINVAL, "rows_dev is null");
  if (m < 2 || m > ekf::LM_MAX_FUSED) return fail(EKF_EINVAL, "m must be in [2, %d]", ekf::LM_MAX_FUSED);
  CU(cudaSetDevice(h->device));
  ekf::LmMultiArgs a{};
  lm_multi_fill(h, a, m, slots_host, nullptr, nullptr, nullptr);
  CU(ekf::lm_launch_export_rows_multi(a, rows_dev, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_lm_update_multi_from_rows(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                                  const double* brg_host, const double* rows_dev, void* stream) {
  const int rc0 = lm_multi_check(h, m, slots_host, rng_host, brg_host);
  if (rc0 != EKF_OK) return rc0;
  if (!rows_dev) return fail(EKF_EINVAL, "rows_dev is null");
  if (m < 2 || m > ekf::LM_MAX_FUSED) return fail(EKF_EINVAL, "m must be in [2, %d]", ekf::LM_MAX_FUSED);
  if (h->flags & ekf::kFlagJoseph) return fail(EKF_EINVAL, "Joseph form is not available on a row-block shard");
  if (!(h->flags & ekf::kFlagMeasurement)) return EKF_OK;
  CU(cudaSetDevice(h->device));
  ekf::LmMultiArgs a{};
  lm_multi_fill(h, a, m, slots_host, rng_host, brg_host, rows_dev);
  CU(ekf::lm_launch_update_multi(a, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_lm_export_rows(ekf_lm_t* h, int slot, double* rows5_dev, void* stream) {
  if (!h || !rows5_dev) return fail(EKF_EINVAL, "null argument");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (slot < 0 || slot >= h->cap) return fail(EKF_EINVAL, "slot %d outside capacity %d", slot, h->cap);
  CU(cudaSetDevice(h->device));
  CU(ekf::lm_launch_export_rows(h->P, h->mu, h->ld, h->row_begin, h->row_end, 3 + 2 * slot, rows5_dev,
                                h->ld + 4, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_lm_update_from_rows(ekf_lm_t* h, int slot, double rng, double brg, const double* rows5_dev,
                            void* stream) {
  if (!rows5_dev) return fail(EKF_EINVAL, "rows5_dev is null");
  return lm_update_impl(h, slot, rng, brg, rows5_dev, stream);
}

// ---- peer-memory row exchange of the row-block shards ----------------------------------------------------
int ekf_lm_p2p_init(ekf_lm_t* h, int world, int rank, unsigned char* handle_out, int64_t handle_bytes) {
  if (!h || !handle_out) return fail(EKF_EINVAL, "null argument");
  if (!h->P) return fail(EKF_ESTATE, "state not bound");
  if (world < 1 || world > ekf::P2P_MAX_WORLD || rank < 0 || rank >= world)
    return fail(EKF_EINVAL, "world must be in [1, %d] and rank in [0, world)", ekf::P2P_MAX_WORLD);
  if (handle_bytes < (int64_t)sizeof(cudaIpcMemHandle_t)) return fail(EKF_EINVAL, "handle buffer too small");
  if (h->p2p_area) return fail(EKF_ESTATE, "p2p already initialised");
  CU(cudaSetDevice(h->device));
  const size_t bytes = (size_t)p2p_area_doubles(h) * sizeof(double);
  CU(cudaMalloc(&h->p2p_area, bytes));
  CU(cudaMemset(h->p2p_area, 0, bytes));
  CU(cudaMalloc(&h->p2p_peers_dev, ekf::P2P_MAX_WORLD * sizeof(double*)));
  CU(cudaMalloc(&h->p2p_err, sizeof(unsigned long long)));
  CU(cudaMemset(h->p2p_err, 0, sizeof(unsigned long long)));
  cudaIpcMemHandle_t hd;
  CU(cudaIpcGetMemHandle(&hd, h->p2p_area));
  memcpy(handle_out, &hd, sizeof(hd));
  h->p2p_world = world;
  h->p2p_rank = rank;
  h->p2p_seq = 0;
  CU(cudaDeviceSynchronize());
  return EKF_OK;
}

// handles: [world][handle_bytes] as gathered from every rank's ekf_lm_p2p_init (own entry ignored).
// local_areas: optional [world] device pointers for shards living in THIS process (NULL entry = open the handle).
int ekf_lm_p2p_connect(ekf_lm_t* h, const unsigned char* handles, int64_t handle_bytes, void* const* local_areas) {
  if (!h || (!handles && !local_areas)) return fail(EKF_EINVAL, "null argument");
  if (!h->p2p_area) return fail(EKF_ESTATE, "call ekf_lm_p2p_init first");
  CU(cudaSetDevice(h->device));
  for (int r = 0; r < h->p2p_world; ++r) {
    if (r == h->p2p_rank) {
      h->p2p_peer_host[r] = h->p2p_area;
    } else if (local_areas && local_areas[r]) {
      h->p2p_peer_host[r] = (double*)local_areas[r];
    } else {
      if (!handles) return fail(EKF_EINVAL, "no handle for rank %d", r);
      cudaIpcMemHandle_t hd;
      memcpy(&hd, handles + (size_t)r * handle_bytes, sizeof(hd));
      void* p = nullptr;
      CU(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
      h->p2p_peer_host[r] = (double*)p;
      h->p2p_opened[r] = true;
    }
  }
  CU(cudaMemcpy(h->p2p_peers_dev, h->p2p_peer_host, h->p2p_world * sizeof(double*), cudaMemcpyHostToDevice));
  return EKF_OK;
}

int ekf_lm_p2p_area(ekf_lm_t* h, void** area_dev) {
  if (!h || !area_dev) return fail(EKF_EINVAL, "null argument");
  *area_dev = h->p2p_area;
  return EKF_OK;
}

int ekf_lm_p2p_update(ekf_lm_t* h, int slot, double rng, double brg, int owner, void* stream) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  if (!h->p2p_peers_dev) return fail(EKF_ESTATE, "p2p exchange not connected");
  if (slot < 0 || slot >= h->cap) return fail(EKF_EINVAL, "slot %d outside capacity %d", slot, h->cap);
  if (owner < 0 || owner >= h->p2p_world) return fail(EKF_EINVAL, "owner out of range");
  if (!(h->flags & ekf::kFlagMeasurement)) return EKF_OK;
  if (h->p2p_world == 1)  // nobody to exchange with: the plain single-GPU update, no push / flag / ack protocol
    return lm_update_impl(h, slot, rng, brg, nullptr, stream);
  const int j = 3 + 2 * slot;
  if (h->p2p_rank == owner && (j < h->row_begin || j + 1 >= h->row_end))
    return fail(EKF_EINVAL, "rank %d does not own slot %d", owner, slot);
  // two launches, as on one GPU: the gather (owners push their rows at its head, everyone else waits on the flags in
  // its own memory) and the streaming update (acknowledges the buffer; rank 0 pre-pushes rows 0-2 for the next update)
  return lm_update_impl_split(h, slot, rng, brg, nullptr, stream, owner);
}

// The fused step over peer memory: the fused gather pushes / waits for the step's 1 + m row parts, then ONE streaming
// pass over the local row block.
int ekf_lm_p2p_update_multi(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                            const double* brg_host, const int32_t* owners_host, void* stream) {
  const int rc0 = lm_multi_check(h, m, slots_host, rng_host, brg_host);
  if (rc0 != EKF_OK) return rc0;
  if (!h->p2p_peers_dev) return fail(EKF_ESTATE, "p2p exchange not connected");
  if (!owners_host) return fail(EKF_EINVAL, "owners_host is null");
  if (m < 2 || m > ekf::LM_MAX_FUSED) return fail(EKF_EINVAL, "m must be in [2, %d]", ekf::LM_MAX_FUSED);
  if (h->p2p_world == 1) return ekf_lm_update_multi(h, m, slots_host, rng_host, brg_host, stream);
  if (!(h->flags & ekf::kFlagMeasurement)) return EKF_OK;
  if (h->flags & ekf::kFlagJoseph) {  // no fused Joseph step on a shard: the step's updates one after the other
    for (int u = 0; u < m; ++u) {
      const int rc = ekf_lm_p2p_update(h, slots_host[u], rng_host[u], brg_host[u], owners_host[u], stream);
      if (rc != EKF_OK) return rc;
    }
    return EKF_OK;
  }
  for (int u = 0; u < m; ++u) {
    const int j = 3 + 2 * slots_host[u];
    if (owners_host[u] < 0 || owners_host[u] >= h->p2p_world) return fail(EKF_EINVAL, "owner out of range");
    if (h->p2p_rank == owners_host[u] && (j < h->row_begin || j + 1 >= h->row_end))
      return fail(EKF_EINVAL, "rank %d does not own slot %d", owners_host[u], slots_host[u]);
  }
  CU(cudaSetDevice(h->device));
  const unsigned long long seq = ++h->p2p_seq;
  ekf::LmMultiArgs g{};
  lm_multi_fill(h, g, m, slots_host, rng_host, brg_host,
                h->p2p_area + (int64_t)(seq & 1) * ekf::P2P_ROWS * (h->ld + 4) * 2, /*own_local=*/true);
  g.x = p2p_exchange(h, seq);
  const bool robot_there = (h->p2p_robot_seq == seq);
  if (h->p2p_rank == 0) {
    if (!robot_there) g.push_mask |= 1;
  } else {
    g.wait_mask |= 1;
  }
  for (int u = 0; u < m; ++u) {
    if (h->p2p_rank == owners_host[u]) g.push_mask |= 2 << u; else g.wait_mask |= 2 << u;
  }
  g.prepush = (h->p2p_rank == 0);
  h->p2p_robot_seq = seq + 1;
  CU(ekf::lm_launch_update_multi(g, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf_lm_p2p_invalidate(ekf_lm_t* h) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  p2p_drop_prepush(h);
  return EKF_OK;
}

int ekf_lm_p2p_error(ekf_lm_t* h, int64_t* flag) {
  if (!h || !flag) return fail(EKF_EINVAL, "null argument");
  *flag = 0;
  if (!h->p2p_err) return EKF_OK;
  CU(cudaSetDevice(h->device));
  unsigned long long v = 0;
  CU(cudaMemcpy(&v, h->p2p_err, sizeof(v), cudaMemcpyDeviceToHost));
  *flag = (int64_t)v;
  return EKF_OK;
}

int ekf_lm_sync(ekf_lm_t* h) {
  if (!h) return fail(EKF_EINVAL, "null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  return EKF_OK;
}

// ---------------------------------------------------------------------------------------------------------
// 3-state EKF
// ---------------------------------------------------------------------------------------------------------
int ekf3_predict(int64_t batch, double* state_dev, double* cov_dev, const double* v_dev, const double* omega_dev,
                 const double* dt_dev, void* stream) {
  if (batch < 0) return fail(EKF_EINVAL, "batch < 0");
  if (!state_dev || !cov_dev || !v_dev || !omega_dev || !dt_dev) return fail(EKF_EINVAL, "null argument");
  CU(ekf::ekf3_launch_predict(batch, state_dev, cov_dev, v_dev, omega_dev, dt_dev, (cudaStream_t)stream));
  return EKF_OK;
}

int ekf3_update(int64_t batch, double* state_dev, double* cov_dev, const double* rng_dev, const double* brg_dev,
                const double* lx_dev, const double* ly_dev, void* stream) {
  if (batch < 0) return fail(EKF_EINVAL, "batch < 0");
  if (!state_dev || !cov_dev || !rng_dev || !brg_dev || !lx_dev || !ly_dev)
    return fail(EKF_EINVAL, "null argument");
  CU(ekf::ekf3_launch_update(batch, state_dev, cov_dev, rng_dev, brg_dev, lx_dev, ly_dev, (cudaStream_t)stream));
  return EKF_OK;
}

}  // extern "C"
