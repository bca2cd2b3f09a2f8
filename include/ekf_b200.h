/* ekf_b200.h — C-ABI of the B200-native EKF-SLAM filter core (libekf_b200.so, sm_100a).
 *
 * Drop-in boundary for the predict/update hot path of AHHHZ975/SLAM-Duckietown. The reference has no FFI of
 * its own (src/ekf_bindings.py is an empty file); the entry points below are what a ctypes binding in that
 * file binds so that the reference's module-level functions keep their signatures:
 *
 *   delta_phi            src/replay_no_ros.py:250-266   -> ekf_delta_phi
 *   displacement         src/replay_no_ros.py:484-497   -> ekf_displacement
 *   EKF_pose_estimation  src/replay_no_ros.py:269-482   -> ekf_lm_* (single filter, any n) /
 *                                                          ekf_batch_* (many filters, fused step)
 *   predict              src/EKF-SLAM.py:29-56          -> ekf3_predict
 *   update               src/EKF-SLAM.py:59-84          -> ekf3_update
 *
 * Conventions
 *   - every function returns 0 on success, a negative EKF_E* code on failure; ekf_last_error() gives the
 *     thread-local message. No C++ exception and no abort() crosses this boundary.
 *   - pointers named *_dev are DEVICE pointers owned by the caller (e.g. torch tensor storage); *_host are
 *     host pointers. `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - all arithmetic is IEEE float64; slots / ids / ticks are integers.
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with EKF_ECUDA.
 */
#ifndef EKF_B200_H
#define EKF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EKF_OK 0
#define EKF_EINVAL (-1)  /* bad argument */
#define EKF_ECUDA (-2)   /* CUDA runtime error (message has cudaGetErrorString) */
#define EKF_ESTATE (-3)  /* handle not bound / wrong mode */

/* covariance update forms */
#define EKF_COV_PLAIN 0  /* P <- P - K (H P): the reference, replay_no_ros.py:480 */
#define EKF_COV_JOSEPH 1 /* P <- (I-KH) P (I-KH)^T + K Q K^T as a one-pass rank-4 update */

int ekf_version(void);
const char* ekf_last_error(void);
int ekf_device_count(int* count);

/* ---- a1/a2: odometry from wheel-encoder ticks (host scalars) ------------------------------------------ */
double ekf_delta_phi(int64_t ticks, int64_t prev_ticks, int64_t resolution); /* replay_no_ros.py:262-266 */
int ekf_displacement(double wheel_radius, double baseline, double dphi_left, double dphi_right,
                     double* angular, double* linear);                        /* replay_no_ros.py:491-497 */

/* ---- batched mode: B independent filters of fixed capacity, one CTA per filter ------------------------- */
typedef struct ekf_batch ekf_batch_t;

/* Device layout of one filter (n = 3 + 2*n_landmarks, ld = n rounded up to 4):
 *   mu  : mu_stride doubles (= ld), entries [n, ld) zero
 *   P   : p_stride doubles (= n*ld), row-major with leading dimension ld, pad columns zero            */
int ekf_batch_layout(int n_landmarks, int* n, int* ld, int64_t* p_stride, int64_t* mu_stride);
int ekf_batch_create(int n_landmarks, int64_t batch, int device, ekf_batch_t** out);
int ekf_batch_destroy(ekf_batch_t* h);
/* mu_dev [batch][mu_stride], p_dev [batch][p_stride], n_known_dev [batch] (landmarks initialised so far) */
int ekf_batch_bind_state(ekf_batch_t* h, double* mu_dev, double* p_dev, int32_t* n_known_dev);
/* per-filter MOTION_MODEL_VARIANCE / MEASUREMENT_MODEL_VARIANCE (replay_no_ros.py:15-16), [batch] each */
int ekf_batch_set_noise(ekf_batch_t* h, const double* motion_var_dev, const double* meas_var_dev);
/* ENABLE_CIRCULAR_INTERPOLATION, ENABLE_MEASUREMENT_MODEL, DISABLE_MOTION_MODEL (:18-19,:28), cov form */
int ekf_batch_set_flags(ekf_batch_t* h, int circular, int measurement, int disable_motion, int cov_form);
/* mu = 0, P = p0_diag * I3 (replay_no_ros.py:69-70), n_known = 0 */
int ekf_batch_init(ekf_batch_t* h, double p0_diag, void* stream);
/* One fused filter-step for every filter (replay_no_ros.py:341-480): augment slots >= n_known from the
 * pre-predict pose, odometry predict, then up to m sequential range-bearing updates in the given order.
 * ang/lin [batch]; slot/rng/brg [batch][m], slot < 0 = no observation.                                   */
int ekf_batch_step(ekf_batch_t* h, const double* ang_dev, const double* lin_dev, const int32_t* slot_dev,
                   const double* rng_dev, const double* brg_dev, int m, void* stream);
/* Same, odometry given as encoder ticks: delta_phi + displacement (a1/a2) are evaluated on the device.  */
int ekf_batch_step_ticks(ekf_batch_t* h, const int32_t* ticks_left_dev, const int32_t* ticks_right_dev,
                         const int32_t* prev_left_dev, const int32_t* prev_right_dev, double wheel_radius,
                         double baseline, int resolution, const int32_t* slot_dev, const double* rng_dev,
                         const double* brg_dev, int m, void* stream);
/* Same, every filter sees the SAME encoder ticks (Monte Carlo replays of one bag): four scalars, no per-filter arrays */
int ekf_batch_step_ticks_common(ekf_batch_t* h, int32_t ticks_left, int32_t ticks_right, int32_t prev_left,
                                int32_t prev_right, double wheel_radius, double baseline, int resolution,
                                const int32_t* slot_dev, const double* rng_dev, const double* brg_dev, int m,
                                void* stream);
/* Reference layout (mean (n,), covariance (n,n) C-order per filter, device memory) <-> device layout.   */
int ekf_batch_pack(ekf_batch_t* h, const double* mean_ref_dev, const double* cov_ref_dev, int64_t first,
                   int64_t count, void* stream);
int ekf_batch_unpack(ekf_batch_t* h, double* mean_ref_dev, double* cov_ref_dev, int64_t first, int64_t count,
                     void* stream);
/* End-to-end call with HOST buffers: pinned staging + H2D of this step's inputs, the fused step, D2H of the
 * robot pose (x, y, theta) of every filter into pose_host [batch][3]; synchronous.
 * Stream contract (both host entry points): the work runs on the handle's own streams; `stream` is the caller's
 * stream that last touched the bound state (ekf_batch_init / step / pack ...). On entry the handle's streams wait
 * for everything enqueued on it so far; on exit `stream` is made to wait for this step's kernel, so later calls
 * on it (ekf_batch_step, unpack, moments) see the state this call produced.                                 */
int ekf_batch_step_host(ekf_batch_t* h, const double* ang_host, const double* lin_host,
                        const int32_t* slot_host, const double* rng_host, const double* brg_host, int m,
                        double* pose_host, void* stream);
/* Pipelined form of the same call: returns once the work is enqueued; step i's H2D copies overlap step i-1's
 * kernel and step i-2's D2H (three streams, two staging sets). Host buffers must stay valid and pose_host
 * unread until ekf_batch_host_sync returns.                                                             */
int ekf_batch_step_host_async(ekf_batch_t* h, const double* ang_host, const double* lin_host,
                              const int32_t* slot_host, const double* rng_host, const double* brg_host, int m,
                              double* pose_host, void* stream);
int ekf_batch_host_sync(ekf_batch_t* h);
/* number of filter-steps since init that were handed a slot beyond the capacity or one that skips first-seen
 * order (the reference raises KeyError there, replay_no_ros.py:359); those observations are not applied       */
int ekf_batch_fault_count(ekf_batch_t* h, int64_t* count);
/* health check on demand: number of filters whose mu or P currently holds a non-finite value (the reference
 * carries inf / nan silently, e.g. q = 0 at :446-469); scans the whole state once, synchronous on `stream`   */
int ekf_batch_nonfinite_count(ekf_batch_t* h, int64_t* count, void* stream);

/* a3/a4 on the device, per filter (replay_no_ros.py:280-337): gate the window's detections on their per-filter
 * perturbed camera-frame position (pose_t[0] + noise_tx, pose_t[2] + noise_tz; strict z^2 + x^2 > 2.25 drops),
 * assign landmark slots in first-seen order in the filter's own slot table, average the window per slot and convert
 * to range / bearing. Writes the [batch][m] (slot, rng, brg) block ekf_batch_step consumes, in first-seen-in-window
 * order. det_tag holds compact tag indices in [0, n_tags) (the host maps tag ids and marks IGNORE_TAGS < 0);
 * noise arrays are [n_det][batch] (or both NULL); tag_slot is [n_tags][batch] (-1 = never seen), n_seen [batch].
 * Gate decisions, slot tables and update order are bit-exact with the host association.                       */
int ekf_batch_associate(ekf_batch_t* h, int n_det, const int32_t* det_tag_dev, const double* det_tx_dev,
                        const double* det_tz_dev, const double* noise_tx_dev, const double* noise_tz_dev,
                        int n_tags, int32_t* tag_slot_dev, int32_t* n_seen_dev, int32_t* slot_dev, double* rng_dev,
                        double* brg_dev, int m, void* stream);
/* ekf_batch_associate with the per-filter detection noise generated INSIDE the kernel: detection d of filter b is
 * perturbed by sigma x the normal pair of (seed, first_filter + b, step, detection stream d) — the numbers
 * ekf_batch_detection_noise would write, without the two [n_det][batch] arrays                                 */
int ekf_batch_associate_generated(ekf_batch_t* h, int n_det, const int32_t* det_tag_dev, const double* det_tx_dev,
                                  const double* det_tz_dev, uint64_t seed, uint64_t step, uint64_t first_filter,
                                  double sigma, int n_tags, int32_t* tag_slot_dev, int32_t* n_seen_dev,
                                  int32_t* slot_dev, double* rng_dev, double* brg_dev, int m, void* stream);
/* windows that held more than m distinct landmarks since create: the extra detections are dropped BEFORE a slot
 * is handed out, so a dropped new tag is first-seen in a later window and slot order stays gap-free            */
int ekf_batch_associate_overflow_count(ekf_batch_t* h, int64_t* count);
/* sse_dev[b] += (x_b - gt_x)^2 + (y_b - gt_y)^2 for every filter: absolute trajectory error of the ensemble
 * against a ground-truth track (the comparison of replay_no_ros.py:519-529 as a number per member)          */
int ekf_batch_accumulate_ate(ekf_batch_t* h, double gt_x, double gt_y, double* sse_dev, void* stream);
/* ensemble moments of the state over this handle's filters, fixed summation order (no atomics):
 * sums_dev[i] = sum_b mu_b[i] (i < n), sums_dev[n + i*n + k] = sum_b mu_b[i] mu_b[k]; n + n*n doubles.
 * The one optional reduction of the batched mode: ranks all-reduce these n + n*n doubles
 * (plus the filter count) to get the ensemble mean and covariance that the reference's ablation plots show per
 * run (README.md:421-492); the filters themselves never interact.                                            */
int ekf_batch_moments(ekf_batch_t* h, double* sums_dev, void* stream);
/* the same for the robot pose only: sums_dev[0..2] = sum_b (x, y, theta), sums_dev[3 + 3 i + k] = sum_b pose_i pose_k
 * (12 doubles) — the per-step result of a Monte Carlo replay (mean trajectory and its spread)                  */
int ekf_batch_pose_moments(ekf_batch_t* h, double* sums_dev, void* stream);
/* NO COUNTERPART IN THE REFERENCE: added for callers that may not allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The result of
 * ekf_batch_pose_moments — bit for bit, same kernels and summation order — for batch records of mu_stride doubles
 * at mu_dev (record b starts with x, y, theta at mu_dev + b * mu_stride), 12 doubles into sums_dev, on `device`
 * and `stream`. Takes no handle and calls no cudaMalloc / cudaFree / allocating routine: the per-CTA partial sums
 * live in a file-scope device array of 592 x 9 doubles (42,624 bytes) that every GPU loading the library holds until
 * exit. Limit: batch <= EKF_POSE_MOMENTS_NOALLOC_MAX_FILTERS, beyond it EKF_EINVAL. All calls on one device share
 * the array (it keeps the last call's contents; every slot a call reads it has written itself, so nothing needs
 * clearing): calls that OVERLAP on the GPU (different streams, or graphs replayed concurrently) race — serialise
 * them. The first call on a device resolves the array's address (cudaGetSymbolAddress); make it before capture. */
#define EKF_POSE_MOMENTS_NOALLOC_MAX_FILTERS 4194304
int ekf_pose_moments_noalloc(const double* mu_dev, int64_t mu_stride, int64_t batch, double* sums_dev, int device,
                             void* stream);

/* ---- Monte Carlo perturbations generated ON the device ---------------------------------------------------------
 * The batched mode replays ONE common trajectory per ensemble member with member-specific perturbations (north star:
 * "Monte Carlo noise replays of the same bag"). Every perturbation is a pure function of (seed, GLOBAL filter index
 * first_filter + b, step, stream): Philox4x32-10 -> two 23-bit uniforms -> a pair of standard normals by Box-Muller in
 * single precision (separately rounded IEEE operations only, widened to double before scaling; tails to 8 sigma), so the
 * host can regenerate any member's inputs bit for bit (oracle/mc_noise.py) whatever the sharding.
 * The common step record is: doubles [ang, lin, rng[0..m), brg[0..m)] followed by int32 slot[0..m)
 * (ekf_batch_mc_record_doubles(m) doubles in all). sigma4 = std-dev of the perturbation of (ang, lin, rng, brg).  */
int ekf_batch_mc_record_doubles(int m, int64_t* doubles);
/* writes the [batch] / [batch][m] blocks ekf_batch_step consumes from a DEVICE record                            */
int ekf_batch_perturb(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                      const double* record_dev, const double* sigma4, double* ang_dev, double* lin_dev,
                      int32_t* slot_dev, double* rng_dev, double* brg_dev, void* stream);
/* the fused filter-step with the per-member inputs drawn INSIDE the step kernel (the numbers ekf_batch_perturb would
 * write, generated by m + 1 lanes while the TMA load of the filter's state is in flight): no per-filter input arrays  */
int ekf_batch_step_generated(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                             const double* record_dev, const double* sigma4, void* stream);
/* per-filter perturbation of the window's detections for ekf_batch_associate: noise_tx / noise_tz [n_det][batch]  */
int ekf_batch_detection_noise(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int n_det,
                              double sigma, double* noise_tx_dev, double* noise_tz_dev, void* stream);
/* One pipelined Monte Carlo step from a HOST record (same pipeline, stream contract and ekf_batch_host_sync as
 * ekf_batch_step_host_async): H2D of the record (< 200 bytes), the fused step drawing its inputs in the kernel, optionally
 * sse_dev[b] += squared position error against (gt_x, gt_y) when score != 0, and the 12 pose moments copied to
 * pose_moments_host (may be NULL). Per step the host sends the common record and receives 96 bytes. ATE and the moments
 * are formed from a compact pose copy on the copy-out stream while the next step runs: sse_dev and pose_moments_host
 * are complete after ekf_batch_host_sync.                                                                      */
int ekf_batch_mc_step_host_async(ekf_batch_t* h, uint64_t seed, uint64_t step, uint64_t first_filter, int m,
                                 const double* record_host, const double* sigma4, int score, double gt_x, double gt_y,
                                 double* sse_dev, double* pose_moments_host, void* stream);

/* ---- large-map mode: ONE filter, P streamed from HBM, optional row-block shard ------------------------- */
typedef struct ekf_lm ekf_lm_t;

int ekf_lm_layout(int capacity_landmarks, int* n, int64_t* ld);
int ekf_lm_create(int capacity_landmarks, int device, ekf_lm_t** out);
int ekf_lm_destroy(ekf_lm_t* h);
/* mu_dev [ld]; p_dev holds rows [row_begin, row_end) of P, row-major with leading dimension ld.
 * Single GPU: row_begin = 0, row_end = n. Pad entries must be zero.                                     */
int ekf_lm_bind_state(ekf_lm_t* h, double* mu_dev, double* p_dev, int64_t ld, int row_begin, int row_end);
/* Reference layout on the HOST (mean (m,), covariance (m,m) C order, m <= n) <-> the bound device state, one call
 * each: upload zeroes the rest of the capacity (landmarks not yet seen) and copies this block's rows; download copies
 * mean[0..m) and the block's live rows [row_begin, min(row_end, m)) x m, and is synchronous on `stream`.          */
int ekf_lm_upload(ekf_lm_t* h, int m, const double* mean_host, const double* cov_host, void* stream);
int ekf_lm_download(ekf_lm_t* h, int m, double* mean_host, double* cov_host, void* stream);
int ekf_lm_set_noise(ekf_lm_t* h, double motion_var, double meas_var);
int ekf_lm_set_flags(ekf_lm_t* h, int circular, int measurement, int disable_motion, int cov_form);
/* a5: new landmark `slot` at (gx, gy): mu entries + 10000 on the two new diagonals (:355-360)            */
int ekf_lm_augment(ekf_lm_t* h, int slot, double gx, double gy, void* stream);
/* a6-a7: odometry predict on the robot block rows/columns of P (:368-430)                               */
int ekf_lm_predict(ekf_lm_t* h, double ang, double lin, void* stream);
/* a8-a11: one range-bearing update of landmark `slot` (:436-480)                                        */
int ekf_lm_update(ekf_lm_t* h, int slot, double rng, double brg, void* stream);
/* The m sequential updates of one step (the reference's update loop :436-480, in the given order) with ONE read +
 * write of P: the 5 rows / 5 columns of update u are rebuilt from the step's initial P and the earlier (K, H P)
 * records, then all m rank-2 terms are applied per element in order. Bitwise equal to m ekf_lm_update calls.
 * slots / rng / brg are HOST arrays of length m; groups of 4 are fused; on a row-block shard or with the Joseph
 * form the call falls back to the sequential kernels.                                                   */
int ekf_lm_update_multi(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                        const double* brg_host, void* stream);
/* Fused step on a row-block shard: export the locally owned ones of the rows {0,1,2} U {j_u, j_u+1 : u < m} (+ mu
 * entries at column ld) into rows_dev [3 + 2m][ld + 4]; after the owners' rows were exchanged, run the m updates on
 * the local block with one read + write of it. 2 <= m <= 4. Sharded result == unsharded ekf_lm_update_multi bitwise. */
int ekf_lm_export_rows_multi(ekf_lm_t* h, int m, const int32_t* slots_host, double* rows_dev, void* stream);
int ekf_lm_update_multi_from_rows(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                                  const double* brg_host, const double* rows_dev, void* stream);
/* Row-block shards: export this shard's share of rows {0,1,2,j,j+1} (+ mu entries) into rows5_dev
 * [5][ld+4] (rows it does not own are left untouched), then — after the owners' rows were exchanged
 * (NCCL broadcast by the host) — run the update on the local block from the replicated buffer.          */
int ekf_lm_export_rows(ekf_lm_t* h, int slot, double* rows5_dev, void* stream);
int ekf_lm_update_from_rows(ekf_lm_t* h, int slot, double rng, double brg, const double* rows5_dev,
                            void* stream);
/* Row exchange over NVLink peer memory instead of host-driven broadcasts (one process per GPU, all stream-ordered):
 * ekf_lm_p2p_init allocates this shard's exchange area and returns its cudaIpc handle (handle_bytes >= 64); the host
 * gathers the handles of all ranks (any transport) and passes them to ekf_lm_p2p_connect ([world][handle_bytes];
 * local_areas optionally gives device pointers of shards living in the same process). ekf_lm_p2p_update then runs
 * one sharded update in the SAME two launches as on one GPU: at the head of the gather kernel the owner of rows
 * {j,j+1} stores them (+ mu entries) into every peer's area as 16-byte cells tagged with the update's sequence
 * number, and every consumer thread polls the cells it needs in its own memory (no flag, no fence); rank 0's
 * streaming kernel stores the rows {0,1,2} it has just updated into the buffer of the NEXT update (a predict, or
 * ekf_lm_p2p_invalidate after an external write to P, makes rank 0's next gather push them instead). With the
 * Joseph form every rank also stores K of its rows into every peer's area (W needs K at every column). With world
 * = 1 the call is ekf_lm_update. Every rank must make the same calls in the same order. owner = the rank whose row
 * block holds rows j, j+1 (any contiguous partition on landmark pairs; EKF_EINVAL if this rank is named but does
 * not hold them). ekf_lm_p2p_error reports a spin that ran out of its 2 s budget (code << 32 | sequence number;
 * code 1/2: acks before a push / pre-push, 3: a cell that never arrived).                                       */
int ekf_lm_p2p_init(ekf_lm_t* h, int world, int rank, unsigned char* handle_out, int64_t handle_bytes);
int ekf_lm_p2p_connect(ekf_lm_t* h, const unsigned char* handles, int64_t handle_bytes, void* const* local_areas);
int ekf_lm_p2p_area(ekf_lm_t* h, void** area_dev);
int ekf_lm_p2p_update(ekf_lm_t* h, int slot, double rng, double brg, int owner, void* stream);
/* the fused step (2 <= m <= 4 updates, one pass over the local row block): the fused gather stores / polls its
 * 1 + m row parts (owners_host[u] = rank owning rows j_u, j_u+1), the streaming kernel pre-pushes rows {0,1,2}.
 * Joseph form: the step's updates run one after the other through ekf_lm_p2p_update                            */
int ekf_lm_p2p_update_multi(ekf_lm_t* h, int m, const int32_t* slots_host, const double* rng_host,
                            const double* brg_host, const int32_t* owners_host, void* stream);
int ekf_lm_p2p_invalidate(ekf_lm_t* h);
int ekf_lm_p2p_error(ekf_lm_t* h, int64_t* flag);
int ekf_lm_sync(ekf_lm_t* h);

/* ---- 3-state localisation EKF of src/EKF-SLAM.py, batched, SoA: state [3][B], cov [9][B] --------------- */
int ekf3_predict(int64_t batch, double* state_dev, double* cov_dev, const double* v_dev,
                 const double* omega_dev, const double* dt_dev, void* stream);
int ekf3_update(int64_t batch, double* state_dev, double* cov_dev, const double* rng_dev,
                const double* brg_dev, const double* lx_dev, const double* ly_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EKF_B200_H */
