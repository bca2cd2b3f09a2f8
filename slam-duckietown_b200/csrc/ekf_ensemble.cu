// ekf_ensemble.cu — the callers' side of the batched Monte Carlo step, on the device (sm_100a):
//
//  * batch_associate_kernel — a3/a4 of the reference (src/replay_no_ros.py:280-337) per filter: gate each detection
//    of the window on its (per-filter perturbed) camera-frame position, assign landmark slots in first-seen order
//    (the filter's own TAG_INDEX), average the window per slot and convert to range / bearing. Output is exactly the
//    [B][m] (slot, range, bearing) block the fused step consumes, so per-filter measurement noise needs no host
//    round trip. Integer bookkeeping (gate decisions, slot tables, update order) is bit-exact
//    with the host association: the gate and the sums use explicitly rounded IEEE operations (no FMA contraction).
//  * batch_ate_kernel — accumulates each filter's squared position error against a ground-truth track
//    (the README's trajectory comparison, replay_no_ros.py:519-529, as a number per ensemble member; §8f-4).
//
// One THREAD per filter: the work per filter is a few dozen scalar operations; every per-filter array is laid out
// [item][B] so that consecutive threads touch consecutive addresses (coalesced), except the [B][m] output block,
// whose layout is the step kernel's.
#include <mutex>

#include "ekf_common.cuh"
#include "ekf_ensemble.h"
#include "ekf_philox.cuh"

namespace ekf {

__global__ void __launch_bounds__(128) batch_associate_kernel(AssocArgs a) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= a.batch) return;
  const int m = a.m;
  int bslot[ASSOC_MAX_M];
  double sx[ASSOC_MAX_M], sz[ASSOC_MAX_M];
  int cnt[ASSOC_MAX_M];
  int nb = 0;                      // buckets in use, first-seen-in-window order (:313 iterates the dict in that order)
  int seen = a.n_seen[b];          // len(TAG_INDEX) of this filter
  bool overflow = false;
  for (int d = 0; d < a.n_det; ++d) {
    const int tag = a.det_tag[d];  // compact tag index; < 0 = ignored tag (IGNORE_TAGS, :286)
    if (tag < 0 || tag >= a.n_tags) continue;
    double tx = a.det_tx[d], tz = a.det_tz[d];
    if (a.gen) {  // the same numbers ekf_batch_detection_noise would have written: sigma * z, added with one rounding
      double z0, z1;
      normal_pair(a.gen_seed, a.gen_first + (uint64_t)b, a.gen_step, kStreamDetection0 + d, &z0, &z1);
      tx = __dadd_rn(tx, __dmul_rn(a.gen_sigma, z0));
      tz = __dadd_rn(tz, __dmul_rn(a.gen_sigma, z1));
    } else if (a.noise_tx) {
      tx = __dadd_rn(tx, a.noise_tx[(int64_t)d * a.batch + b]);
      tz = __dadd_rn(tz, a.noise_tz[(int64_t)d * a.batch + b]);
    }
    // pose_t[2]^2 + pose_t[0]^2 > 1.5^2 drops the detection (:289), strict, on the raw per-detection value
    if (__dadd_rn(__dmul_rn(tz, tz), __dmul_rn(tx, tx)) > 2.25) continue;
    int32_t* cell = a.tag_slot + (int64_t)tag * a.batch + b;
    int s = *cell;
    int k = 0;
    if (s >= 0)
      while (k < nb && bslot[k] != s) ++k;
    else
      k = nb;                      // a tag never seen before cannot have a bucket yet
    if (k == nb) {
      if (nb == m) {               // more distinct landmarks in one window than the step block holds: the
        overflow = true;           // detection is dropped BEFORE a slot is handed out, so a dropped new tag is
        continue;                  // simply "first seen" in a later window and slot order stays gap-free
      }
      if (s < 0) {                 // first sighting ever: slot = len(TAG_INDEX) (:294-295)
        s = seen++;
        *cell = s;
      }
      bslot[nb] = s; sx[nb] = 0.0; sz[nb] = 0.0; cnt[nb] = 0;
      ++nb;
    }
    sx[k] = __dadd_rn(sx[k], tx);  // sequential sums == np.mean over the window axis (:311)
    sz[k] = __dadd_rn(sz[k], tz);
    ++cnt[k];
  }
  a.n_seen[b] = seen;
  if (overflow) atomicAdd(a.overflow, 1ULL);
  for (int k = 0; k < m; ++k) {
    int s = -1;
    double r = 0.0, be = 0.0;
    if (k < nb) {
      const double n = (double)cnt[k];
      const double x = sz[k] / n, y = -(sx[k] / n);              // x = t[2], y = -t[0] (:316)
      r = sqrt(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)));      // :327
      be = atan2(y, x);                                           // :328
      s = bslot[k];
    }
    a.slot[b * m + k] = s;
    a.rng[b * m + k] = r;
    a.brg[b * m + k] = be;
  }
}

__global__ void __launch_bounds__(256) batch_ate_kernel(const double* __restrict__ mu, int64_t mu_stride,
                                                        int64_t batch, double gx, double gy,
                                                        double* __restrict__ sse) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const double ex = mu[b * mu_stride] - gx, ey = mu[b * mu_stride + 1] - gy;
  sse[b] += ex * ex + ey * ey;
}

// Ensemble moments of the state (batched mode: the one optional reduction of the Monte Carlo batch):
// sums[i] = sum_b mu_b[i], sums[n + i n + k] = sum_b mu_b[i] mu_b[k]. Two passes with a FIXED summation order (no
// atomics): CTA c reduces its contiguous chunk of filters, staging MOM_TILE mean vectors at a time in shared memory
// (coalesced rows), thread e owning output element e (and e + 256, ...) with one running sum per element; the
// second kernel adds the MOM_CTAS partial vectors in index order. mu is 0.35 KB per filter, so the pass costs
// 1/40th of a filter-step's traffic.
constexpr int MOM_TILE = 32;
__global__ void __launch_bounds__(256) batch_moments_partial_kernel(const double* __restrict__ mu, int64_t mu_stride,
                                                                    int64_t batch, int n, int64_t chunk,
                                                                    double* __restrict__ partial) {
  extern __shared__ double tile[];  // [MOM_TILE][n]
  const int E = n + n * n;
  const int64_t b0 = blockIdx.x * chunk, b1 = (b0 + chunk < batch) ? b0 + chunk : batch;
  constexpr int MAXE = 16;  // elements per thread: n <= 63 -> E <= 4032 <= 16 x 256
  double acc[MAXE];
#pragma unroll
  for (int q = 0; q < MAXE; ++q) acc[q] = 0.0;
  for (int64_t t0 = b0; t0 < b1; t0 += MOM_TILE) {
    const int nb = (int)((b1 - t0 < MOM_TILE) ? b1 - t0 : MOM_TILE);
    __syncthreads();
    for (int x = threadIdx.x; x < nb * n; x += blockDim.x) {
      const int f = x / n, i = x - f * n;
      tile[f * n + i] = mu[(t0 + f) * mu_stride + i];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < MAXE; ++q) {
      const int e = threadIdx.x + q * 256;
      if (e >= E) break;
      int i, k;
      if (e < n) { i = e; k = -1; } else { i = (e - n) / n; k = (e - n) - i * n; }
      double a = acc[q];
      for (int f = 0; f < nb; ++f) {
        const double vi = tile[f * n + i];
        a = (k < 0) ? __dadd_rn(a, vi) : fma(vi, tile[f * n + k], a);
      }
      acc[q] = a;
    }
  }
#pragma unroll
  for (int q = 0; q < MAXE; ++q) {
    const int e = threadIdx.x + q * 256;
    if (e < E) partial[(int64_t)blockIdx.x * E + e] = acc[q];
  }
}

__global__ void __launch_bounds__(256) batch_moments_final_kernel(const double* __restrict__ partial, int ctas, int E,
                                                                  double* __restrict__ sums) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  double a = 0.0;
  for (int c = 0; c < ctas; ++c) a = __dadd_rn(a, partial[(int64_t)c * E + e]);
  sums[e] = a;
}

// Health check on demand (not on the hot path): counts[0] = filters whose mu or P holds a non-finite value (the
// reference would carry inf / nan silently, e.g. q = 0 when a landmark estimate coincides with the pose, :446-469).
// One warp per filter, coalesced reads of its record.
__global__ void __launch_bounds__(256) batch_nonfinite_kernel(const double* __restrict__ mu, const double* __restrict__ P,
                                                              int64_t mu_stride, int64_t p_stride, int64_t batch,
                                                              unsigned long long* __restrict__ count) {
  const int64_t b = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= batch) return;
  bool bad = false;
  for (int64_t i = lane; i < mu_stride; i += 32) bad |= !isfinite(mu[b * mu_stride + i]);
  for (int64_t i = lane; i < p_stride; i += 32) bad |= !isfinite(P[b * p_stride + i]);
  if (__any_sync(0xffffffffu, bad) && lane == 0) atomicAdd(count, 1ULL);
}

cudaError_t batch_launch_nonfinite(const double* mu, const double* P, int64_t mu_stride, int64_t p_stride,
                                   int64_t batch, unsigned long long* count, cudaStream_t st) {
  const int64_t threads = batch * 32;
  batch_nonfinite_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(mu, P, mu_stride, p_stride, batch, count);
  return cudaGetLastError();
}

cudaError_t batch_launch_moments(const double* mu, int64_t mu_stride, int64_t batch, int n, double* partial,
                                 double* sums, cudaStream_t st) {
  if (n < 1 || n > 63 || batch < 1) return cudaErrorInvalidValue;
  const int E = n + n * n;
  const int64_t chunk = (batch + MOM_CTAS - 1) / MOM_CTAS;
  const int ctas = (int)((batch + chunk - 1) / chunk);
  batch_moments_partial_kernel<<<ctas, 256, MOM_TILE * n * sizeof(double), st>>>(mu, mu_stride, batch, n, chunk,
                                                                                  partial);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  batch_moments_final_kernel<<<(E + 255) / 256, 256, 0, st>>>(partial, ctas, E, sums);
  return cudaGetLastError();
}


// -------------------------------------------------------------------------------------------------------
// Monte Carlo perturbations on the device. One thread per (filter, stream): stream 0 -> (ang, lin), stream 1 + k ->
// (rng_k, brg_k). ~60 integer + ~50 fp64 instructions per normal, no memory traffic besides the outputs
// (16 + 20 m bytes per filter, < 1 % of a filter-step's bytes).
// -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) batch_perturb_kernel(PerturbArgs a) {
  const int per = 1 + a.m;
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t b = tid / per;
  if (b >= a.batch) return;
  const int q = (int)(tid - b * per);
  double z0, z1;
  normal_pair(a.seed, a.first_filter + (uint64_t)b, a.step, q == 0 ? kStreamOdometry : kStreamObservation0 + (q - 1),
              &z0, &z1);
  if (q == 0) {
    a.ang[b] = __dadd_rn(a.record[0], __dmul_rn(a.s_ang, z0));
    a.lin[b] = __dadd_rn(a.record[1], __dmul_rn(a.s_lin, z1));
  } else {
    const int k = q - 1;
    const int32_t* slots = reinterpret_cast<const int32_t*>(a.record + 2 + 2 * a.m);
    a.slot[b * a.m + k] = slots[k];
    a.rng[b * a.m + k] = __dadd_rn(a.record[2 + k], __dmul_rn(a.s_rng, z0));
    a.brg[b * a.m + k] = __dadd_rn(a.record[2 + a.m + k], __dmul_rn(a.s_brg, z1));
  }
}

cudaError_t batch_launch_perturb(const PerturbArgs& a, cudaStream_t st) {
  if (a.m < 0 || a.m > ASSOC_MAX_M) return cudaErrorInvalidValue;
  const int64_t threads = a.batch * (1 + a.m);
  batch_perturb_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(a);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) batch_detection_noise_kernel(int64_t batch, uint64_t seed, uint64_t first_filter,
                                                                    uint32_t step, int n_det, double sigma,
                                                                    double* __restrict__ ntx, double* __restrict__ ntz) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int d = blockIdx.y;
  if (b >= batch || d >= n_det) return;
  double z0, z1;
  normal_pair(seed, first_filter + (uint64_t)b, step, kStreamDetection0 + d, &z0, &z1);
  ntx[(int64_t)d * batch + b] = __dmul_rn(sigma, z0);
  ntz[(int64_t)d * batch + b] = __dmul_rn(sigma, z1);
}

cudaError_t batch_launch_detection_noise(int64_t batch, uint64_t seed, uint64_t first_filter, uint32_t step, int n_det,
                                         double sigma, double* noise_tx, double* noise_tz, cudaStream_t st) {
  if (n_det <= 0) return cudaSuccess;
  dim3 grid((unsigned)((batch + 255) / 256), (unsigned)n_det);
  batch_detection_noise_kernel<<<grid, 256, 0, st>>>(batch, seed, first_filter, step, n_det, sigma, noise_tx, noise_tz);
  return cudaGetLastError();
}

// -------------------------------------------------------------------------------------------------------
// Ensemble moments of the robot pose (the per-step result of a Monte Carlo replay: mean trajectory + its spread).
// Pass 1: thread per filter (grid-stride), 9 running sums (3 first, 6 distinct second moments), fixed-order warp
// shuffle tree, then the warps of the CTA in index order. Pass 2: one CTA adds the per-CTA partials, lane-strided
// then by the same shuffle tree. No atomics: two calls give identical bits.
// -------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_tree_sum(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, off));
  return v;
}

__global__ void __launch_bounds__(256) batch_pose_moments_partial_kernel(const double* __restrict__ mu, int64_t mu_stride,
                                                                         int64_t batch, double* __restrict__ partial) {
  __shared__ double sw[8][9];
  double acc[9];
#pragma unroll
  for (int q = 0; q < 9; ++q) acc[q] = 0.0;
#pragma unroll 4
  for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < batch; b += (int64_t)gridDim.x * blockDim.x) {
    const double x = mu[b * mu_stride], y = mu[b * mu_stride + 1], t = mu[b * mu_stride + 2];
    acc[0] = __dadd_rn(acc[0], x);
    acc[1] = __dadd_rn(acc[1], y);
    acc[2] = __dadd_rn(acc[2], t);
    acc[3] = fma(x, x, acc[3]);
    acc[4] = fma(x, y, acc[4]);
    acc[5] = fma(x, t, acc[5]);
    acc[6] = fma(y, y, acc[6]);
    acc[7] = fma(y, t, acc[7]);
    acc[8] = fma(t, t, acc[8]);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 9; ++q) {
    const double v = warp_tree_sum(acc[q]);
    if (lane == 0) sw[warp][q] = v;
  }
  __syncthreads();
  if (threadIdx.x < 9) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) v = __dadd_rn(v, sw[w][threadIdx.x]);
    partial[(int64_t)blockIdx.x * 9 + threadIdx.x] = v;
  }
}

__global__ void __launch_bounds__(288) batch_pose_moments_final_kernel(const double* __restrict__ partial, int ctas,
                                                                       double* __restrict__ sums) {
  const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;  // warp q sums element q
  double v = 0.0;
  for (int c = lane; c < ctas; c += 32) v = __dadd_rn(v, partial[(int64_t)c * 9 + q]);
  v = warp_tree_sum(v);
  if (lane == 0) {
    // sums[0..2] first moments; sums[3 + 3 i + k] the symmetric 3x3 second-moment block
    static const int where[9][2] = {{0, -1}, {1, -1}, {2, -1}, {3, 3}, {4, 6}, {5, 9}, {7, 7}, {8, 10}, {11, 11}};
    sums[where[q][0]] = v;
    if (where[q][1] >= 0) sums[where[q][1]] = v;
  }
}

cudaError_t batch_launch_pose_moments(const double* mu, int64_t mu_stride, int64_t batch, double* partial, double* sums,
                                      cudaStream_t st) {
  if (batch < 1) return cudaErrorInvalidValue;
  const int64_t want = (batch + 255) / 256;
  const int ctas = (int)(want < POSE_MOM_CTAS ? want : POSE_MOM_CTAS);
  batch_pose_moments_partial_kernel<<<ctas, 256, 0, st>>>(mu, mu_stride, batch, partial);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  batch_pose_moments_final_kernel<<<1, 288, 0, st>>>(partial, ctas, sums);
  return cudaGetLastError();
}

// -------------------------------------------------------------------------------------------------------
// The same two passes with the per-CTA partials in a file-scope device array instead of a cudaMalloc'ed buffer, for
// callers that may not allocate (CUDA graph capture, serving loops). Every GPU that loads the library holds the
// array (POSE_MOM_CTAS x 9 doubles = 42,624 bytes) until exit. In HOST code `g_pose_partial_static` is only a
// stand-in for the symbol: its device address is resolved once per device with cudaGetSymbolAddress (that device
// current) and cached; the kernels get the resolved pointer. All calls on a device share the array: pass 2 reads
// only the slots pass 1 of the same call wrote, so nothing needs clearing, but calls that overlap on the GPU race.
// -------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_pose_partial_static[POSE_MOM_CTAS * 9];

namespace {
constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mu;
double* g_pose_partial_addr[kMaxStaticDevices] = {};
}  // namespace

cudaError_t batch_launch_pose_moments_static(const double* mu, int64_t mu_stride, int64_t batch, double* sums,
                                             int device, cudaStream_t st) {
  if (device < 0 || device >= kMaxStaticDevices) return cudaErrorInvalidDevice;
  if (batch < 1 || batch > POSE_MOM_STATIC_MAX_FILTERS) return cudaErrorInvalidValue;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return e;
  double* partial = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_pose_partial_addr[device]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_pose_partial_static);  // this device's copy of the array
      if (e != cudaSuccess) return e;
      g_pose_partial_addr[device] = static_cast<double*>(p);
    }
    partial = g_pose_partial_addr[device];
  }
  return batch_launch_pose_moments(mu, mu_stride, batch, partial, sums, st);
}

cudaError_t batch_launch_associate(const AssocArgs& a, cudaStream_t st) {
  if (a.m < 1 || a.m > ASSOC_MAX_M) return cudaErrorInvalidValue;
  batch_associate_kernel<<<(unsigned)((a.batch + 127) / 128), 128, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t batch_launch_ate(const double* mu, int64_t mu_stride, int64_t batch, double gx, double gy, double* sse,
                             cudaStream_t st) {
  batch_ate_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, st>>>(mu, mu_stride, batch, gx, gy, sse);
  return cudaGetLastError();
}

}  // namespace ekf
