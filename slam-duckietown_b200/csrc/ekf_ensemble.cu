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
#include "ekf_common.cuh"
#include "ekf_ensemble.h"

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
    if (a.noise_tx) {
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
