// ekf_largemap_multi.cu — large-map filter: the m sequential range-bearing updates of ONE step fused into a
// single pass over P (reference update loop src/replay_no_ros.py:436-480).
//
// The reference applies the updates one after the other, each a dense (I - K H) P product. In rank form update u
// touches P only through its five rows / five columns idx5(u) = {0, 1, 2, j_u, j_u + 1}; those 10 n entries of the
// INTERMEDIATE covariance P_u can be rebuilt from the step's initial P0 and the earlier (K_w, W_w = H_w P_w):
//     P_u[i, c] = P0[i, c] - sum_{w < u} K_w[i, :] W_w[:, c]        (applied in order, same rounding sequence)
// so the O(n) part of every update (innovation, S^-1, K_u, W_u, the five mu entries) runs first — m small gather
// kernels, update u costing O(n (10 + 4u)) — and ONE streaming kernel then applies all m rank-2 terms to every
// element in order. HBM traffic per step drops from m x 16 n^2 to 16 n^2 bytes; the per-element operation
// sequence is that of the sequential kernels (lm_gather_kernel + lm_rank_update_kernel), so the result is bitwise
// the same (tests/test_gpu_parity.py::test_large_map_fused_multi_update_bitwise).
// Plain covariance form; whole matrix on one GPU or a row-block shard fed by exchanged rows.
#include "ekf_common.cuh"
#include "ekf_largemap.h"

namespace ekf {

namespace {

__device__ __forceinline__ double rank2e(double v, double k0, double k1, double w0, double w1) {
  return fma(-k1, w1, fma(-k0, w0, v));  // == lm_rank_update_kernel's fma(-k1, w1, fma(-k0, w0, v))
}

}  // namespace

// exchanged-row bookkeeping of a fused step: R = {0, 1, 2} U {j_u, j_u + 1 : u < m}; rho indexes R
__device__ __forceinline__ int rho_row(const LmMultiArgs& a, int rho) {
  return rho < 3 ? rho : a.j[(rho - 3) >> 1] + ((rho - 3) & 1);
}
// row rho of the step's INITIAL covariance: the exchange buffer on a row-block shard, P itself otherwise
__device__ __forceinline__ const double* rho_ptr(const LmMultiArgs& a, int rho) {
  return a.xrows ? a.xrows + (int64_t)rho * a.ldx : a.P + (int64_t)rho_row(a, rho) * a.ld;
}

// a8-a9 for update u of the step, on the covariance as it stands after updates 0..u-1 (never materialised).
// Every rank computes W_u over all n columns (from the five full rows) and K_u for its local rows (from their five
// columns); K_u at the exchanged rows R — which later updates need as uniform operands on EVERY rank — is rebuilt
// from the exchanged rows themselves by block 0 (a.KR), with the operation sequence of the owning thread.
//
// The O(1) part (five mu entries, innovation, the 5 x 5 block of P_u, S^-1) is on the critical path of the whole
// step — m of these kernels run back to back before the streaming pass — so it is spread over the CTA instead of
// being one thread's serial chain of dependent global loads: threads rho < 3 + 2m of warp 0 rebuild the five
// columns idx5(u) of exchanged row rho (all loads in flight together), lanes 0-4 of warp 1 rebuild the five mu
// entries and lane 0 of warp 1 runs the innovation meanwhile; thread 0 then only contracts shared-memory values.
// Every value keeps the operation sequence it had in the one-thread version, so results are unchanged bitwise.
__global__ void __launch_bounds__(256) lm_gather_multi_kernel(LmMultiArgs a, int u) {
  __shared__ double sh[2][5];
  __shared__ double ssi[2][2];
  __shared__ double skw[LM_MAX_FUSED][5][2];  // K_w[idx5(u), :] of the earlier updates
  __shared__ double sww[LM_MAX_FUSED][2][5];  // W_w[:, idx5(u)]
  __shared__ double syw[LM_MAX_FUSED][2];     // innovations of the earlier updates
  __shared__ double spk[LM_MAX_ROWS][5];      // P_u[R, idx5(u)]
  __shared__ double smu[5];
  const int j = a.j[u];
  const int idx5[5] = {0, 1, 2, j, j + 1};
  const int rho5[5] = {0, 1, 2, 3 + 2 * u, 4 + 2 * u};
  const int nr = 3 + 2 * a.m;
  const int tid = threadIdx.x;
  pdl_wait();
  if (a.wait_flags && u == 0) {  // peer-memory exchange: the step's 1 + m row parts carry this step's number
    if (tid == 0) {
      const volatile unsigned long long* f = a.wait_flags;
      const long long t0 = clock64();
      for (int part = 0; part <= a.m; ++part)
        while (f[part] < a.wait_seq) {
          if (clock64() - t0 > 4000000000LL) {
            atomicExch(a.wait_err, 1ULL);
            break;
          }
          __nanosleep(32);
        }
      __threadfence_system();
    }
    __syncthreads();
  }
  if (tid < 5 * u) {  // uniform operands of the rebuild
    const int w = tid / 5, k = tid - 5 * w;
    skw[w][k][0] = a.KR[(w * LM_MAX_ROWS + rho5[k]) * 2];
    skw[w][k][1] = a.KR[(w * LM_MAX_ROWS + rho5[k]) * 2 + 1];
    sww[w][0][k] = a.W[(2 * w) * a.ldw + idx5[k]];
    sww[w][1][k] = a.W[(2 * w + 1) * a.ldw + idx5[k]];
  } else if (tid >= 64 && tid < 64 + 2 * u) {
    syw[(tid - 64) >> 1][tid & 1] = a.y[tid - 64];
  }
  // operands that do not depend on the staged values are requested before the barrier
  const int t = blockIdx.x * blockDim.x + tid;
  const bool live = t < a.n;
  const bool local = live && t >= a.row_begin && t < a.row_end;
  double pr[5], pc[5];
  double kt0[LM_MAX_FUSED], kt1[LM_MAX_FUSED], wt0[LM_MAX_FUSED], wt1[LM_MAX_FUSED];
  if (live) {
#pragma unroll
    for (int w = 0; w < LM_MAX_FUSED; ++w)
      if (w < u) {
        kt0[w] = local ? a.K[(2 * w) * a.ldw + t] : 0.0;
        kt1[w] = local ? a.K[(2 * w + 1) * a.ldw + t] : 0.0;
        wt0[w] = a.W[(2 * w) * a.ldw + t];
        wt1[w] = a.W[(2 * w + 1) * a.ldw + t];
      }
    const double* prow = a.P + (int64_t)(t - a.row_begin) * a.ld;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      pr[k] = rho_ptr(a, rho5[k])[t];            // row idx5[k], column t
      pc[k] = local ? prow[idx5[k]] : 0.0;       // row t, column idx5[k]
    }
  }
  const bool kr_thread = tid < nr;  // exchanged row rho = tid: its five columns idx5(u)
  double pk[5], kr0[LM_MAX_FUSED], kr1[LM_MAX_FUSED];
  if (kr_thread) {
    const double* xr = rho_ptr(a, tid);
#pragma unroll
    for (int k = 0; k < 5; ++k) pk[k] = xr[idx5[k]];
#pragma unroll
    for (int w = 0; w < LM_MAX_FUSED; ++w)
      if (w < u) {
        kr0[w] = a.KR[(w * LM_MAX_ROWS + tid) * 2];
        kr1[w] = a.KR[(w * LM_MAX_ROWS + tid) * 2 + 1];
      }
  }
  double mu0 = 0.0;
  const bool mu_thread = tid >= 32 && tid < 37;
  if (mu_thread) {
    // the step's initial mean: riding at column ld of the exchanged row on a shard (local copies of non-local
    // entries are stale there), mu itself otherwise (it is written by the streaming kernel only)
    const int k = tid - 32;
    mu0 = a.xrows ? rho_ptr(a, rho5[k])[a.ld] : a.mu[idx5[k]];
  }
  __syncthreads();
  if (live) {
#pragma unroll
    for (int k = 0; k < 5; ++k)
#pragma unroll
      for (int w = 0; w < LM_MAX_FUSED; ++w)
        if (w < u) {
          pr[k] = rank2e(pr[k], skw[w][k][0], skw[w][k][1], wt0[w], wt1[w]);
          pc[k] = rank2e(pc[k], kt0[w], kt1[w], sww[w][0][k], sww[w][1][k]);
        }
  }
  if (kr_thread) {
#pragma unroll
    for (int w = 0; w < LM_MAX_FUSED; ++w)
      if (w < u) {
#pragma unroll
        for (int k = 0; k < 5; ++k) pk[k] = rank2e(pk[k], kr0[w], kr1[w], sww[w][0][k], sww[w][1][k]);
      }
#pragma unroll
    for (int k = 0; k < 5; ++k) spk[tid][k] = pk[k];
  }
  if (tid >= 32 && tid < 64) {  // warp 1: the five mu entries, then the innovation
    if (mu_thread) {
      const int k = tid - 32;
      double v = mu0;
      for (int w = 0; w < u; ++w) v = mu_inc(v, skw[w][k][0], skw[w][k][1], syw[w][0], syw[w][1]);
      smu[k] = v;
    }
    __syncwarp();
    if (tid == 32) {
      const Innov in = innovation(smu[0], smu[1], smu[2], smu[3], smu[4], a.rng[u], a.brg[u]);
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int k = 0; k < 5; ++k) sh[r][k] = in.h[r][k];
      if (blockIdx.x == 0) {
        a.y[2 * u] = in.y0;
        a.y[2 * u + 1] = in.y1;
      }
    }
  }
  __syncthreads();
  if (tid == 0) {
    double hp5[2][5];
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double p = spk[rho5[k]][c];
        s0 = fma(sh[0][k], p, s0);
        s1 = fma(sh[1][k], p, s1);
      }
      hp5[0][c] = s0;
      hp5[1][c] = s1;
    }
    double s[2][2], si[2][2];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        double acc = 0.0;
#pragma unroll
        for (int c = 0; c < 5; ++c) acc = fma(hp5[r][c], sh[b][c], acc);
        s[r][b] = acc;
      }
    const double qn = a.mm * a.mm;
    s[0][0] += qn;
    s[1][1] += qn;
    inv2(s, si);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      ssi[r][0] = si[r][0];
      ssi[r][1] = si[r][1];
    }
  }
  __syncthreads();
  if (kr_thread && blockIdx.x == 0) {  // K_u at exchanged row rho (a.KR), a uniform operand of later updates
    double ph0 = 0.0, ph1 = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      ph0 = fma(pk[k], sh[0][k], ph0);
      ph1 = fma(pk[k], sh[1][k], ph1);
    }
    a.KR[(u * LM_MAX_ROWS + tid) * 2] = dot2(ph0, ssi[0][0], ph1, ssi[1][0]);
    a.KR[(u * LM_MAX_ROWS + tid) * 2 + 1] = dot2(ph0, ssi[0][1], ph1, ssi[1][1]);
  }
  if (!live) return;
  double hp0 = 0.0, hp1 = 0.0, pht0 = 0.0, pht1 = 0.0;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    hp0 = fma(sh[0][k], pr[k], hp0);
    hp1 = fma(sh[1][k], pr[k], hp1);
    pht0 = fma(pc[k], sh[0][k], pht0);
    pht1 = fma(pc[k], sh[1][k], pht1);
  }
  if (local) {
    a.K[(2 * u) * a.ldw + t] = dot2(pht0, ssi[0][0], pht1, ssi[1][0]);
    a.K[(2 * u + 1) * a.ldw + t] = dot2(pht0, ssi[0][1], pht1, ssi[1][1]);
  }
  a.W[(2 * u) * a.ldw + t] = hp0;
  a.W[(2 * u + 1) * a.ldw + t] = hp1;
}

// a10-a11 for all M updates of the step in one pass: P[i, c] <- P0[i, c] - sum_u K_u[i, :] W_u[:, c] in update
// order, mu[i] += K_u[i, :] y_u in update order. Tiling as lm_rank_update_kernel: a CTA owns LM_THREADS column
// groups of 4 doubles x RB rows; W_u of a thread's 4 columns stays in registers (8 M doubles), K_u of the RB rows
// in shared memory, UNROLL rows of P in flight per thread (256-bit streaming loads, in-place 256-bit stores).
template <int RB, int UNROLL, int M>
__global__ void __launch_bounds__(LM_THREADS) lm_rank_update_multi_kernel(LmMultiArgs a) {
  __shared__ double sk[RB][2 * M];
  const int tid = threadIdx.x;
  pdl_wait();
  if (a.ack_peers && blockIdx.x == 0 && blockIdx.y == 0 && tid < a.ack_world) {
    // this launch follows the gathers on the stream: the exchange buffer of step wait_seq is no longer read here
    volatile unsigned long long* ack =
        reinterpret_cast<unsigned long long*>(a.ack_peers[tid] + a.ack_flag_off) + 2 * P2P_PARTS + a.ack_rank;
    *ack = a.wait_seq;
  }
  const int r0 = a.row_begin + blockIdx.y * RB;  // first global row of this CTA
  const int nrows = min(RB, a.row_end - r0);
  if (a.xrows && a.row_begin != 0 && blockIdx.x == 0 && blockIdx.y == 0 && tid < 3) {
    // row-block shard that does not own the robot rows: keep the replicated pose entries of mu current
    double v = a.mu[tid];
#pragma unroll
    for (int u = 0; u < M; ++u)
      v = mu_inc(v, a.KR[(u * LM_MAX_ROWS + tid) * 2], a.KR[(u * LM_MAX_ROWS + tid) * 2 + 1], a.y[2 * u], a.y[2 * u + 1]);
    a.mu[tid] = v;
  }
  for (int e = tid; e < nrows * 2 * M; e += LM_THREADS) {
    const int r = e / (2 * M), q = e - r * (2 * M);
    sk[r][q] = a.K[(int64_t)q * a.ldw + r0 + r];
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    for (int r = tid; r < nrows; r += LM_THREADS) {
      double v = a.mu[r0 + r];
#pragma unroll
      for (int u = 0; u < M; ++u) v = mu_inc(v, sk[r][2 * u], sk[r][2 * u + 1], a.y[2 * u], a.y[2 * u + 1]);
      a.mu[r0 + r] = v;
    }
  }
  double* __restrict__ Pblk = a.P + (int64_t)(r0 - a.row_begin) * a.ld;
  if ((int)blockIdx.x < a.full_tiles) {
    const int c = (blockIdx.x * LM_THREADS + tid) * 4;
    double4x w0[M], w1[M];
#pragma unroll
    for (int u = 0; u < M; ++u) {
      w0[u] = ldg256(a.W + (int64_t)(2 * u) * a.ldw + c);
      w1[u] = ldg256(a.W + (int64_t)(2 * u + 1) * a.ldw + c);
    }
    double* p = Pblk + c;
    int r = 0;
    for (; r + UNROLL <= nrows; r += UNROLL) {
      double4x v[UNROLL];
#pragma unroll
      for (int q = 0; q < UNROLL; ++q) v[q] = ldg256_stream(p + (int64_t)(r + q) * a.ld);
#pragma unroll
      for (int u = 0; u < M; ++u)
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
          const double k0 = sk[r + q][2 * u], k1 = sk[r + q][2 * u + 1];
          v[q].a = rank2e(v[q].a, k0, k1, w0[u].a, w1[u].a);
          v[q].b = rank2e(v[q].b, k0, k1, w0[u].b, w1[u].b);
          v[q].c = rank2e(v[q].c, k0, k1, w0[u].c, w1[u].c);
          v[q].d = rank2e(v[q].d, k0, k1, w0[u].d, w1[u].d);
        }
#pragma unroll
      for (int q = 0; q < UNROLL; ++q) stg256(p + (int64_t)(r + q) * a.ld, v[q]);
    }
    for (; r < nrows; ++r) {
      double4x v = ldg256_stream(p + (int64_t)r * a.ld);
#pragma unroll
      for (int u = 0; u < M; ++u) {
        const double k0 = sk[r][2 * u], k1 = sk[r][2 * u + 1];
        v.a = rank2e(v.a, k0, k1, w0[u].a, w1[u].a);
        v.b = rank2e(v.b, k0, k1, w0[u].b, w1[u].b);
        v.c = rank2e(v.c, k0, k1, w0[u].c, w1[u].c);
        v.d = rank2e(v.d, k0, k1, w0[u].d, w1[u].d);
      }
      stg256(p + (int64_t)r * a.ld, v);
    }
  } else {
    // column tail: a.tail_groups (< LM_THREADS) groups per row, (row, group) mapping
    const int tg = a.tail_groups;
    const int cbase = a.full_tiles * LM_THREADS * 4;
    for (int e = tid; e < nrows * tg; e += LM_THREADS) {
      const int r = e / tg, c = cbase + (e - r * tg) * 4;
      double4x v = ldg256_stream(Pblk + (int64_t)r * a.ld + c);
#pragma unroll
      for (int u = 0; u < M; ++u) {
        const double4x w0 = ldg256(a.W + (int64_t)(2 * u) * a.ldw + c);
        const double4x w1 = ldg256(a.W + (int64_t)(2 * u + 1) * a.ldw + c);
        const double k0 = sk[r][2 * u], k1 = sk[r][2 * u + 1];
        v.a = rank2e(v.a, k0, k1, w0.a, w1.a);
        v.b = rank2e(v.b, k0, k1, w0.b, w1.b);
        v.c = rank2e(v.c, k0, k1, w0.c, w1.c);
        v.d = rank2e(v.d, k0, k1, w0.d, w1.d);
      }
      stg256(Pblk + (int64_t)r * a.ld + c, v);
    }
  }
}

// shard export of a fused step: the locally owned ones of the rows R (and their mu entries at column ld) into the
// exchange buffer rows [3 + 2m][ldx]
__global__ void lm_export_rows_multi_kernel(LmMultiArgs a, double* __restrict__ rows) {
  const int rho = blockIdx.y;
  const int row = rho_row(a, rho);
  if (row < a.row_begin || row >= a.row_end) return;
  const double* src = a.P + (int64_t)(row - a.row_begin) * a.ld;
  double* dst = rows + (int64_t)rho * a.ldx;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < a.ld; c += (int64_t)gridDim.x * blockDim.x)
    dst[c] = src[c];
  if (blockIdx.x == 0 && threadIdx.x == 0) dst[a.ld] = a.mu[row];
}

cudaError_t lm_launch_export_rows_multi(const LmMultiArgs& a, double* rows, cudaStream_t st) {
  dim3 grid(8, 3 + 2 * a.m);
  lm_export_rows_multi_kernel<<<grid, 256, 0, st>>>(a, rows);
  return cudaGetLastError();
}

template <int M>
static cudaError_t launch_multi_t(const LmMultiArgs& a, cudaStream_t st) {
  constexpr int RB = 32, UNROLL = 8;
  const int rows = a.row_end - a.row_begin;
  if (rows <= 0) return cudaSuccess;
  dim3 grid(a.full_tiles + (a.tail_groups ? 1 : 0), (rows + RB - 1) / RB);
  return launch_pdl(lm_rank_update_multi_kernel<RB, UNROLL, M>, grid, dim3(LM_THREADS), 0, st, a);
}

cudaError_t lm_launch_update_multi(const LmMultiArgs& a0, cudaStream_t st) {
  LmMultiArgs a = a0;
  if (a.m < 2 || a.m > LM_MAX_FUSED) return cudaErrorInvalidValue;
  const int groups = (a.n + 3) / 4;
  a.full_tiles = groups / LM_THREADS;
  a.tail_groups = groups - a.full_tiles * LM_THREADS;
  for (int u = 0; u < a.m; ++u) {
    cudaError_t e = launch_pdl(lm_gather_multi_kernel, dim3((a.n + 255) / 256), dim3(256), 0, st, a, u);
    if (e != cudaSuccess) return e;
  }
  switch (a.m) {
    case 2: return launch_multi_t<2>(a, st);
    case 3: return launch_multi_t<3>(a, st);
    default: return launch_multi_t<4>(a, st);
  }
}

}  // namespace ekf
