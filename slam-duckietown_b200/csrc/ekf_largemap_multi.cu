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
#include "ekf_largemap_p2p.cuh"

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
// row rho of the step's INITIAL covariance: this rank's own P where it owns the row, the exchange buffer otherwise
// (a.rowp / a.mup are filled by the host; the whole matrix on one GPU: all of them point into P / mu)
__device__ __forceinline__ const double* rho_ptr(const LmMultiArgs& a, int rho) { return a.rowp[rho]; }
// exchange part of row rho (0: rows 0-2, 1 + u: rows j_u, j_u + 1) and whether it is read as CELLS of the exchange
// buffer (it arrives from a peer) rather than as plain doubles (own P / replicated rows buffer)
__device__ __forceinline__ int rho_part(int rho) { return rho < 3 ? 0 : 1 + ((rho - 3) >> 1); }
__device__ __forceinline__ bool rho_ll(const LmMultiArgs& a, int rho) {
  return a.x.peers && ((a.wait_mask >> rho_part(rho)) & 1);
}

// a8-a9 for ALL M updates of the step in one launch. Update u works on the covariance as it stands after updates
// 0..u-1 (never materialised). Thread t needs, over the whole step, only column t of the exchanged rows R
// (P0[R, t], for W_u[:, t]) and, for a local row t, that row's entries in the columns R (P0[t, R], for K_u[t, :]),
// plus its OWN earlier K_w[t, :], W_w[:, t] — and a handful of uniform operands per update: H5_u, S_u^-1, y_u,
// K_w[R, :] and W_w[:, R]. Those depend only on the (3 + 2M)^2 block P0[R, R] and mu[R], so every CTA rebuilds
// them itself in shared memory: lanes rho < 3 + 2M of warp 0 act as the threads owning row / column R[rho]
// ("virtual threads", same operation sequence as the real thread t = R[rho]), lane 0 runs innovation, S and S^-1
// between them. No CTA waits for another one, so the m dependent gather launches of the first version (a global
// round trip and a launch between consecutive updates: 36 us of a 196 us step at 4096 landmarks) collapse into one
// launch whose serial part is M innovations. Every value keeps its operation sequence: results unchanged bitwise.
// Block 0 publishes K_u[R, :] (a.KR) and y_u (a.y) for the streaming kernel.
// JOSEPH (whole matrix on one GPU only: K at every column must be local): W_u = H P - Q K^T, AHt_u = P H^T - K (S - Q)
// and every rebuilt entry takes the second rank-2 term (AHt of its row, K of its column) after the first, update by
// update — the operation sequence of the sequential Joseph kernels.
template <int M, bool JOSEPH>
__global__ void __launch_bounds__(256) lm_gather_fused_kernel(LmMultiArgs a) {
  constexpr int NR = 3 + 2 * M;
  __shared__ double sAR[JOSEPH ? M : 1][NR][2];  // AHt_u[R, :]
  __shared__ double ssq[JOSEPH ? M : 1][2][2];   // S_u - Q
  __shared__ double sB[NR][NR + 1];  // P0[R, R]
  __shared__ double smuR[NR];        // mu[R], kept current through the updates
  __shared__ double sKR[M][NR][2];   // K_u[R, :]
  __shared__ double sWR[M][2][NR];   // W_u[:, R]
  __shared__ double sh[M][2][5];
  __shared__ double ssi[M][2][2];
  __shared__ double sy[M][2];
  __shared__ double spk[NR][5];      // P_u[R, idx5(u)] of the update in progress
  const int tid = threadIdx.x;
  pdl_wait();
  const int t = blockIdx.x * blockDim.x + tid;
  const bool live = t < a.n;
  const bool local = live && t >= a.row_begin && t < a.row_end;
  // this thread's column operands touch only the local block: requested first, they fly under the exchange and
  // under warp 0's chain
  double pr0[NR], pc0[NR];
  if (live) {
    const double* prow = a.P + (int64_t)(t - a.row_begin) * a.ld;
#pragma unroll
    for (int rho = 0; rho < NR; ++rho) pc0[rho] = local ? prow[rho_row(a, rho)] : 0.0;  // row t, column R[rho]
  }
  if (a.x.peers && a.push_mask) {  // peer-memory exchange inside this kernel: tagged cells into every peer's buffer
    if (tid < 32) x_wait_acks(a.x, a.x.seq > 2 ? a.x.seq - 2 : 0);
    __syncthreads();
#pragma unroll
    for (int rho = 0; rho < NR; ++rho) {
      if (!((a.push_mask >> rho_part(rho)) & 1)) continue;
      const double v = live ? a.rowp[rho][t] : 0.0;  // this rank's own P
      const double m = (t == 0) ? *a.mup[rho] : 0.0;
      for (int d = 0; d < a.x.world; ++d) {
        if (d == a.x.rank) continue;
        double* dst = x_row(a.x.peers[d], a.x, a.x.seq, rho);
        if (live) ll_store(dst, t, v, a.x.seq);
        if (t == 0) ll_store(dst, a.ld, m, a.x.seq);
      }
    }
  }
  if (tid < NR * NR) {
    const int r = tid / NR, c = tid - r * NR;
    sB[r][c] = x_read(rho_ptr(a, r), rho_row(a, c), rho_ll(a, r), a.x);
  } else if (tid >= 128 && tid < 128 + NR) {
    // the step's initial mean: riding in cell ld of the exchanged row on a shard (local copies of non-local
    // entries are stale there), mu itself otherwise (it is written by the streaming kernel only)
    const int rho = tid - 128;
    smuR[rho] = x_read(a.mup[rho], 0, rho_ll(a, rho), a.x);
  }
  if (live) {  // row R[rho], column t: every request in flight before the first is examined
    LLRaw raw[NR];
#pragma unroll
    for (int rho = 0; rho < NR; ++rho)
      if (rho_ll(a, rho)) raw[rho] = ll_peek(rho_ptr(a, rho), t);
#pragma unroll
    for (int rho = 0; rho < NR; ++rho)
      pr0[rho] = rho_ll(a, rho) ? ll_resolve(raw[rho], rho_ptr(a, rho), t, a.x.seq, a.x.err) : rho_ptr(a, rho)[t];
  }
  __syncthreads();
  if (tid < 32) {
    const int rho = tid < NR ? tid : 0;
    const bool vt = tid < NR;
    double vK[M][2], vW[M][2], vA[M][2];
#pragma unroll
    for (int u = 0; u < M; ++u) {
      const int rho5[5] = {0, 1, 2, 3 + 2 * u, 4 + 2 * u};
      double pr[5], pc[5];
      if (vt) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          pr[k] = sB[rho5[k]][rho];
          pc[k] = sB[rho][rho5[k]];
        }
#pragma unroll
        for (int k = 0; k < 5; ++k)
#pragma unroll
          for (int w = 0; w < u; ++w) {
            pr[k] = rank2e(pr[k], sKR[w][rho5[k]][0], sKR[w][rho5[k]][1], vW[w][0], vW[w][1]);
            pc[k] = rank2e(pc[k], vK[w][0], vK[w][1], sWR[w][0][rho5[k]], sWR[w][1][rho5[k]]);
            if (JOSEPH) {
              pr[k] = rank2e(pr[k], sAR[w][rho5[k]][0], sAR[w][rho5[k]][1], vK[w][0], vK[w][1]);
              pc[k] = rank2e(pc[k], vA[w][0], vA[w][1], sKR[w][rho5[k]][0], sKR[w][rho5[k]][1]);
            }
          }
#pragma unroll
        for (int k = 0; k < 5; ++k) spk[rho][k] = pc[k];
      }
      __syncwarp();
      if (tid == 0) {
        const Innov in = innovation(smuR[0], smuR[1], smuR[2], smuR[rho5[3]], smuR[rho5[4]], a.rng[u], a.brg[u]);
        double hp5[2][5];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            const double p = spk[rho5[k]][c];
            s0 = fma(in.h[0][k], p, s0);
            s1 = fma(in.h[1][k], p, s1);
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
            for (int c = 0; c < 5; ++c) acc = fma(hp5[r][c], in.h[b][c], acc);
            s[r][b] = acc;
          }
        const double qn = a.mm * a.mm;
        if (JOSEPH) {
          ssq[u][0][0] = s[0][0]; ssq[u][0][1] = s[0][1]; ssq[u][1][0] = s[1][0]; ssq[u][1][1] = s[1][1];
        }
        s[0][0] += qn;
        s[1][1] += qn;
        inv2(s, si);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
          for (int k = 0; k < 5; ++k) sh[u][r][k] = in.h[r][k];
          ssi[u][r][0] = si[r][0];
          ssi[u][r][1] = si[r][1];
        }
        sy[u][0] = in.y0;
        sy[u][1] = in.y1;
      }
      __syncwarp();
      if (vt) {
        double hp0 = 0.0, hp1 = 0.0, pht0 = 0.0, pht1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          hp0 = fma(sh[u][0][k], pr[k], hp0);
          hp1 = fma(sh[u][1][k], pr[k], hp1);
          pht0 = fma(pc[k], sh[u][0][k], pht0);
          pht1 = fma(pc[k], sh[u][1][k], pht1);
        }
        vK[u][0] = dot2(pht0, ssi[u][0][0], pht1, ssi[u][1][0]);
        vK[u][1] = dot2(pht0, ssi[u][0][1], pht1, ssi[u][1][1]);
        if (JOSEPH) {
          const double qn = a.mm * a.mm;
          vA[u][0] = joseph_a(pht0, vK[u][0], vK[u][1], ssq[u][0][0], ssq[u][1][0]);
          vA[u][1] = joseph_a(pht1, vK[u][0], vK[u][1], ssq[u][0][1], ssq[u][1][1]);
          hp0 = joseph_w(hp0, qn, vK[u][0]);
          hp1 = joseph_w(hp1, qn, vK[u][1]);
          sAR[u][rho][0] = vA[u][0];
          sAR[u][rho][1] = vA[u][1];
        }
        vW[u][0] = hp0;
        vW[u][1] = hp1;
        sKR[u][rho][0] = vK[u][0];
        sKR[u][rho][1] = vK[u][1];
        sWR[u][0][rho] = hp0;
        sWR[u][1][rho] = hp1;
        smuR[rho] = mu_inc(smuR[rho], vK[u][0], vK[u][1], sy[u][0], sy[u][1]);
        if (blockIdx.x == 0) {
          a.KR[(u * LM_MAX_ROWS + rho) * 2] = vK[u][0];
          a.KR[(u * LM_MAX_ROWS + rho) * 2 + 1] = vK[u][1];
        }
      }
      if (tid == 0 && blockIdx.x == 0) {
        a.y[2 * u] = sy[u][0];
        a.y[2 * u + 1] = sy[u][1];
      }
      __syncwarp();
    }
  }
  __syncthreads();
  if (!live) return;
  double hK[M][2], hW[M][2], hA[M][2];
#pragma unroll
  for (int u = 0; u < M; ++u) {
    const int rho5[5] = {0, 1, 2, 3 + 2 * u, 4 + 2 * u};
    double pr[5], pc[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      pr[k] = pr0[rho5[k]];
      pc[k] = pc0[rho5[k]];
    }
#pragma unroll
    for (int k = 0; k < 5; ++k)
#pragma unroll
      for (int w = 0; w < u; ++w) {
        pr[k] = rank2e(pr[k], sKR[w][rho5[k]][0], sKR[w][rho5[k]][1], hW[w][0], hW[w][1]);
        pc[k] = rank2e(pc[k], hK[w][0], hK[w][1], sWR[w][0][rho5[k]], sWR[w][1][rho5[k]]);
        if (JOSEPH) {
          pr[k] = rank2e(pr[k], sAR[w][rho5[k]][0], sAR[w][rho5[k]][1], hK[w][0], hK[w][1]);
          pc[k] = rank2e(pc[k], hA[w][0], hA[w][1], sKR[w][rho5[k]][0], sKR[w][rho5[k]][1]);
        }
      }
    double hp0 = 0.0, hp1 = 0.0, pht0 = 0.0, pht1 = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      hp0 = fma(sh[u][0][k], pr[k], hp0);
      hp1 = fma(sh[u][1][k], pr[k], hp1);
      pht0 = fma(pc[k], sh[u][0][k], pht0);
      pht1 = fma(pc[k], sh[u][1][k], pht1);
    }
    hK[u][0] = local ? dot2(pht0, ssi[u][0][0], pht1, ssi[u][1][0]) : 0.0;
    hK[u][1] = local ? dot2(pht0, ssi[u][0][1], pht1, ssi[u][1][1]) : 0.0;
    if (JOSEPH) {
      const double qn = a.mm * a.mm;
      hA[u][0] = joseph_a(pht0, hK[u][0], hK[u][1], ssq[u][0][0], ssq[u][1][0]);
      hA[u][1] = joseph_a(pht1, hK[u][0], hK[u][1], ssq[u][0][1], ssq[u][1][1]);
      hp0 = joseph_w(hp0, qn, hK[u][0]);
      hp1 = joseph_w(hp1, qn, hK[u][1]);
      a.A[(2 * u) * a.ldw + t] = hA[u][0];
      a.A[(2 * u + 1) * a.ldw + t] = hA[u][1];
    }
    hW[u][0] = hp0;
    hW[u][1] = hp1;
    if (local) {
      a.K[(2 * u) * a.ldw + t] = hK[u][0];
      a.K[(2 * u + 1) * a.ldw + t] = hK[u][1];
    }
    a.W[(2 * u) * a.ldw + t] = hp0;
    a.W[(2 * u + 1) * a.ldw + t] = hp1;
  }
}

// a10-a11 for all M updates of the step in one pass: P[i, c] <- P0[i, c] - sum_u K_u[i, :] W_u[:, c] in update
// order, mu[i] += K_u[i, :] y_u in update order. Tiling as lm_rank_update_kernel: a CTA owns LM_THREADS column
// groups of 4 doubles x RB rows; W_u of a thread's 4 columns stays in registers (8 M doubles), K_u of the RB rows
// in shared memory, UNROLL rows of P in flight per thread (256-bit streaming loads, in-place 256-bit stores).
__device__ __forceinline__ void xm_prepush4(const LmExchange& x, int row, int c, const double4x& v) {
  for (int d = 0; d < x.world; ++d) {
    if (d == x.rank) continue;
    double* dst = x_row(x.peers[d], x, x.seq + 1, row);
    ll_store2(dst, c, v.a, v.b, x.seq + 1);
    ll_store2(dst, c + 2, v.c, v.d, x.seq + 1);
  }
}

// PRE: rank 0 of a peer-memory exchange — row tile 0 also pushes the updated robot rows for the NEXT step
// second rank-2 term of the one-pass Joseph form on 4 columns: AHt of the row (a0, a1), K of the columns (kc0, kc1)
__device__ __forceinline__ void joseph4(double4x& v, double a0, double a1, const double4x& kc0, const double4x& kc1) {
  v.a = rank2e(v.a, a0, a1, kc0.a, kc1.a);
  v.b = rank2e(v.b, a0, a1, kc0.b, kc1.b);
  v.c = rank2e(v.c, a0, a1, kc0.c, kc1.c);
  v.d = rank2e(v.d, a0, a1, kc0.d, kc1.d);
}

template <int RB, int UNROLL, int M, bool PRE, bool JOSEPH>
__global__ void __launch_bounds__(LM_THREADS) lm_rank_update_multi_kernel(LmMultiArgs a) {
  __shared__ double sk[RB][(JOSEPH ? 4 : 2) * M];  // K_u of the rows, then (Joseph) AHt_u of the rows
  const int tid = threadIdx.x;
  pdl_wait();
  // this launch follows the gather on the stream: the exchange buffer of step x.seq is no longer read here
  if (a.x.peers && blockIdx.x == 0 && blockIdx.y == 0) x_ack(a.x);
  const bool pre = PRE && blockIdx.y == 0;
  if (pre) {
    if (tid < 32) x_wait_acks(a.x, a.x.seq > 1 ? a.x.seq - 1 : 0, 2);
    __syncthreads();
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
    if (JOSEPH) sk[r][2 * M + q] = a.A[(int64_t)q * a.ldw + r0 + r];
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    for (int r = tid; r < nrows; r += LM_THREADS) {
      double v = a.mu[r0 + r];
#pragma unroll
      for (int u = 0; u < M; ++u) v = mu_inc(v, sk[r][2 * u], sk[r][2 * u + 1], a.y[2 * u], a.y[2 * u + 1]);
      a.mu[r0 + r] = v;
      if (pre && r < 3)
        for (int d = 0; d < a.x.world; ++d)
          if (d != a.x.rank) ll_store(x_row(a.x.peers[d], a.x, a.x.seq + 1, r), a.ld, v, a.x.seq + 1);
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
    double4x kc0[JOSEPH ? M : 1], kc1[JOSEPH ? M : 1];  // Joseph: K_u of this thread's 4 columns
    if (JOSEPH) {
#pragma unroll
      for (int u = 0; u < M; ++u) {
        kc0[u] = ldg256(a.K + (int64_t)(2 * u) * a.ldw + c);
        kc1[u] = ldg256(a.K + (int64_t)(2 * u + 1) * a.ldw + c);
      }
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
          if (JOSEPH) joseph4(v[q], sk[r + q][2 * M + 2 * u], sk[r + q][2 * M + 2 * u + 1], kc0[u], kc1[u]);
        }
#pragma unroll
      for (int q = 0; q < UNROLL; ++q) stg256(p + (int64_t)(r + q) * a.ld, v[q]);
      if (pre && r == 0) {
#pragma unroll
        for (int q = 0; q < 3 && q < UNROLL; ++q) xm_prepush4(a.x, q, c, v[q]);
      }
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
        if (JOSEPH) joseph4(v, sk[r][2 * M + 2 * u], sk[r][2 * M + 2 * u + 1], kc0[u], kc1[u]);
      }
      stg256(p + (int64_t)r * a.ld, v);
      if (pre && r < 3 && nrows < UNROLL) xm_prepush4(a.x, r, c, v);
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
        if (JOSEPH)
          joseph4(v, sk[r][2 * M + 2 * u], sk[r][2 * M + 2 * u + 1], ldg256(a.K + (int64_t)(2 * u) * a.ldw + c),
                  ldg256(a.K + (int64_t)(2 * u + 1) * a.ldw + c));
      }
      stg256(Pblk + (int64_t)r * a.ld + c, v);
      if (pre && r < 3) xm_prepush4(a.x, r, c, v);
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

template <int M, bool JOSEPH, int RB>
static cudaError_t launch_multi_rb(const LmMultiArgs& a, cudaStream_t st) {
  constexpr int UNROLL = JOSEPH ? 4 : 8;
  const int rows = a.row_end - a.row_begin;
  if (rows <= 0) return cudaSuccess;
  dim3 grid(a.full_tiles + (a.tail_groups ? 1 : 0), (rows + RB - 1) / RB);
  if constexpr (!JOSEPH) {
    if (a.prepush && a.x.peers)
      return launch_pdl(lm_rank_update_multi_kernel<RB, UNROLL, M, true, false>, grid, dim3(LM_THREADS), 0, st, a);
  }
  return launch_pdl(lm_rank_update_multi_kernel<RB, UNROLL, M, false, JOSEPH>, grid, dim3(LM_THREADS), 0, st, a);
}

// variant 1: 16-row tiles for row blocks that fit L2 (see ekf_lm_bind_state), else 32-row tiles
template <int M, bool JOSEPH>
static cudaError_t launch_multi_t(const LmMultiArgs& a, cudaStream_t st) {
  if (a.variant == 1) return launch_multi_rb<M, JOSEPH, 16>(a, st);
  return launch_multi_rb<M, JOSEPH, 32>(a, st);
}

template <int M, bool JOSEPH>
static cudaError_t launch_fused_t(const LmMultiArgs& a, cudaStream_t st) {
  const dim3 ggrid((a.n + 255) / 256), gblock(256);
  cudaError_t e = launch_pdl(lm_gather_fused_kernel<M, JOSEPH>, ggrid, gblock, 0, st, a);
  if (e != cudaSuccess) return e;
  return launch_multi_t<M, JOSEPH>(a, st);
}

cudaError_t lm_launch_update_multi(const LmMultiArgs& a0, cudaStream_t st) {
  LmMultiArgs a = a0;
  if (a.m < 2 || a.m > LM_MAX_FUSED) return cudaErrorInvalidValue;
  if (a.joseph && (a.xrows || a.row_begin != 0 || a.row_end != a.n)) return cudaErrorInvalidValue;
  const int groups = (a.n + 3) / 4;
  a.full_tiles = groups / LM_THREADS;
  a.tail_groups = groups - a.full_tiles * LM_THREADS;
  switch (a.m) {
    case 2: return a.joseph ? launch_fused_t<2, true>(a, st) : launch_fused_t<2, false>(a, st);
    case 3: return a.joseph ? launch_fused_t<3, true>(a, st) : launch_fused_t<3, false>(a, st);
    default: return a.joseph ? launch_fused_t<4, true>(a, st) : launch_fused_t<4, false>(a, st);
  }
}

}  // namespace ekf
