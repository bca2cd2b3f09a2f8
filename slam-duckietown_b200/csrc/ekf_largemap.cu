// ekf_largemap.cu — large-map single-filter EKF-SLAM kernels (sm_100a, fp64, HBM-streaming).
//
// One filter whose dense covariance P (n x n, n = 3 + 2N, row-major, leading dimension ld = n rounded up
// to 16 doubles so every row starts on a 128-byte line) lives in HBM. A GPU holds rows [row_begin, row_end)
// of P (the whole matrix on one GPU, a row block when sharded).
//
// Reference semantics (src/replay_no_ros.py): augment :341-360, predict :368-430, update :436-480.
// The reference multiplies dense n x n matrices (O(n^3)); the same algebra is a gather of 5 rows and
// 5 columns plus a rank-2 (plain) or rank-4 (Joseph, one pass) in-place update of P:
//
//   K  = P[:, idx5] H5^T S^-1      from COLUMNS of P        idx5 = {0, 1, 2, j, j+1}, j = 3 + 2*slot
//   HP = H5 P[idx5, :]             from ROWS of P
//   plain : P -= K HP                                   Joseph: P -= K W + AHt K^T,
//                                                        W = HP - Q K^T, AHt = P H^T - K (S - Q)
// Rows and columns are both read for real: substituting (P H^T)^T for H P on full, slightly asymmetric
// storage is numerically unstable.
//
// Roofline: HBM. Algorithmic bytes per update = 16 n^2 (read + write P) + 96 n (gathers + mu); the
// streaming kernel `lm_rank_update_kernel` moves the 16 n^2 with 256-bit loads/stores.
#include "ekf_common.cuh"
#include "ekf_largemap.h"

#ifdef EKF_LM_TIMELINE
// Diagnostic build only (make timeline -> libekf_b200_tl.so, tools/lm_p2p_timeline.py): globaltimer stamps of the
// phases of the sharded single update, indexed by the update's sequence number. Never compiled into libekf_b200.so.
namespace ekf {
__device__ unsigned long long g_lm_tl[256][10];
__device__ __forceinline__ unsigned long long lm_gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
}  // namespace ekf
#define EKF_TL(seq, k) (ekf::g_lm_tl[(seq) & 255][k] = ekf::lm_gtime())
#define EKF_TLMAX(seq, k) atomicMax(&ekf::g_lm_tl[(seq) & 255][k], ekf::lm_gtime())
#endif
#include "ekf_largemap_p2p.cuh"

namespace ekf {

// -------------------------------------------------------------------------------------------------------
// a5 augment (replay_no_ros.py:355-360): rows/columns of a not-yet-seen landmark are zero by invariant.
// -------------------------------------------------------------------------------------------------------
__global__ void lm_augment_kernel(double* __restrict__ mu, double* __restrict__ P, int64_t ld, int row_begin,
                                  int row_end, int j, double gx, double gy) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  mu[j] = gx;
  mu[j + 1] = gy;
  if (j >= row_begin && j < row_end) P[(int64_t)(j - row_begin) * ld + j] = kNewLandmarkVar;
  if (j + 1 >= row_begin && j + 1 < row_end) P[(int64_t)(j + 1 - row_begin) * ld + j + 1] = kNewLandmarkVar;
}

// -------------------------------------------------------------------------------------------------------
// a6-a7 predict (:368-430) in ONE launch. Mean: every CTA's thread 0 evaluates the motion model from the
// pre-predict heading mu[2] (a few hundred cycles, redundant per CTA) and shares g0 = G[0,2], g1 = G[1,2]
// through shared memory. Covariance: P <- G_F P G_F^T + F^T R F touches rows 0-1, columns 0-1 and three
// diagonal entries; thread t >= 3 does column t of rows 0-1 and row t of columns 0-1, thread 0 the 3x3 robot
// block in the same order (rows first, then columns with the updated rows, then the noise).
// mu[0..2] is written by CTA 0 only AFTER every CTA has read the old heading: each CTA bumps an arrival counter
// once it holds g0/g1, CTA 0's thread 0 waits for gridDim.x arrivals (nobody waits on CTA 0, so this cannot
// deadlock whatever the residency) and re-arms the counter. Round 1 used two plain launches (mean, covariance):
// 22 us per predict at 4096 landmarks, of which 8 us were kernels.
// -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lm_predict_kernel(double* __restrict__ mu, double* __restrict__ P, int64_t ld,
                                                         int n, int row_begin, int row_end,
                                                         unsigned long long* __restrict__ arrive, double ang,
                                                         double lin, double mv, int flags) {
  __shared__ double sg[5];
  pdl_wait();
  if (threadIdx.x == 0) {
    const double th = *reinterpret_cast<volatile double*>(mu + 2);
    const Motion m = motion_step(th, ang, lin, flags);
    sg[0] = m.g0; sg[1] = m.g1; sg[2] = m.dx; sg[3] = m.dy; sg[4] = m.theta;
    __threadfence();
    atomicAdd(arrive, 1ULL);
  }
  __syncthreads();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const double g0 = sg[0], g1 = sg[1];
  const bool own_robot = (row_begin == 0);
  if (t >= 3 && t < n) {
    if (own_robot) {  // rows 0-1, column t
      const double p2 = P[2 * ld + t];
      P[0 * ld + t] += g0 * p2;
      P[1 * ld + t] += g1 * p2;
    }
    if (t >= row_begin && t < row_end) {  // row t, columns 0-1
      double* row = P + (int64_t)(t - row_begin) * ld;
      const double p2 = row[2];
      row[0] += g0 * p2;
      row[1] += g1 * p2;
    }
  } else if (t == 0) {
    if (own_robot) {
      double b[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) b[i][c] = P[i * ld + c];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        b[0][c] += g0 * b[2][c];
        b[1][c] += g1 * b[2][c];
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        b[i][0] += g0 * b[i][2];
        b[i][1] += g1 * b[i][2];
      }
      b[0][0] += mv * mv;
      b[1][1] += mv * mv;
      b[2][2] += (mv / 2.0) * (mv / 2.0);
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) P[i * ld + c] = b[i][c];
    }
    // the mean, last: every CTA has taken its copy of the old heading
    const long long t0 = clock64();
    while (*reinterpret_cast<volatile unsigned long long*>(arrive) < (unsigned long long)gridDim.x) {
      if (clock64() - t0 > 4000000000LL) break;  // cannot happen (no CTA waits on another); never hang the GPU
      __nanosleep(20);
    }
    *arrive = 0ULL;
    mu[0] += sg[2];
    mu[1] += sg[3];
    mu[2] = sg[4];
  }
}

// -------------------------------------------------------------------------------------------------------
// a8-a9 gather (:443-473): every CTA recomputes the innovation, H5, S and S^-1 (25 + 5 loads), then
// thread t produces HP[:, t] from the 5 rows (coalesced) and, for a local row t, PHt[t, :] from the
// 5 columns (one or two 32-byte sectors per row) and K[t, :]. Nothing in P or mu is written here, so the
// streaming kernel that follows can update P in place.
// -------------------------------------------------------------------------------------------------------
// bid / nblocks: this CTA's index among the CTAs running the gather and their number (the kernel's own blockIdx.x /
// gridDim.x when launched alone; the leading CTAs of the single-launch update otherwise)
__device__ __forceinline__ void lm_gather_body(const LmGatherArgs& a, int bid, int nblocks) {
  __shared__ double sh[2][5];
  __shared__ double ssi[2][2];
  __shared__ double ssq[2][2];  // S - Q
  __shared__ double sy[2];
  __shared__ double sb5[5][5], sm5[5];  // P[idx5, idx5], mu[idx5]
  const int j = a.j;
  const int idx5[5] = {0, 1, 2, j, j + 1};
  const int t = bid * blockDim.x + threadIdx.x;
  const bool live = t < a.n;
  const bool local = live && (t >= a.row_begin && t < a.row_end);
  // this thread's column gathers touch only the local block: issued first, they fly under the exchange below and
  // under thread 0's scalar chain (innovation, S^-1)
  double pr[5], pc[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) pc[k] = local ? a.P[(int64_t)(t - a.row_begin) * a.ld + idx5[k]] : 0.0;
  const bool tl = a.x.peers && bid == 0 && threadIdx.x == 0;
  if (tl) EKF_TL(a.x.seq, 0);
  // which of the two row parts (0: rows 0-2, 1: rows j, j+1) are read as cells of the exchange buffer
  const bool ll0 = a.x.peers && (a.wait_mask & 1), ll1 = a.x.peers && (a.wait_mask & 2);
  if (a.x.peers && (a.push_mask || a.share_k)) {
    // peer-memory exchange inside this kernel (LmExchange, ekf_largemap.h): the owner stores its rows as tagged
    // cells into every peer's buffer; nothing to publish, the consumers poll the cells themselves
    if (threadIdx.x < 32) x_wait_acks(a.x, a.x.seq > 2 ? a.x.seq - 2 : 0);
    __syncthreads();
    if (tl) EKF_TL(a.x.seq, 1);
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      if (!((a.push_mask >> (k < 3 ? 0 : 1)) & 1)) continue;
      const double v = live ? a.row[k][t] : 0.0;                     // a.row[k] points into this rank's own P
      const double m = (t == 0) ? *a.mu5[k] : 0.0;
      for (int d = 0; d < a.x.world; ++d) {
        if (d == a.x.rank) continue;
        double* dst = x_row(a.x.peers[d], a.x, a.x.seq, k);
        if (live) ll_store(dst, t, v, a.x.seq);
        if (t == 0) ll_store(dst, a.ld, m, a.x.seq);
      }
    }
    if (tl) EKF_TL(a.x.seq, 2);
  }
  // the 25 entries P[idx5, idx5] and the 5 mu entries every CTA's scalar chain needs: one per lane of warp 0
  if (threadIdx.x < 30) {
    const int k = threadIdx.x < 25 ? threadIdx.x / 5 : threadIdx.x - 25;
    const bool ll = k < 3 ? ll0 : ll1;
    if (threadIdx.x < 25) sb5[k][threadIdx.x - 5 * k] = x_read(a.row[k], idx5[threadIdx.x - 5 * k], ll, a.x);
    else sm5[k] = x_read(a.mu5[k], 0, ll, a.x);
  }
  {  // this thread's column of the five rows: all five requests in flight before the first is examined
    LLRaw raw[5];
#pragma unroll
    for (int k = 0; k < 5; ++k)
      if (live && (k < 3 ? ll0 : ll1)) raw[k] = ll_peek(a.row[k], t);
#pragma unroll
    for (int k = 0; k < 5; ++k)
      pr[k] = !live ? 0.0 : (k < 3 ? ll0 : ll1) ? ll_resolve(raw[k], a.row[k], t, a.x.seq, a.x.err) : a.row[k][t];
  }
  __syncthreads();
  if (tl) EKF_TL(a.x.seq, 3);
  if (threadIdx.x == 0) {
    const Innov in = innovation(sm5[0], sm5[1], sm5[2], sm5[3], sm5[4], a.rng, a.brg);
    double hp5[2][5];
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double p = sb5[k][c];
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
    ssq[0][0] = s[0][0]; ssq[0][1] = s[0][1]; ssq[1][0] = s[1][0]; ssq[1][1] = s[1][1];
    s[0][0] += qn;
    s[1][1] += qn;
    inv2(s, si);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
#pragma unroll
      for (int k = 0; k < 5; ++k) sh[r][k] = in.h[r][k];
      ssi[r][0] = si[r][0];
      ssi[r][1] = si[r][1];
    }
    sy[0] = in.y0;
    sy[1] = in.y1;
    if (bid == 0) {
      a.scal[kScalY0] = in.y0;
      a.scal[kScalY1] = in.y1;
    }
    if (tl) EKF_TL(a.x.seq, 4);
  }
  __syncthreads();
  if (!live) return;
  if (a.shard && t < 3 && !(t >= a.row_begin && t < a.row_end)) {
    // row-block shard that does not own the robot rows: keep the replicated pose entries of mu current.
    // K[t, :] of rows 0-2 is rebuilt from the exchanged rows (mu5 points into the exchange buffer, so
    // writing mu here does not race with the innovation above).
    double ph0 = 0.0, ph1 = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const double p = sb5[t][k];  // == a.row[t][idx5[k]]
      ph0 = fma(p, sh[0][k], ph0);
      ph1 = fma(p, sh[1][k], ph1);
    }
    const double kk0 = dot2(ph0, ssi[0][0], ph1, ssi[1][0]);
    const double kk1 = dot2(ph0, ssi[0][1], ph1, ssi[1][1]);
    a.mu_out[t] = mu_inc(a.mu_out[t], kk0, kk1, sy[0], sy[1]);
  }
  double hp0 = 0.0, hp1 = 0.0;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    hp0 = fma(sh[0][k], pr[k], hp0);
    hp1 = fma(sh[1][k], pr[k], hp1);
  }
  double k0 = 0.0, k1 = 0.0, pht0 = 0.0, pht1 = 0.0;
  if (local) {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      pht0 = fma(pc[k], sh[0][k], pht0);
      pht1 = fma(pc[k], sh[1][k], pht1);
    }
    k0 = dot2(pht0, ssi[0][0], pht1, ssi[1][0]);
    k1 = dot2(pht0, ssi[0][1], pht1, ssi[1][1]);
    a.K0[t] = k0;
    a.K1[t] = k1;
  }
  if (a.flags & kFlagJoseph) {
    // W needs K at EVERY column t. Whole matrix on one GPU: K[t] is local. Row-block shard (peer-memory exchange):
    // every rank stores K of its rows into every peer's buffer and takes the others' from its own (tagged cells).
    if (a.share_k) {
      if (local) {
        for (int d = 0; d < a.x.world; ++d) {
          if (d == a.x.rank) continue;
          ll_store(x_row(a.x.peers[d], a.x, a.x.seq, P2P_KROW), t, k0, a.x.seq);
          ll_store(x_row(a.x.peers[d], a.x, a.x.seq, P2P_KROW + 1), t, k1, a.x.seq);
        }
      } else {
        double* mine = a.x.peers[a.x.rank];
        const LLRaw r0 = ll_peek(x_row(mine, a.x, a.x.seq, P2P_KROW), t);
        const LLRaw r1 = ll_peek(x_row(mine, a.x, a.x.seq, P2P_KROW + 1), t);
        k0 = ll_resolve(r0, x_row(mine, a.x, a.x.seq, P2P_KROW), t, a.x.seq, a.x.err);
        k1 = ll_resolve(r1, x_row(mine, a.x, a.x.seq, P2P_KROW + 1), t, a.x.seq, a.x.err);
        a.K0[t] = k0;  // the streaming kernel reads K at its tile's columns
        a.K1[t] = k1;
      }
    }
    const double qn = a.mm * a.mm;
    a.W0[t] = joseph_w(hp0, qn, k0);
    a.W1[t] = joseph_w(hp1, qn, k1);
    a.A0[t] = joseph_a(pht0, k0, k1, ssq[0][0], ssq[1][0]);
    a.A1[t] = joseph_a(pht1, k0, k1, ssq[0][1], ssq[1][1]);
  } else {
    a.W0[t] = hp0;
    a.W1[t] = hp1;
  }
  if (a.x.peers && threadIdx.x == 0) EKF_TLMAX(a.x.seq, 5);
}

__global__ void __launch_bounds__(256) lm_gather_kernel(LmGatherArgs a) {
  pdl_wait();
  lm_gather_body(a, blockIdx.x, gridDim.x);
}

// -------------------------------------------------------------------------------------------------------
// a10-a11 streaming update (:476-480): P[i, c] -= K[i,0] W[0,c] + K[i,1] W[1,c]  (+ Joseph terms), in place.
//
// Tiling: a CTA of LM_THREADS threads owns LM_THREADS column groups of 4 doubles (one 256-bit access per
// thread per row) and RB rows; the two W values of a thread's 4 columns stay in registers for all RB rows,
// K of the RB rows is staged in shared memory, and UNROLL rows of P are in flight per thread
// (UNROLL x 32 B x threads x resident CTAs per SM of loads outstanding). Columns beyond the last full tile
// are done by an extra column of CTAs with a (row, group) mapping. CTAs of column tile 0 also apply
// mu += K y for their rows.
// -------------------------------------------------------------------------------------------------------
// push of rows 0-2 (just updated, still in registers) into every peer's buffer of the NEXT update: tagged cells,
// two per 256-bit store
__device__ __forceinline__ void x_prepush4(const LmExchange& x, int row, int c, const double4x& v) {
  for (int d = 0; d < x.world; ++d) {
    if (d == x.rank) continue;
    double* dst = x_row(x.peers[d], x, x.seq + 1, row);
    ll_store2(dst, c, v.a, v.b, x.seq + 1);
    ll_store2(dst, c + 2, v.c, v.d, x.seq + 1);
  }
}

// PRE: rank 0 of a peer-memory exchange — row tile 0 also pushes the updated robot rows (see LmExchange)
// (bx, by) of nbx column tiles: this CTA's tile. ready != null (single-launch update): the CTA first requests its first
// UNROLL rows of P — they do not depend on the gather — and only then waits until the gather CTAs of the same launch
// have counted in (`*ready >= target`); nothing of P is stored before that, so the gather still sees the old P.
template <int RB, int UNROLL, bool JOSEPH, bool PRE>
__device__ __forceinline__ void lm_rank_update_body(const LmUpdateArgs& a, int bx, int by, int nbx,
                                                    const unsigned long long* ready, unsigned long long target) {
  __shared__ double sk[RB][4];
  const int tid = threadIdx.x;
  const bool pre = PRE && by == 0;
  const int r0 = a.row_begin + by * RB;          // first global row of this CTA
  const int nrows = min(RB, a.row_end - r0);
  double* __restrict__ Pblk = a.P + (int64_t)(r0 - a.row_begin) * a.ld;
  double4x v0[UNROLL];                            // rows 0..UNROLL-1 of a full tile, requested ahead of the wait
  const bool ahead = ready != nullptr && bx < a.full_tiles && nrows >= UNROLL;
  if (ahead) {
    const double* p = Pblk + (bx * LM_THREADS + tid) * 4;
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v0[u] = ldg256_stream(p + (int64_t)u * a.ld);
  }
  if (ready) {
    if (tid == 0) {
      const long long t0 = clock64();
      while (*reinterpret_cast<const volatile unsigned long long*>(ready) < target) {
        if (clock64() - t0 > kSpinBudget) break;  // cannot happen (the gather CTAs wait on nothing local)
        __nanosleep(20);
      }
      __threadfence();
    }
    __syncthreads();
  }
  // the gather of this update is complete on this rank: its exchange buffer is no longer read here
  if (a.x.peers && bx == 0 && by == 0) {
    if (tid == 0) EKF_TL(a.x.seq, 6);
    x_ack(a.x);
  }
  if (pre) {  // the buffer of update seq + 1 was last read by the gathers of update seq - 1
    if (tid < 32) x_wait_acks(a.x, a.x.seq > 1 ? a.x.seq - 1 : 0, 2);
    __syncthreads();
  }
  for (int r = tid; r < nrows; r += LM_THREADS) {
    sk[r][0] = a.K0[r0 + r];
    sk[r][1] = a.K1[r0 + r];
    if (JOSEPH) {
      sk[r][2] = a.A0[r0 + r];
      sk[r][3] = a.A1[r0 + r];
    }
  }
  __syncthreads();
  if (bx == 0) {  // a10: mu += K y for this CTA's rows
    const double y0 = a.scal[kScalY0], y1 = a.scal[kScalY1];
    for (int r = tid; r < nrows; r += LM_THREADS) {
      const double m = mu_inc(a.mu[r0 + r], sk[r][0], sk[r][1], y0, y1);
      a.mu[r0 + r] = m;
      if (pre && r < 3)  // r0 == 0 here: the pose entries ride at column ld of the pushed rows
        for (int d = 0; d < a.x.world; ++d)
          if (d != a.x.rank) ll_store(x_row(a.x.peers[d], a.x, a.x.seq + 1, r), a.ld, m, a.x.seq + 1);
    }
  }
  if (bx < a.full_tiles) {
    const int c = (bx * LM_THREADS + tid) * 4;
    const double4x w0 = ldg256(a.W0 + c), w1 = ldg256(a.W1 + c);
    double4x kc0, kc1;
    if (JOSEPH) {
      kc0 = ldg256(a.K0 + c);
      kc1 = ldg256(a.K1 + c);
    }
    double* p = Pblk + c;
    int r = 0;
    for (; r + UNROLL <= nrows; r += UNROLL) {
      double4x v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) v[u] = (ahead && r == 0) ? v0[u] : ldg256_stream(p + (int64_t)(r + u) * a.ld);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const double k0 = -sk[r + u][0], k1 = -sk[r + u][1];
        v[u].a = fma(k1, w1.a, fma(k0, w0.a, v[u].a));
        v[u].b = fma(k1, w1.b, fma(k0, w0.b, v[u].b));
        v[u].c = fma(k1, w1.c, fma(k0, w0.c, v[u].c));
        v[u].d = fma(k1, w1.d, fma(k0, w0.d, v[u].d));
        if (JOSEPH) {
          const double a0 = -sk[r + u][2], a1 = -sk[r + u][3];
          v[u].a = fma(a1, kc1.a, fma(a0, kc0.a, v[u].a));
          v[u].b = fma(a1, kc1.b, fma(a0, kc0.b, v[u].b));
          v[u].c = fma(a1, kc1.c, fma(a0, kc0.c, v[u].c));
          v[u].d = fma(a1, kc1.d, fma(a0, kc0.d, v[u].d));
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) stg256(p + (int64_t)(r + u) * a.ld, v[u]);
      if (pre && r == 0) {
#pragma unroll
        for (int u = 0; u < 3 && u < UNROLL; ++u) x_prepush4(a.x, u, c, v[u]);
      }
    }
    for (; r < nrows; ++r) {
      double4x v = ldg256_stream(p + (int64_t)r * a.ld);
      const double k0 = -sk[r][0], k1 = -sk[r][1];
      v.a = fma(k1, w1.a, fma(k0, w0.a, v.a));
      v.b = fma(k1, w1.b, fma(k0, w0.b, v.b));
      v.c = fma(k1, w1.c, fma(k0, w0.c, v.c));
      v.d = fma(k1, w1.d, fma(k0, w0.d, v.d));
      if (JOSEPH) {
        const double a0 = -sk[r][2], a1 = -sk[r][3];
        v.a = fma(a1, kc1.a, fma(a0, kc0.a, v.a));
        v.b = fma(a1, kc1.b, fma(a0, kc0.b, v.b));
        v.c = fma(a1, kc1.c, fma(a0, kc0.c, v.c));
        v.d = fma(a1, kc1.d, fma(a0, kc0.d, v.d));
      }
      stg256(p + (int64_t)r * a.ld, v);
      if (pre && r < 3 && nrows < UNROLL) x_prepush4(a.x, r, c, v);
    }
  } else {
    // column tail: a.tail_groups (< LM_THREADS) groups per row, (row, group) mapping
    const int tg = a.tail_groups;
    const int cbase = a.full_tiles * LM_THREADS * 4;
    for (int e = tid; e < nrows * tg; e += LM_THREADS) {
      const int r = e / tg, c = cbase + (e - r * tg) * 4;
      const double4x w0 = ldg256(a.W0 + c), w1 = ldg256(a.W1 + c);
      double4x v = ldg256_stream(Pblk + (int64_t)r * a.ld + c);
      const double k0 = -sk[r][0], k1 = -sk[r][1];
      v.a = fma(k1, w1.a, fma(k0, w0.a, v.a));
      v.b = fma(k1, w1.b, fma(k0, w0.b, v.b));
      v.c = fma(k1, w1.c, fma(k0, w0.c, v.c));
      v.d = fma(k1, w1.d, fma(k0, w0.d, v.d));
      if (JOSEPH) {
        const double4x kc0 = ldg256(a.K0 + c), kc1 = ldg256(a.K1 + c);
        const double a0 = -sk[r][2], a1 = -sk[r][3];
        v.a = fma(a1, kc1.a, fma(a0, kc0.a, v.a));
        v.b = fma(a1, kc1.b, fma(a0, kc0.b, v.b));
        v.c = fma(a1, kc1.c, fma(a0, kc0.c, v.c));
        v.d = fma(a1, kc1.d, fma(a0, kc0.d, v.d));
      }
      stg256(Pblk + (int64_t)r * a.ld + c, v);
      if (pre && r < 3) x_prepush4(a.x, r, c, v);
    }
  }
  if (pre && bx == 0 && tid == 0) EKF_TL(a.x.seq, 8);
  if (a.x.peers && tid == 0) EKF_TLMAX(a.x.seq, 7);
}


template <int RB, int UNROLL, bool JOSEPH, bool PRE>
__global__ void __launch_bounds__(LM_THREADS) lm_rank_update_kernel(LmUpdateArgs a) {
  pdl_wait();
  lm_rank_update_body<RB, UNROLL, JOSEPH, PRE>(a, blockIdx.x, blockIdx.y, gridDim.x, nullptr, 0ULL);
}

// ONE launch per update (round 2): the first `gather_ctas` CTAs — dispatched first — run the gather and count in on
// `ready`; the others are the streaming tiles, which pre-load their first rows and then wait for that count. The
// gather's dependent chain (loads -> innovation -> S^-1 -> dot products -> stores, 8-9 us at 4096 landmarks) and one
// kernel boundary disappear behind the first wave of P loads. `ready` only ever grows: `target` = gather_ctas x the
// number of single-launch updates issued on this handle, so it needs no reset.
struct LmFusedArgs {
  LmGatherArgs g;
  LmUpdateArgs u;
  unsigned long long* ready;
  unsigned long long target;
  int gather_ctas, nbx;
};
template <int RB, int UNROLL, bool PRE>
__global__ void __launch_bounds__(LM_THREADS) lm_update_single_launch_kernel(LmFusedArgs f) {
  pdl_wait();
  const int bid = blockIdx.x;
  if (bid < f.gather_ctas) {
    lm_gather_body(f.g, bid, f.gather_ctas);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(f.ready, 1ULL);
  } else {
    const int idx = bid - f.gather_ctas;
    lm_rank_update_body<RB, UNROLL, false, PRE>(f.u, idx % f.nbx, idx / f.nbx, f.nbx, f.ready, f.target);
  }
}

// shard export: copy the locally owned ones of rows {0,1,2,j,j+1} (and their mu entries, stashed at column
// ld of each exported row) into the replicated exchange buffer.
__global__ void lm_export_rows_kernel(const double* __restrict__ P, const double* __restrict__ mu, int64_t ld,
                                      int row_begin, int row_end, int j, double* __restrict__ rows5,
                                      int64_t ldx) {
  const int k = blockIdx.y;
  const int row = (k < 3) ? k : j + (k - 3);
  if (row < row_begin || row >= row_end) return;
  const double* src = P + (int64_t)(row - row_begin) * ld;
  double* dst = rows5 + (int64_t)k * ldx;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < ld; c += (int64_t)gridDim.x * blockDim.x)
    dst[c] = src[c];
  if (blockIdx.x == 0 && threadIdx.x == 0) dst[ld] = mu[row];
}

// after a sharded update every rank refreshes its replicated copy of the 5 mu entries from K rows it can
// rebuild out of the exchanged rows: K[r, :] = (rows5[k][idx5] . H5^T) S^-1 — done inside the gather kernel
// of the NEXT update through the stash, so nothing is needed here.

// -------------------------------------------------------------------------------------------------------
// host launchers
// -------------------------------------------------------------------------------------------------------
cudaError_t lm_launch_augment(double* mu, double* P, int64_t ld, int row_begin, int row_end, int j, double gx,
                              double gy, cudaStream_t st) {
  lm_augment_kernel<<<1, 32, 0, st>>>(mu, P, ld, row_begin, row_end, j, gx, gy);
  return cudaGetLastError();
}

cudaError_t lm_launch_predict(double* mu, double* P, int64_t ld, int n, int row_begin, int row_end,
                              double* scal, double ang, double lin, double mv, int flags, cudaStream_t st) {
  const int threads = 256;
  unsigned long long* arrive = reinterpret_cast<unsigned long long*>(scal + kScalArrive);
  return launch_pdl(lm_predict_kernel, dim3((n + threads - 1) / threads), dim3(threads), 0, st, mu, P, ld, n,
                    row_begin, row_end, arrive, ang, lin, mv, flags);
}

cudaError_t lm_launch_gather(const LmGatherArgs& a, cudaStream_t st) {
  // peer-memory exchange: smaller CTAs put the owner's stores to its peers on twice as many SMs (the timeline showed
  // ~17 GB/s of peer stores per SM, so 33 CTAs could not fill the NVLink ports)
  const int threads = a.x.peers ? 128 : 256;
  return launch_pdl(lm_gather_kernel, dim3((a.n + threads - 1) / threads), dim3(threads), 0, st, a);
}

template <int RB, int UNROLL>
static cudaError_t launch_update_t(const LmUpdateArgs& a0, bool joseph, cudaStream_t st) {
  LmUpdateArgs a = a0;
  const int groups = (a.n + 3) / 4;  // groups holding at least one real column
  a.full_tiles = groups / LM_THREADS;
  a.tail_groups = groups - a.full_tiles * LM_THREADS;
  const int rows = a.row_end - a.row_begin;
  if (rows <= 0) return cudaSuccess;
  dim3 grid(a.full_tiles + (a.tail_groups ? 1 : 0), (rows + RB - 1) / RB);
  if (a.prepush && a.x.peers && !joseph)
    return launch_pdl(lm_rank_update_kernel<RB, UNROLL, false, true>, grid, dim3(LM_THREADS), 0, st, a);
  if (joseph) return launch_pdl(lm_rank_update_kernel<RB, UNROLL, true, false>, grid, dim3(LM_THREADS), 0, st, a);
  return launch_pdl(lm_rank_update_kernel<RB, UNROLL, false, false>, grid, dim3(LM_THREADS), 0, st, a);
}

cudaError_t lm_launch_update(const LmUpdateArgs& a, bool joseph, int variant, cudaStream_t st) {
  switch (variant) {
    case 1: return launch_update_t<16, 8>(a, joseph, st);
    case 2: return launch_update_t<64, 8>(a, joseph, st);
    case 3: return launch_update_t<32, 4>(a, joseph, st);
    case 4: return launch_update_t<128, 8>(a, joseph, st);
    case 5: return launch_update_t<32, 16>(a, joseph, st);
    default: return launch_update_t<32, 8>(a, joseph, st);
  }
}

template <int RB, int UNROLL>
static cudaError_t launch_update_single_t(const LmGatherArgs& g, const LmUpdateArgs& u0, unsigned long long* ready,
                                          unsigned long long* calls, cudaStream_t st) {
  LmFusedArgs f{};
  f.g = g;
  f.u = u0;
  const int groups = (u0.n + 3) / 4;
  f.u.full_tiles = groups / LM_THREADS;
  f.u.tail_groups = groups - f.u.full_tiles * LM_THREADS;
  const int rows = u0.row_end - u0.row_begin;
  f.nbx = f.u.full_tiles + (f.u.tail_groups ? 1 : 0);
  const int nby = rows > 0 ? (rows + RB - 1) / RB : 0;
  f.gather_ctas = (g.n + LM_THREADS - 1) / LM_THREADS;
  f.ready = ready;
  f.target = (unsigned long long)f.gather_ctas * ++(*calls);
  dim3 grid(f.gather_ctas + f.nbx * nby);
  if (u0.prepush && u0.x.peers)
    return launch_pdl(lm_update_single_launch_kernel<RB, UNROLL, true>, grid, dim3(LM_THREADS), 0, st, f);
  return launch_pdl(lm_update_single_launch_kernel<RB, UNROLL, false>, grid, dim3(LM_THREADS), 0, st, f);
}

// variant 0: 32-row tiles (blocks that stream from HBM); variant 1: 16-row tiles (row blocks that fit L2)
cudaError_t lm_launch_update_single(const LmGatherArgs& g, const LmUpdateArgs& u, int variant, unsigned long long* ready,
                                    unsigned long long* calls, cudaStream_t st) {
  if (variant == 1) return launch_update_single_t<16, 8>(g, u, ready, calls, st);
  return launch_update_single_t<32, 8>(g, u, ready, calls, st);
}

#ifdef EKF_LM_TIMELINE
}  // namespace ekf
extern "C" int ekf_lm_debug_timeline(unsigned long long* out_host, int clear) {
  if (out_host && cudaMemcpyFromSymbol(out_host, ekf::g_lm_tl, sizeof(ekf::g_lm_tl)) != cudaSuccess) return -1;
  if (clear) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, ekf::g_lm_tl) != cudaSuccess) return -1;
    if (cudaMemset(p, 0, sizeof(ekf::g_lm_tl)) != cudaSuccess) return -1;
  }
  return 0;
}
namespace ekf {
#endif

cudaError_t lm_launch_export_rows(const double* P, const double* mu, int64_t ld, int row_begin, int row_end,
                                  int j, double* rows5, int64_t ldx, cudaStream_t st) {
  dim3 grid(8, 5);
  lm_export_rows_kernel<<<grid, 256, 0, st>>>(P, mu, ld, row_begin, row_end, j, rows5, ldx);
  return cudaGetLastError();
}

}  // namespace ekf
