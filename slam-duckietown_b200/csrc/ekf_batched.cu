// ekf_batched.cu — fused EKF-SLAM filter-step for B independent filters, one CTA per filter (sm_100a, fp64).
//
// Reference semantics: one EKF_pose_estimation call (src/replay_no_ros.py:341-480) on an already associated
// window: augment new slots from the PRE-predict pose (:355-360), odometry predict (:368-430), then the
// observations as SEQUENTIAL range-bearing updates in the given order (:436-480).
//
// HBM layout (structure of arrays, one contiguous, 32-byte aligned record per filter and field):
//   mu [B][LD]        LD = n rounded up to 4 doubles, pad zero
//   P  [B][N*LD]      row-major, leading dimension LD, pad column(s) zero
// so a filter's covariance is ONE contiguous block: it is brought into shared memory with a single TMA
// bulk copy (cp.async.bulk + mbarrier) and leaves through coalesced 256-bit stores. Algorithmic traffic per
// filter-step: one read + one write of mu and P regardless of the number of updates.
//
// Algorithm (bitwise equal to applying the updates one after the other):
//   phase 0  TMA bulk load of P, mu; per-filter inputs into registers/shared memory
//   phase 1  augment + motion model (thread 0), predict of the robot rows/columns of P in shared memory
//   phase 2  for each update u: the 5 rows / 5 columns it touches are rebuilt from the staged P0 and the
//            earlier (K_v, W_v), giving H5, S^-1, K_u (from columns), W_u = H5 P[idx5,:] (from rows),
//            mu += K_u y_u. O(n) work per update, P0 is not modified.
//   phase 3  one pass over P: P[i,c] = P0[i,c] - sum_u K_u[i,:] W_u[:,c] (plain; rank-4 per update for
//            Joseph), each thread owning a fixed group of 4 columns (W in registers) and PASSES rows,
//            streamed to HBM with 256-bit stores.
// Rows and columns are both gathered for real — (P H^T)^T is never used for H P.
#include "ekf_batched.h"
#include "ekf_common.cuh"

namespace ekf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on the mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// one rank-2 (plain) step on a single element; the SAME sequence is used wherever an entry of an
// intermediate P_u is needed, so phase 2's rebuilt rows/columns equal what phase 3 finally writes.
__device__ __forceinline__ double rank2(double v, double k0, double k1, double w0, double w1) {
  return fma(-k1, w1, fma(-k0, w0, v));
}

constexpr int kScalDoubles = 96;

struct __align__(16) d2 {
  double x, y;
};

template <int NLM, int THREADS, bool JOSEPH>
__global__ void __launch_bounds__(THREADS) batch_step_kernel(BatchStepArgs a) {
  constexpr int N = 3 + 2 * NLM;
  constexpr int LD = (N + 3) & ~3;
  constexpr int G = LD / 4;
  constexpr int R = THREADS / G;
  constexpr int PASSES = (N + R - 1) / R;
  constexpr int HALF = THREADS / 2;
  static_assert(N <= HALF, "row and column workers are the two halves of the CTA");
  static_assert(R >= 1, "CTA narrower than one row of column groups");

  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sP = reinterpret_cast<double*>(smem_raw);  // [N][LD]
  double* sMu = sP + N * LD;                         // [LD]
  double* sScal = sMu + LD;                          // kScalDoubles doubles of scalars
  double* sK = sScal + kScalDoubles;                          // [m][LD][2]
  const int m = a.m;
  double* sW = sK + m * LD * 2;                      // [m][2][LD]
  double* sA = sW + m * 2 * LD;                      // [m][LD][2] (Joseph only)
  double* sH = sScal;                                // [2][5]
  double* sSi = sScal + 10;                          // [2][2]
  double* sSq = sScal + 14;                          // [2][2] S - Q
  double* sY = sScal + 18;                           // [2]
  double* sG = sScal + 20;                           // g0, g1
  double* sV = sScal + 24;                           // [5][5] rebuilt P[idx5, idx5]
  double* sRng = sScal + 56;                         // [BATCH_MAX_M]
  double* sBrg = sScal + 64;                         // [BATCH_MAX_M]
  int* sSlot = reinterpret_cast<int*>(sScal + 72);   // [BATCH_MAX_M]
  uint64_t* bar = reinterpret_cast<uint64_t*>(sScal + 76);

  const int tid = threadIdx.x;
  const int64_t b = blockIdx.x;
  double* __restrict__ gP = a.P + b * (int64_t)(N * LD);
  double* __restrict__ gMu = a.mu + b * (int64_t)LD;

  // ---- phase 0: TMA bulk load + per-filter inputs --------------------------------------------------
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)((N * LD + LD) * sizeof(double)));
    tma_bulk_g2s(sP, gP, (uint32_t)(N * LD * sizeof(double)), bar);
    tma_bulk_g2s(sMu, gMu, (uint32_t)(LD * sizeof(double)), bar);
  }
  double ang = 0.0, lin = 0.0, mv = 0.0;
  int nk = 0;
  if (tid == 0) {
    if (a.ticks_l) {  // a1/a2 on the device (replay_no_ros.py:262-266, :491-497)
      const double dl = delta_phi(a.ticks_l[b], a.prev_l[b], a.resolution);
      const double dr = delta_phi(a.ticks_r[b], a.prev_r[b], a.resolution);
      displacement(a.wheel_radius, a.baseline, dl, dr, &ang, &lin);
    } else {
      ang = a.ang[b];
      lin = a.lin[b];
    }
    mv = a.motion_var[b];
    nk = a.n_known[b];
  }
  const double mm = a.meas_var[b];
  double o_rng = 0.0, o_brg = 0.0;
  int o_slot = -1;
  if (tid < m) {
    o_slot = a.slot[b * m + tid];
    o_rng = a.rng[b * m + tid];
    o_brg = a.brg[b * m + tid];
    sSlot[tid] = o_slot;
    sRng[tid] = o_rng;
    sBrg[tid] = o_brg;
  }
  // zero the pads of K / W / A so pad columns of P stay exactly zero
  for (int e = tid; e < m * LD * 2; e += THREADS) {
    sK[e] = 0.0;
    sW[e] = 0.0;
    if (JOSEPH) sA[e] = 0.0;
  }
  __syncthreads();  // barrier init + staged observations visible
  mbar_wait(bar, 0);

  // ---- phase 1: augment (:355-360) + motion model (:368-417), thread 0 ------------------------------
  if (tid == 0) {
    const double px = sMu[0], py = sMu[1], pth = sMu[2];
    int fault = 0;
    for (int u = 0; u < m; ++u) {
      const int s = sSlot[u];
      if (s < 0) continue;
      if (s >= NLM || s > nk) {  // beyond capacity, or a slot that skips first-seen order (:359 KeyError)
        sSlot[u] = -1;
        fault = 1;
        continue;
      }
      if (s == nk) {
        const int j = 3 + 2 * s;
        const double r = sRng[u], be = sBrg[u];
        double sn, cs;
        sincos(be + pth, &sn, &cs);
        sMu[j] = px + r * cs;
        sMu[j + 1] = py + r * sn;
        sP[j * LD + j] = kNewLandmarkVar;
        sP[(j + 1) * LD + j + 1] = kNewLandmarkVar;
        ++nk;
      }
    }
    if (!(a.flags & kFlagMeasurement))
      for (int u = 0; u < m; ++u) sSlot[u] = -1;
    const Motion mo = motion_step(pth, ang, lin, a.flags);
    sMu[0] = px + mo.dx;
    sMu[1] = py + mo.dy;
    sMu[2] = mo.theta;
    sG[0] = mo.g0;
    sG[1] = mo.g1;
    sG[2] = mv;
    a.n_known[b] = nk;
    if (fault) atomicAdd(a.faults, 1ULL);
  }
  __syncthreads();

  // ---- predict covariance in shared memory (:420-430) -----------------------------------------------
  {
    const double g0 = sG[0], g1 = sG[1];
    if (tid >= 3 && tid < N) {  // column tid of rows 0-1
      const double p2 = sP[2 * LD + tid];
      sP[0 * LD + tid] += g0 * p2;
      sP[1 * LD + tid] += g1 * p2;
    } else if (tid >= HALF + 3 && tid < HALF + N) {  // row i of columns 0-1
      double* row = sP + (tid - HALF) * LD;
      const double p2 = row[2];
      row[0] += g0 * p2;
      row[1] += g1 * p2;
    } else if (tid == 0) {
      const double mvv = sG[2];
      double blk[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) blk[i][c] = sP[i * LD + c];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        blk[0][c] += g0 * blk[2][c];
        blk[1][c] += g1 * blk[2][c];
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        blk[i][0] += g0 * blk[i][2];
        blk[i][1] += g1 * blk[i][2];
      }
      blk[0][0] += mvv * mvv;
      blk[1][1] += mvv * mvv;
      blk[2][2] += (mvv / 2.0) * (mvv / 2.0);
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) sP[i * LD + c] = blk[i][c];
    }
  }
  __syncthreads();

  // ---- phase 2: sequential updates on the touched rows / columns only --------------------------------
  for (int u = 0; u < m; ++u) {
    const int s = sSlot[u];
    if (s < 0) continue;  // uniform across the CTA; K_u = W_u = 0 so phase 3 is a no-op for this u
    const int j = 3 + 2 * s;
    // (a) rebuilt 5x5 block, innovation, S, S^-1 — warp 0
    if (tid < 25) {
      const int k = tid / 5, c = tid - 5 * k;
      const int rk = (k < 3) ? k : j + (k - 3);
      const int cc = (c < 3) ? c : j + (c - 3);
      double v = sP[rk * LD + cc];
      for (int w = 0; w < u; ++w) {
        const double* kw = sK + (w * LD + rk) * 2;
        const double* ww = sW + w * 2 * LD;
        v = rank2(v, kw[0], kw[1], ww[cc], ww[LD + cc]);
        if (JOSEPH) {
          const double* aw = sA + (w * LD + rk) * 2;
          const double* kc = sK + (w * LD + cc) * 2;
          v = rank2(v, aw[0], aw[1], kc[0], kc[1]);
        }
      }
      sV[tid] = v;
    }
    if (tid < 32) {
      __syncwarp();
      if (tid == 0) {
        const Innov in = innovation(sMu[0], sMu[1], sMu[2], sMu[j], sMu[j + 1], sRng[u], sBrg[u]);
        double hp5[2][5];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            s0 = fma(in.h[0][k], sV[k * 5 + c], s0);
            s1 = fma(in.h[1][k], sV[k * 5 + c], s1);
          }
          hp5[0][c] = s0;
          hp5[1][c] = s1;
        }
        double S[2][2], Si[2][2];
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < 5; ++c) acc = fma(hp5[r][c], in.h[q][c], acc);
            S[r][q] = acc;
          }
        sSq[0] = S[0][0]; sSq[1] = S[0][1]; sSq[2] = S[1][0]; sSq[3] = S[1][1];
        S[0][0] += mm * mm;
        S[1][1] += mm * mm;
        inv2(S, Si);
        sSi[0] = Si[0][0]; sSi[1] = Si[0][1]; sSi[2] = Si[1][0]; sSi[3] = Si[1][1];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          sH[k] = in.h[0][k];
          sH[5 + k] = in.h[1][k];
        }
        sY[0] = in.y0;
        sY[1] = in.y1;
      }
    }
    __syncthreads();
    // (b) first half of the CTA: W_u[:, t] from the 5 ROWS; second half: K_u[i, :] from the 5 COLUMNS
    if (tid < N) {
      const int t = tid;
      double hp0 = 0.0, hp1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int rk = (k < 3) ? k : j + (k - 3);
        double v = sP[rk * LD + t];
        for (int w = 0; w < u; ++w) {
          const double* kw = sK + (w * LD + rk) * 2;
          const double* ww = sW + w * 2 * LD;
          v = rank2(v, kw[0], kw[1], ww[t], ww[LD + t]);
          if (JOSEPH) {
            const double* aw = sA + (w * LD + rk) * 2;
            const double* kc = sK + (w * LD + t) * 2;
            v = rank2(v, aw[0], aw[1], kc[0], kc[1]);
          }
        }
        hp0 = fma(sH[k], v, hp0);
        hp1 = fma(sH[5 + k], v, hp1);
      }
      sW[u * 2 * LD + t] = hp0;
      sW[u * 2 * LD + LD + t] = hp1;
    } else if (tid >= HALF && tid < HALF + N) {
      const int i = tid - HALF;
      double pht0 = 0.0, pht1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int ck = (k < 3) ? k : j + (k - 3);
        double v = sP[i * LD + ck];
        for (int w = 0; w < u; ++w) {
          const double* kw = sK + (w * LD + i) * 2;
          const double* ww = sW + w * 2 * LD;
          v = rank2(v, kw[0], kw[1], ww[ck], ww[LD + ck]);
          if (JOSEPH) {
            const double* aw = sA + (w * LD + i) * 2;
            const double* kc = sK + (w * LD + ck) * 2;
            v = rank2(v, aw[0], aw[1], kc[0], kc[1]);
          }
        }
        pht0 = fma(v, sH[k], pht0);
        pht1 = fma(v, sH[5 + k], pht1);
      }
      const double k0 = pht0 * sSi[0] + pht1 * sSi[2];
      const double k1 = pht0 * sSi[1] + pht1 * sSi[3];
      sK[(u * LD + i) * 2 + 0] = k0;
      sK[(u * LD + i) * 2 + 1] = k1;
      if (JOSEPH) {
        sA[(u * LD + i) * 2 + 0] = pht0 - (k0 * sSq[0] + k1 * sSq[2]);
        sA[(u * LD + i) * 2 + 1] = pht1 - (k0 * sSq[1] + k1 * sSq[3]);
      }
      sMu[i] += k0 * sY[0] + k1 * sY[1];  // a10 (:476)
    }
    __syncthreads();
    if (JOSEPH) {  // W = H P - Q K^T (App. A.4)
      if (tid < N) {
        const double qn = mm * mm;
        sW[u * 2 * LD + tid] -= qn * sK[(u * LD + tid) * 2 + 0];
        sW[u * 2 * LD + LD + tid] -= qn * sK[(u * LD + tid) * 2 + 1];
      }
      __syncthreads();
    }
  }

  // ---- phase 3: one pass over P, all updates applied in order, streamed to HBM -----------------------
  if (tid < G * R) {
    const int g = tid % G, rs = tid / G;
    const int c = 4 * g;
    double4x p[PASSES];
#pragma unroll
    for (int q = 0; q < PASSES; ++q) {
      const int i = rs + q * R;
      if (i < N) p[q] = *reinterpret_cast<const double4x*>(sP + i * LD + c);
    }
    for (int u = 0; u < m; ++u) {
      if (sSlot[u] < 0) continue;
      const double4x w0 = *reinterpret_cast<const double4x*>(sW + u * 2 * LD + c);
      const double4x w1 = *reinterpret_cast<const double4x*>(sW + u * 2 * LD + LD + c);
      d2 kc[4];
      if (JOSEPH) {
#pragma unroll
        for (int e = 0; e < 4; ++e) kc[e] = *reinterpret_cast<const d2*>(sK + (u * LD + c + e) * 2);
      }
#pragma unroll
      for (int q = 0; q < PASSES; ++q) {
        const int i = rs + q * R;
        if (i < N) {
          const d2 k = *reinterpret_cast<const d2*>(sK + (u * LD + i) * 2);
          p[q].a = rank2(p[q].a, k.x, k.y, w0.a, w1.a);
          p[q].b = rank2(p[q].b, k.x, k.y, w0.b, w1.b);
          p[q].c = rank2(p[q].c, k.x, k.y, w0.c, w1.c);
          p[q].d = rank2(p[q].d, k.x, k.y, w0.d, w1.d);
          if (JOSEPH) {
            const d2 av = *reinterpret_cast<const d2*>(sA + (u * LD + i) * 2);
            p[q].a = rank2(p[q].a, av.x, av.y, kc[0].x, kc[0].y);
            p[q].b = rank2(p[q].b, av.x, av.y, kc[1].x, kc[1].y);
            p[q].c = rank2(p[q].c, av.x, av.y, kc[2].x, kc[2].y);
            p[q].d = rank2(p[q].d, av.x, av.y, kc[3].x, kc[3].y);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < PASSES; ++q) {
      const int i = rs + q * R;
      if (i < N) stg256(gP + i * LD + c, p[q]);
    }
  }
  if (tid < LD) gMu[tid] = sMu[tid];
}

// -------------------------------------------------------------------------------------------------------
// init / pack / unpack
// -------------------------------------------------------------------------------------------------------
__global__ void batch_init_kernel(double* __restrict__ mu, double* __restrict__ P, int* __restrict__ n_known,
                                  int64_t batch, int n, int ld, double p0) {
  const int64_t per = (int64_t)n * ld;
  const int64_t total = batch * per;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e % per;
    const int i = (int)(r / ld), c = (int)(r % ld);
    P[e] = (i == c && i < 3) ? p0 : 0.0;
    if (e < batch * ld) mu[e] = 0.0;
    if (e < batch) n_known[e] = 0;
  }
}

__global__ void batch_pack_kernel(double* __restrict__ mu, double* __restrict__ P,
                                  const double* __restrict__ mean_ref, const double* __restrict__ cov_ref,
                                  int64_t first, int64_t count, int n, int ld) {
  const int64_t per = (int64_t)n * ld;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < count * per;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t f = e / per, r = e % per;
    const int i = (int)(r / ld), c = (int)(r % ld);
    P[(first + f) * per + r] = (c < n) ? cov_ref[f * (int64_t)n * n + (int64_t)i * n + c] : 0.0;
    if (i == 0) mu[(first + f) * ld + c] = (c < n) ? mean_ref[f * n + c] : 0.0;
  }
}

__global__ void batch_unpack_kernel(const double* __restrict__ mu, const double* __restrict__ P,
                                    double* __restrict__ mean_ref, double* __restrict__ cov_ref, int64_t first,
                                    int64_t count, int n, int ld) {
  const int64_t per = (int64_t)n * n;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < count * per;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t f = e / per, r = e % per;
    const int i = (int)(r / n), c = (int)(r % n);
    cov_ref[e] = P[(first + f) * (int64_t)n * ld + (int64_t)i * ld + c];
    if (i == 0) mean_ref[f * n + c] = mu[(first + f) * ld + c];
  }
}

__global__ void batch_pose_kernel(const double* __restrict__ mu, double* __restrict__ pose, int64_t batch,
                                  int ld) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e < batch * 3) pose[e] = mu[(e / 3) * ld + (e % 3)];
}

// -------------------------------------------------------------------------------------------------------
// launchers
// -------------------------------------------------------------------------------------------------------
size_t batch_step_smem_bytes(int n_landmarks, int m, bool joseph) {
  const int n = 3 + 2 * n_landmarks, ld = (n + 3) & ~3;
  size_t d = (size_t)n * ld + ld + kScalDoubles + (size_t)m * ld * 2 * (joseph ? 3 : 2);
  return d * sizeof(double);
}

template <int NLM, int THREADS>
static cudaError_t launch_step_t(const BatchStepArgs& a, int64_t batch, bool joseph, cudaStream_t st) {
  const size_t smem = batch_step_smem_bytes(NLM, a.m, joseph);
  cudaError_t e;
  if (joseph) {
    e = cudaFuncSetAttribute(batch_step_kernel<NLM, THREADS, true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    batch_step_kernel<NLM, THREADS, true><<<(unsigned)batch, THREADS, smem, st>>>(a);
  } else {
    e = cudaFuncSetAttribute(batch_step_kernel<NLM, THREADS, false>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    batch_step_kernel<NLM, THREADS, false><<<(unsigned)batch, THREADS, smem, st>>>(a);
  }
  return cudaGetLastError();
}

bool batch_capacity_supported(int n_landmarks) {
  switch (n_landmarks) {
    case 4: case 8: case 12: case 16: case 20: case 24: case 28: return true;
    default: return false;
  }
}

cudaError_t batch_launch_step(int n_landmarks, const BatchStepArgs& a, int64_t batch, bool joseph,
                              cudaStream_t st) {
  switch (n_landmarks) {
    case 4: return launch_step_t<4, 64>(a, batch, joseph, st);
    case 8: return launch_step_t<8, 64>(a, batch, joseph, st);
    case 12: return launch_step_t<12, 64>(a, batch, joseph, st);
    case 16: return launch_step_t<16, 128>(a, batch, joseph, st);
    case 20: return launch_step_t<20, 128>(a, batch, joseph, st);
    case 24: return launch_step_t<24, 128>(a, batch, joseph, st);
    case 28: return launch_step_t<28, 128>(a, batch, joseph, st);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t batch_launch_init(double* mu, double* P, int* n_known, int64_t batch, int n, int ld, double p0,
                              cudaStream_t st) {
  batch_init_kernel<<<148 * 8, 256, 0, st>>>(mu, P, n_known, batch, n, ld, p0);
  return cudaGetLastError();
}
cudaError_t batch_launch_pack(double* mu, double* P, const double* mean_ref, const double* cov_ref,
                              int64_t first, int64_t count, int n, int ld, cudaStream_t st) {
  batch_pack_kernel<<<148 * 8, 256, 0, st>>>(mu, P, mean_ref, cov_ref, first, count, n, ld);
  return cudaGetLastError();
}
cudaError_t batch_launch_unpack(const double* mu, const double* P, double* mean_ref, double* cov_ref,
                                int64_t first, int64_t count, int n, int ld, cudaStream_t st) {
  batch_unpack_kernel<<<148 * 8, 256, 0, st>>>(mu, P, mean_ref, cov_ref, first, count, n, ld);
  return cudaGetLastError();
}
cudaError_t batch_launch_pose(const double* mu, double* pose, int64_t batch, int ld, cudaStream_t st) {
  const int64_t tot = batch * 3;
  batch_pose_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(mu, pose, batch, ld);
  return cudaGetLastError();
}

}  // namespace ekf
