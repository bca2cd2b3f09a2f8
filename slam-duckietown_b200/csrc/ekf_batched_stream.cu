// ekf_batched_stream.cu — fused EKF-SLAM filter-step, one warp per filter, STREAMING variant (sm_100a, fp64).
//
// Same reference semantics and the same per-element arithmetic as batch_step_kernel (ekf_batched.cu; reference
// src/replay_no_ros.py:341-480), different residency: only the rows and columns of P that the step's predict and
// its m updates TOUCH are staged in shared memory —
//     rows / columns {0, 1, 2} U {j_u, j_u + 1 : u < m}        (3 + 2m of each, 6.3 KB at n = 43, m = 3)
// — and the bulk of P goes HBM -> registers -> HBM in one streaming pass that applies all m rank-2 updates.
// Round-1 ncu of the fully staged kernel showed a latency-bound chain (11 resident warps per SM, limited by
// 19.7 KB of shared memory per filter, issue slots 41 % busy). Staging 10.9 KB per filter instead keeps ~19
// filters in flight per SM.
//
// Consistency rules that make the result bit-identical to the staged kernel:
//  * predict (:420-430) changes rows 0-1, columns 0-1 and P[2][2] only. It is applied to the staged rows and
//    columns with the same FMA operands wherever an element lives in both; the streaming pass takes rows 0-2 and
//    every other staged row from shared memory, and columns 0-1 of every row from the staged column pair.
//  * augment (:355-360) writes P[j][j], P[j+1][j+1] of a landmark that is updated in this very step, so those
//    rows are staged and carry the new diagonal.
//  * rows/columns of update u are rebuilt from the staged P0 and the earlier (K_v, W_v) with rank2(), the same
//    sequence the streaming pass applies to every element.
// Plain covariance form only (the Joseph form runs on the staged kernel).
#include "ekf_batched.h"
#include "ekf_batched_dev.cuh"

namespace ekf {

__device__ __forceinline__ d2 ldg128(const double* p) {
  d2 v;
  asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

template <int NLM, int MCAP>
__global__ void __launch_bounds__(32) batch_step_stream_kernel(BatchStepArgs a) {
  constexpr int N = 3 + 2 * NLM;
  constexpr int LD = (N + 3) & ~3;
  constexpr int G = LD / 4;
  constexpr int HIST = 2 * LD;
  constexpr int NT = (LD + 31) / 32;
  static_assert(N <= 64, "staged-row mask is 64 bits");

  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int m = a.m;
  const int RS = 3 + 2 * m;                            // staged rows / columns
  double* sRow = reinterpret_cast<double*>(smem_raw);  // [RS][LD]  sRow[k][c] = P[idx_k][c]
  double* sC01 = sRow + RS * LD;                       // [LD][2]   (P[i][0], P[i][1]) pairs: staged columns 0-1
  double* sCol = sC01 + 2 * LD;                        // [RS-2][LD] sCol[k-2][i] = P[i][idx_k], k >= 2
  double* sMu = sCol + (RS - 2) * LD;                  // [LD]
  double* sK = sMu + LD;                               // [m][LD][2]
  double* sW = sK + m * HIST;                          // [m][2][LD]
  uint64_t* bar = reinterpret_cast<uint64_t*>(sW + m * HIST);
  int* sIdx = reinterpret_cast<int*>(bar + 1);         // [RS] row index of staged slot k, -1 = unused
  signed char* sRowOf = reinterpret_cast<signed char*>(sIdx + RS);  // [64] staged slot of row r, -1 = not staged

  const int lane = threadIdx.x;
  const int64_t b = blockIdx.x;
  double* __restrict__ gP = a.P + b * (int64_t)(N * LD);
  double* __restrict__ gMu = a.mu + b * (int64_t)LD;

  // ---- phase 0: inputs, slot bookkeeping, staging of the touched rows (TMA) and columns (gather) ------
  if (lane == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // start moving this filter's record HBM -> L2 now; the row copies, the column gather and the streaming pass
    // below then find it there
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gP), "r"((uint32_t)(N * LD * sizeof(double)))
                 : "memory");
  }
  double ang, lin;
  if (a.ticks_l) {  // a1/a2 on the device (replay_no_ros.py:262-266, :491-497)
    const double dl = delta_phi(a.ticks_l[b], a.prev_l[b], a.resolution);
    const double dr = delta_phi(a.ticks_r[b], a.prev_r[b], a.resolution);
    displacement(a.wheel_radius, a.baseline, dl, dr, &ang, &lin);
  } else {
    ang = a.ang[b];
    lin = a.lin[b];
  }
  const double mv = a.motion_var[b];
  const double mm = a.meas_var[b];
  const double qn = mm * mm;
  int nk = a.n_known[b];
  double o_rng = 0.0, o_brg = 0.0;
  int o_slot = -1;
  if (lane < m) {
    o_slot = a.slot[b * m + lane];
    o_rng = a.rng[b * m + lane];
    o_brg = a.brg[b * m + lane];
  }
  // validity / first-sighting bookkeeping (:294-295, :359), warp-uniform
  unsigned valid = 0, aug = 0;
  int fault = 0;
  for (int u = 0; u < m; ++u) {
    const int s = __shfl_sync(kFull, o_slot, u);
    if (s < 0) continue;
    if (s >= NLM || s > nk) {  // beyond capacity, or a slot that skips first-seen order (:359 KeyError)
      fault = 1;
      continue;
    }
    valid |= 1u << u;
    if (s == nk) {
      aug |= 1u << u;
      ++nk;
    }
  }
  {
    const int u = (lane >= 3) ? ((lane - 3) >> 1) : 0;
    const int s = __shfl_sync(kFull, o_slot, u & 31);
    int r = lane;
    if (lane >= 3) r = ((valid >> u) & 1u) ? 3 + 2 * s + ((lane - 3) & 1) : -1;
    sRowOf[lane] = -1;
    sRowOf[lane + 32] = -1;
    __syncwarp();
    if (lane < RS) {
      sIdx[lane] = r;
      if (r >= 0) sRowOf[r] = (signed char)lane;  // a landmark observed twice: either copy serves
    }
  }
  __syncwarp();
  if (lane == 0) {
    int nv = __popc(valid);
    mbar_expect_tx(bar, (uint32_t)((4 + 2 * nv) * LD * sizeof(double)));
    tma_bulk_g2s(sMu, gMu, (uint32_t)(LD * sizeof(double)), bar);
    tma_bulk_g2s(sRow, gP, (uint32_t)(3 * LD * sizeof(double)), bar);
    for (int u = 0; u < m; ++u)
      if ((valid >> u) & 1u) {
        const int j = sIdx[3 + 2 * u];
        tma_bulk_g2s(sRow + (3 + 2 * u) * LD, gP + j * LD, (uint32_t)(2 * LD * sizeof(double)), bar);
      }
  }
  // columns: lane i reads P[i][0..3] (one aligned 256-bit load) and P[i][j_u], P[i][j_u + 1]
  {
    int ju[MCAP];
#pragma unroll
    for (int u = 0; u < MCAP; ++u) ju[u] = (u < m) ? sIdx[3 + 2 * u] : -1;
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int i = lane + 32 * p;
      if (i < LD) {
        double4x c03 = {0.0, 0.0, 0.0, 0.0};
        double cj[MCAP][2];
        if (i < N) {
          const double* row = gP + i * LD;
          c03 = ldg256(row);
#pragma unroll
          for (int u = 0; u < MCAP; ++u)
            if (ju[u] >= 0) {
              cj[u][0] = row[ju[u]];
              cj[u][1] = row[ju[u] + 1];
            }
        }
        d2 c01 = {c03.a, c03.b};  // rows >= N (the pad) stay zero: K of the pad row is zero
        *reinterpret_cast<d2*>(sC01 + 2 * i) = c01;
        sCol[i] = c03.c;
#pragma unroll
        for (int u = 0; u < MCAP; ++u)
          if (ju[u] >= 0) {
            sCol[(1 + 2 * u) * LD + i] = (i < N) ? cj[u][0] : 0.0;
            sCol[(2 + 2 * u) * LD + i] = (i < N) ? cj[u][1] : 0.0;
          }
      }
    }
  }
  __syncwarp();
  mbar_wait(bar, 0);

  // ---- phase 1: augment (:355-360) + motion model (:368-417), predict on the staged rows / columns ----
  {
    const double px = sMu[0], py = sMu[1], pth = sMu[2];
    double sn, cs;
    fm_sincos((lane & 1) ? pth + ang : pth, &sn, &cs);
    const double s1 = __shfl_sync(kFull, sn, 0), c1 = __shfl_sync(kFull, cs, 0);
    const double s2 = __shfl_sync(kFull, sn, 1), c2 = __shfl_sync(kFull, cs, 1);
    for (int u = 0; u < m; ++u) {
      if (!((aug >> u) & 1u)) continue;  // first sighting: initialise from the PRE-predict pose (:331-332)
      const int s = __shfl_sync(kFull, o_slot, u);
      const double r = __shfl_sync(kFull, o_rng, u), be = __shfl_sync(kFull, o_brg, u);
      const int j = 3 + 2 * s;
      double sb, cb;
      fm_sincos(be + pth, &sb, &cb);
      if (lane == 0) {
        sMu[j] = __dadd_rn(px, __dmul_rn(r, cb));
        sMu[j + 1] = __dadd_rn(py, __dmul_rn(r, sb));
      }
      // every staged copy of rows / columns j, j+1 gets the new diagonal (a slot may be observed twice)
      if (lane < m && o_slot == s && ((valid >> lane) & 1u)) {
        const int k = 3 + 2 * lane;
        sRow[k * LD + j] = kNewLandmarkVar;
        sRow[(k + 1) * LD + j + 1] = kNewLandmarkVar;
        sCol[(k - 2) * LD + j] = kNewLandmarkVar;
        sCol[(k - 1) * LD + j + 1] = kNewLandmarkVar;
      }
    }
    if (!(a.flags & kFlagMeasurement)) valid = 0;
    const Motion mo = motion_from_sincos(pth, ang, lin, a.flags, s1, c1, s2, c2);
    if (lane == 0) {
      sMu[0] = px + mo.dx;
      sMu[1] = py + mo.dy;
      sMu[2] = mo.theta;
      a.n_known[b] = nk;
      if (fault) atomicAdd(a.faults, 1ULL);
    }
    __syncwarp();
    const double g0 = mo.g0, g1 = mo.g1;
    // row operation P[0,:] += g0 P[2,:], P[1,:] += g1 P[2,:] — staged rows 0-1 (all columns) and rows 0-1 of
    // every staged column
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      if (t < N) {
        const double p2 = sRow[2 * LD + t];
        sRow[0 * LD + t] = fma(g0, p2, sRow[0 * LD + t]);
        sRow[1 * LD + t] = fma(g1, p2, sRow[1 * LD + t]);
      }
    }
    if (lane < RS) {  // staged column k = lane: elements i = 0, 1 += g * element i = 2
      double* e0 = (lane < 2) ? sC01 + lane : sCol + (lane - 2) * LD;
      const int st = (lane < 2) ? 2 : 1;  // stride between consecutive i
      const double p2 = e0[2 * st];
      e0[0] = fma(g0, p2, e0[0]);
      e0[st] = fma(g1, p2, e0[st]);
    }
    __syncwarp();
    // column operation P[:,0] += g0 P[:,2], P[:,1] += g1 P[:,2] (with the updated rows 0-1), then F^T R F
    if (lane < RS) {
      double* row = sRow + lane * LD;
      const double p2 = row[2];
      double r0 = fma(g0, p2, row[0]), r1 = fma(g1, p2, row[1]);
      if (lane == 0) r0 = __dadd_rn(r0, __dmul_rn(mv, mv));
      if (lane == 1) r1 = __dadd_rn(r1, __dmul_rn(mv, mv));
      row[0] = r0;
      row[1] = r1;
      if (lane == 2) row[2] = __dadd_rn(p2, __dmul_rn(mv / 2.0, mv / 2.0));
    }
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int i = lane + 32 * p;
      if (i < N) {
        const double p2 = sCol[i];
        d2 c01 = lds128(sC01 + 2 * i);
        c01.x = fma(g0, p2, c01.x);
        c01.y = fma(g1, p2, c01.y);
        if (i == 0) c01.x = __dadd_rn(c01.x, __dmul_rn(mv, mv));
        if (i == 1) c01.y = __dadd_rn(c01.y, __dmul_rn(mv, mv));
        *reinterpret_cast<d2*>(sC01 + 2 * i) = c01;
        if (i == 2) sCol[i] = __dadd_rn(p2, __dmul_rn(mv / 2.0, mv / 2.0));
      }
    }
    __syncwarp();
  }

  // ---- phase 2: sequential updates on the staged rows / columns --------------------------------------
  // Update u touches rows/columns {0,1,2,j,j+1} = staged slots {0,1,2,3+2u,4+2u}. Lane t owns column t of those
  // rows (-> W_u[:, t]) and row t of those columns (-> K_u[t, :]) for t = lane, lane + 32.
  for (int u = 0; u < m; ++u) {
    if (!((valid >> u) & 1u)) continue;  // warp-uniform
    const int s = __shfl_sync(kFull, o_slot, u);
    const double z_rng = __shfl_sync(kFull, o_rng, u), z_brg = __shfl_sync(kFull, o_brg, u);
    const int j = 3 + 2 * s;
    const Innov in = innovation(sMu[0], sMu[1], sMu[2], sMu[j], sMu[j + 1], z_rng, z_brg);
    double vr[NT][5], vc[NT][5];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int sk = (k < 3) ? k : 2 * u + k;  // staged slot of row/column idx5[k]
        vr[p][k] = (t < LD) ? sRow[sk * LD + t] : 0.0;
        if (k >= 2) vc[p][k] = (t < LD) ? sCol[(sk - 2) * LD + t] : 0.0;
      }
      const d2 c01 = (t < LD) ? lds128(sC01 + 2 * t) : d2{0.0, 0.0};
      vc[p][0] = c01.x;
      vc[p][1] = c01.y;
    }
    for (int w = 0; w < u; ++w) {
      if (!((valid >> w) & 1u)) continue;
      const double* kw = sK + w * HIST;  // [LD][2]
      const double* ww = sW + w * HIST;  // [2][LD]
      d2 kr[5];
      double wr0[5], wr1[5];
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int rk = (k < 3) ? k : j + (k - 3);
        kr[k] = lds128(kw + 2 * rk);
        wr0[k] = ww[rk];
        wr1[k] = ww[LD + rk];
      }
#pragma unroll
      for (int p = 0; p < NT; ++p) {
        const int t = lane + 32 * p;
        if (t < LD) {
          const double wt0 = ww[t], wt1 = ww[LD + t];
          const d2 kt = lds128(kw + 2 * t);
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            vr[p][k] = rank2(vr[p][k], kr[k].x, kr[k].y, wt0, wt1);
            vc[p][k] = rank2(vc[p][k], kt.x, kt.y, wr0[k], wr1[k]);
          }
        }
      }
    }
    double pht0[NT], pht1[NT];
    double* wu = sW + u * HIST;
    double* ku = sK + u * HIST;
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        a0 = fma(in.h[0][k], vr[p][k], a0);
        a1 = fma(in.h[1][k], vr[p][k], a1);
        b0 = fma(vc[p][k], in.h[0][k], b0);
        b1 = fma(vc[p][k], in.h[1][k], b1);
      }
      pht0[p] = b0;
      pht1[p] = b1;
      const int t = lane + 32 * p;
      if (t < LD) {
        wu[t] = a0;
        wu[LD + t] = a1;
      }
    }
    __syncwarp();
    double S[2][2], Si[2][2];
    {
      double acc00 = 0.0, acc01 = 0.0, acc10 = 0.0, acc11 = 0.0;
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int cc = (c < 3) ? c : j + (c - 3);
        const double x0 = wu[cc], x1 = wu[LD + cc];
        acc00 = fma(x0, in.h[0][c], acc00);
        acc01 = fma(x0, in.h[1][c], acc01);
        acc10 = fma(x1, in.h[0][c], acc10);
        acc11 = fma(x1, in.h[1][c], acc11);
      }
      S[0][0] = acc00 + qn; S[0][1] = acc01; S[1][0] = acc10; S[1][1] = acc11 + qn;
    }
    inv2(S, Si);
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      if (t < LD) {
        d2 kk;
        kk.x = dot2(pht0[p], Si[0][0], pht1[p], Si[1][0]);
        kk.y = dot2(pht0[p], Si[0][1], pht1[p], Si[1][1]);
        *reinterpret_cast<d2*>(ku + 2 * t) = kk;  // zero for the pad row
        if (t < N) sMu[t] = mu_inc(sMu[t], kk.x, kk.y, in.y0, in.y1);
      }
    }
    __syncwarp();
  }

  // ---- phase 3: stream P through registers, all updates applied in order ----------------------------------
#pragma unroll
  for (int p = 0; p < NT; ++p) {
    const int t = lane + 32 * p;
    if (t < LD) gMu[t] = sMu[t];
  }
  {
    constexpr int RP = 32 / G;              // rows per iteration
    constexpr int NIT = (N + RP - 1) / RP;  // row iterations
    constexpr int CH = 4;                   // row iterations per chunk
    constexpr int NCH = (NIT + CH - 1) / CH;
    constexpr int HALF = LD / 2;
    const int ph = lane / G, g = lane - ph * G;
    const int ca = 2 * g, cb = HALF + 2 * g;
    if (ph < RP) {
      // Every row in one software-pipelined pass: the loads of chunk c+1 are in flight while chunk c is updated.
      // A row is always fetched from HBM with immediate offsets off one base pointer; a staged row (predicted /
      // augmented copy in shared memory) then overrides it, and the lanes of columns 0-1 (g == 0) always take the
      // staged column pair.
      const double* gbase = gP + ph * LD;
      d2 pa[2][CH], pb[2][CH];
      auto load_chunk = [&](int c, d2* xa, d2* xb) {
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          const int it = c * CH + q;
          const int r = it * RP + ph;
          if (it < NIT && r < N) {
            xa[q] = ldg128(gbase + it * RP * LD + ca);
            xb[q] = ldg128(gbase + it * RP * LD + cb);
          }
        }
      };
      // staged copies override the streamed values only when the chunk is about to be used: overwriting the
      // destination register of a pending global load would stall on it and undo the pipelining
      auto override_chunk = [&](int c, d2* xa, d2* xb) {
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          const int it = c * CH + q;
          const int r = it * RP + ph;
          if (it < NIT && r < N) {
            const int k = sRowOf[r];
            if (k >= 0) {
              xa[q] = lds128(sRow + k * LD + ca);
              xb[q] = lds128(sRow + k * LD + cb);
            }
            if (g == 0) xa[q] = lds128(sC01 + 2 * r);
          }
        }
      };
      load_chunk(0, pa[0], pb[0]);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        d2* xa = pa[c & 1];
        d2* xb = pb[c & 1];
        if (c + 1 < NCH) load_chunk(c + 1, pa[(c + 1) & 1], pb[(c + 1) & 1]);
        override_chunk(c, xa, xb);
        for (int u = 0; u < m; ++u) {
          if (!((valid >> u) & 1u)) continue;
          const double* ku = sK + u * HIST;
          const double* wu = sW + u * HIST;
          const d2 w0a = lds128(wu + ca), w0b = lds128(wu + cb);
          const d2 w1a = lds128(wu + LD + ca), w1b = lds128(wu + LD + cb);
#pragma unroll
          for (int q = 0; q < CH; ++q) {
            const int it = c * CH + q;
            const int r = it * RP + ph;
            if (it < NIT && r < N) {
              const d2 k = lds128(ku + 2 * r);
              xa[q].x = rank2(xa[q].x, k.x, k.y, w0a.x, w1a.x);
              xa[q].y = rank2(xa[q].y, k.x, k.y, w0a.y, w1a.y);
              xb[q].x = rank2(xb[q].x, k.x, k.y, w0b.x, w1b.x);
              xb[q].y = rank2(xb[q].y, k.x, k.y, w0b.y, w1b.y);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          const int it = c * CH + q;
          const int r = it * RP + ph;
          if (it < NIT && r < N) {
            stg128(gP + r * LD + ca, xa[q]);
            stg128(gP + r * LD + cb, xb[q]);
          }
        }
      }
    }
  }
}

// -------------------------------------------------------------------------------------------------------
size_t batch_stream_smem_bytes(int n_landmarks, int m) {
  const int n = 3 + 2 * n_landmarks, ld = (n + 3) & ~3;
  const int rs = 3 + 2 * m;
  size_t bytes = ((size_t)2 * rs * ld + ld + (size_t)m * 4 * ld) * sizeof(double);
  bytes += 8 + (size_t)rs * sizeof(int) + 64;
  return (bytes + 15) & ~(size_t)15;
}

template <int NLM>
static cudaError_t launch_stream_t(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  const size_t smem = batch_stream_smem_bytes(NLM, a.m);
  cudaError_t e;
  if (a.m <= 4) {
    auto kernel = batch_step_stream_kernel<NLM, 4>;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
      return e;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared)) != cudaSuccess)
      return e;
    kernel<<<(unsigned)batch, 32, smem, st>>>(a);
  } else {
    auto kernel = batch_step_stream_kernel<NLM, BATCH_MAX_M>;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
      return e;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared)) != cudaSuccess)
      return e;
    kernel<<<(unsigned)batch, 32, smem, st>>>(a);
  }
  return cudaGetLastError();
}

cudaError_t batch_launch_step_stream(int n_landmarks, const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  switch (n_landmarks) {
    case 4: return launch_stream_t<4>(a, batch, st);
    case 8: return launch_stream_t<8>(a, batch, st);
    case 12: return launch_stream_t<12>(a, batch, st);
    case 16: return launch_stream_t<16>(a, batch, st);
    case 20: return launch_stream_t<20>(a, batch, st);
    case 24: return launch_stream_t<24>(a, batch, st);
    case 28: return launch_stream_t<28>(a, batch, st);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace ekf
