// ekf_batched_group.cu — fused EKF-SLAM filter-step for SMALL filters: a GROUP of GW lanes (8 or 16) per filter,
// 32 / GW filters per warp (sm_100a, fp64).
//
// Same reference semantics (src/replay_no_ros.py:341-480), the same algorithm and the same per-element operation
// sequence as batch_step_regw_kernel (ekf_batched_regw.cu) — results are bitwise equal. What changes is the SIMD
// mapping. With a whole warp per filter the group-uniform scalar chain of a step (sincos, m x {innovation, atan2,
// sqrt, reciprocals, S, S^-1}) is issued once per FILTER and computed 32 times redundantly; at n = 19 that chain is
// ~45 % of the 1 376 warp-instructions of a filter-step and the kernel is issue/latency bound at 47 % of the HBM
// peak (ncu, profiles/r1h_batchM.*). Here one instruction stream serves 32 / GW filters: lanes of a group own
// columns t = gl, gl + GW, ... of THEIR filter's covariance (staged in the group's slice of shared memory), the
// scalar chain of the 32 / GW filters runs in the same instructions, and shuffles / barriers use the group's
// lane mask, so groups whose observation patterns differ may diverge without deadlock.
// Plain covariance form, compile-time observation count m <= GW.
#include <cstdlib>
#include "ekf_batched.h"
#include "ekf_batched_dev.cuh"

namespace ekf {

namespace {

template <int GW>
__device__ __forceinline__ double gshfl(unsigned gmask, double v, int src) {
  return __shfl_sync(gmask, v, src, GW);
}
template <int GW>
__device__ __forceinline__ int gshfl(unsigned gmask, int v, int src) {
  return __shfl_sync(gmask, v, src, GW);
}

// value of column c (group-uniform) from the lane of the group that owns it (column c lives in lane c % GW, half
// c / GW): every lane offers its value of that half
template <int GW, int NT>
__device__ __forceinline__ double gcol(unsigned gmask, const double (&x)[NT], int c) {
  const int p = c / GW;
  double mine = x[0];
#pragma unroll
  for (int q = 1; q < NT; ++q)
    if (p == q) mine = x[q];
  return gshfl<GW>(gmask, mine, c - p * GW);
}

__host__ __device__ constexpr int group_slice_doubles(int nlm, int m, int gw, bool joseph) {
  const int n = 3 + 2 * nlm, ld = (n + 3) & ~3;
  const int raw = n * ld + ld + m * 2 * ld * (joseph ? 2 : 1) + 2;
  return gw == 8 ? raw + ((8 - raw % 16) + 16) % 16 : raw;
}

}  // namespace

// 16 resident one-warp CTAs per SM: 128 registers per thread = four warps per scheduler's register file
// JOSEPH: one-pass rank-4 Joseph form, same per-element operation sequence as the staged kernel
template <int NLM, int MT, int GW, bool JOSEPH, bool GEN>
__global__ void __launch_bounds__(32, JOSEPH ? 12 : 16) batch_step_group_kernel(BatchStepArgs a, int64_t batch, int pf) {
  constexpr int N = 3 + 2 * NLM;
  constexpr int LD = (N + 3) & ~3;
  constexpr int HIST = 2 * LD;          // doubles of one update's K record: [LD][2]
  constexpr int NT = (LD + GW - 1) / GW;  // lane gl handles indices gl, gl + GW, ...
  constexpr int FPW = 32 / GW;          // filters per warp
  constexpr int m = MT;
  // doubles of shared memory per filter; with four groups the slices start 64 bytes apart modulo the 128-byte bank
  // period, so the four 64-byte row segments the groups read in one instruction fall into two full wavefronts
  constexpr int SLICE = group_slice_doubles(NLM, MT, GW, JOSEPH);
  static_assert(GW == 8 || GW == 16, "group width");
  static_assert(MT >= 1 && MT <= GW && MT <= BATCH_MAX_M, "observation count");
  static_assert(SLICE % 2 == 0, "16-byte aligned slices");

  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int g = lane / GW, gl = lane - g * GW;
  const unsigned gmask = (GW == 16 ? 0xffffu : 0xffu) << (g * GW);
  const int64_t b = (int64_t)blockIdx.x * FPW + g;
  if (b >= batch) return;  // the whole group leaves; the other groups only ever synchronise among themselves
  double* sP = reinterpret_cast<double*>(smem_raw) + g * SLICE;  // [N][LD]
  double* sMu = sP + N * LD;                                     // [LD]
  double* sK = sMu + LD;                                         // [m][LD][2]  (K_u[i,0], K_u[i,1]) pairs
  double* sA = sK + m * HIST;                                    // [m][LD][2]  AHt_u pairs (Joseph only)
  uint64_t* bar = reinterpret_cast<uint64_t*>(sA + (JOSEPH ? m * HIST : 0));  // [2]: mu, P
  double* __restrict__ gP = a.P + b * (int64_t)(N * LD);
  double* __restrict__ gMu = a.mu + b * (int64_t)LD;

  // ---- phase 0: TMA bulk loads (mu on its own mbarrier) + per-filter inputs ------------------------------
  if (gl == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)(LD * sizeof(double)));
    tma_bulk_g2s(sMu, gMu, (uint32_t)(LD * sizeof(double)), bar);
    mbar_expect_tx(bar + 1, (uint32_t)(N * LD * sizeof(double)));
    tma_bulk_g2s(sP, gP, (uint32_t)(N * LD * sizeof(double)), bar + 1);
  }
  // L2 prefetch for the filters that will take this CTA's place on the SM (pf CTAs ahead, ekf_batched_dev.cuh): one
  // request per array and filter, spread over the lanes of the group
  if (pf > 0) {
    const int64_t bn = b + (int64_t)pf * FPW;
    if (bn < batch) {
      const void* src = nullptr;
      switch (gl) {
        case 1: src = a.ticks_l ? (const void*)(a.ticks_l + bn) : (a.ang ? (const void*)(a.ang + bn) : nullptr); break;
        case 2: src = a.ticks_l ? (const void*)(a.ticks_r + bn) : (a.lin ? (const void*)(a.lin + bn) : nullptr); break;
        case 3: src = a.motion_var + bn; break;
        case 4: src = a.meas_var + bn; break;
        case 5: src = a.n_known + bn; break;
        case 6: src = a.slot ? a.slot + bn * m : nullptr; break;
        case 7: src = a.rng ? a.rng + bn * m : nullptr; break;
        default: break;
      }
      if (gl == 0) {
        l2_prefetch_bulk(a.P + bn * (int64_t)(N * LD), (uint32_t)(N * LD * sizeof(double)));
        l2_prefetch_bulk(a.mu + bn * (int64_t)LD, (uint32_t)(LD * sizeof(double)));
        if (a.brg) l2_prefetch(a.brg + bn * m);
      } else if (src) {
        l2_prefetch(src);
      }
    }
  }
  double ang, lin;
  if (a.ticks_common) {  // a1/a2 on the device from ticks every filter shares (Monte Carlo replay of one bag)
    const double dl = delta_phi(a.c_tl, a.c_pl, a.resolution);
    const double dr = delta_phi(a.c_tr, a.c_pr, a.resolution);
    displacement(a.wheel_radius, a.baseline, dl, dr, &ang, &lin);
  } else if (a.ticks_l) {  // a1/a2 on the device (replay_no_ros.py:262-266, :491-497)
    const double dl = delta_phi(a.ticks_l[b], a.prev_l[b], a.resolution);
    const double dr = delta_phi(a.ticks_r[b], a.prev_r[b], a.resolution);
    displacement(a.wheel_radius, a.baseline, dl, dr, &ang, &lin);
  } else if (!GEN) {
    ang = a.ang[b];
    lin = a.lin[b];
  }
  const double mv = a.motion_var[b];
  const double mm = a.meas_var[b];
  const double qn = mm * mm;
  int nk = a.n_known[b];
  double o_rng = 0.0, o_brg = 0.0;  // observation u lives in lane u of the group
  int o_slot = -1;
  if constexpr (GEN) {  // this filter's Monte Carlo inputs, drawn here while its state is still in flight
    double z0 = 0.0, z1 = 0.0;
    if (gl <= m)
      normal_pair(a.gen_seed, a.gen_first + (unsigned long long)b, a.gen_step,
                  gl < m ? kStreamObservation0 + gl : kStreamOdometry, &z0, &z1);
    if (gl < m) {
      o_slot = reinterpret_cast<const int32_t*>(a.gen_record + 2 + 2 * m)[gl];
      o_rng = __dadd_rn(a.gen_record[2 + gl], __dmul_rn(a.gen_s_rng, z0));
      o_brg = __dadd_rn(a.gen_record[2 + m + gl], __dmul_rn(a.gen_s_brg, z1));
    }
    ang = gshfl<GW>(gmask, __dadd_rn(a.gen_record[0], __dmul_rn(a.gen_s_ang, z0)), m);
    lin = gshfl<GW>(gmask, __dadd_rn(a.gen_record[1], __dmul_rn(a.gen_s_lin, z1)), m);
  } else if (gl < m) {
    o_slot = a.slot[b * m + gl];
    o_rng = a.rng[b * m + gl];
    o_brg = a.brg[b * m + gl];
  }
  __syncwarp(gmask);
  mbar_wait(bar, 0);

  // robot rows / columns of the CURRENT covariance, lane t: rr[k][p] = P[k][t], rc[k][p] = P[t][k], t = gl + GW p
  double rr[3][NT], rc[3][NT];

  // ---- phase 1: augment (:355-360) + motion model (:368-417), group-uniform ------------------------------
  unsigned valid = 0;   // bit u: update u is applied
  unsigned fresh = 0;   // bit u: update u is the first sighting of its landmark
  int sl[m];            // slot of observation u (group-uniform)
  {
    const double px = sMu[0], py = sMu[1], pth = sMu[2];
    double sn, cs;  // even lanes sincos(theta), odd lanes sincos(theta + ang): one call instead of two
    fm_sincos((gl & 1) ? pth + ang : pth, &sn, &cs);
    const double s1 = gshfl<GW>(gmask, sn, 0), c1 = gshfl<GW>(gmask, cs, 0);
    const double s2 = gshfl<GW>(gmask, sn, 1), c2 = gshfl<GW>(gmask, cs, 1);
    int fault = 0;
#pragma unroll
    for (int u = 0; u < m; ++u) {
      const int s = gshfl<GW>(gmask, o_slot, u);
      sl[u] = s;
      if (s < 0) continue;
      if (s >= NLM || s > nk) {  // beyond capacity, or a slot that skips first-seen order (:359 KeyError)
        fault = 1;
        continue;
      }
      valid |= 1u << u;
      if (s == nk) {  // first sighting: initialise from the PRE-predict pose (:331-332, :359-360)
        const int j = 3 + 2 * s;
        const double r = gshfl<GW>(gmask, o_rng, u), be = gshfl<GW>(gmask, o_brg, u);
        double sb, cb;
        fm_sincos(be + pth, &sb, &cb);
        if (gl == 0) {
          sMu[j] = __dadd_rn(px, __dmul_rn(r, cb));
          sMu[j + 1] = __dadd_rn(py, __dmul_rn(r, sb));
        }
        fresh |= 1u << u;
        ++nk;
      }
    }
    if (!(a.flags & kFlagMeasurement)) valid = 0;
    const Motion mo = motion_from_sincos(pth, ang, lin, a.flags, s1, c1, s2, c2);
    if (gl == 0) {
      sMu[0] = px + mo.dx;
      sMu[1] = py + mo.dy;
      sMu[2] = mo.theta;
      a.n_known[b] = nk;
      if (fault) atomicAdd(a.faults, 1ULL);
    }
    mbar_wait(bar + 1, 0);
    if (fresh && gl == 0) {  // new landmarks' diagonals (:356-357)
#pragma unroll
      for (int u = 0; u < m; ++u)
        if ((fresh >> u) & 1u) {
          const int j = 3 + 2 * sl[u];
          sP[j * LD + j] = kNewLandmarkVar;
          sP[(j + 1) * LD + j + 1] = kNewLandmarkVar;
        }
    }
    __syncwarp(gmask);
    // predict covariance (:420-430): rows 0-1 += g * row 2 over ALL columns, then columns 0-1 += g * column 2
    // over ALL rows (using the updated rows 0-1), then the three noise terms
    const double g0 = mo.g0, g1 = mo.g1;
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = gl + GW * p;
      rr[0][p] = rr[1][p] = rr[2][p] = 0.0;
      if (t < N) {
        const double p2 = sP[2 * LD + t];
        rr[0][p] = fma(g0, p2, sP[0 * LD + t]);
        rr[1][p] = fma(g1, p2, sP[1 * LD + t]);
        rr[2][p] = p2;
        sP[0 * LD + t] = rr[0][p];
        sP[1 * LD + t] = rr[1][p];
      }
    }
    __syncwarp(gmask);
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int i = gl + GW * p;
      rc[0][p] = rc[1][p] = rc[2][p] = 0.0;
      if (i < N) {
        double* row = sP + i * LD;
        const d2 r01 = lds128(row);
        const double p2 = row[2];
        double r0 = fma(g0, p2, r01.x), r1 = fma(g1, p2, r01.y), r2 = p2;
        if (i == 0) r0 = __dadd_rn(r0, __dmul_rn(mv, mv));  // F^T R F (:421-430)
        if (i == 1) r1 = __dadd_rn(r1, __dmul_rn(mv, mv));
        if (i == 2) r2 = __dadd_rn(p2, __dmul_rn(mv / 2.0, mv / 2.0));
        d2 o;
        o.x = r0; o.y = r1;
        *reinterpret_cast<d2*>(row) = o;
        if (i == 2) row[2] = r2;
        rc[0][p] = r0; rc[1][p] = r1; rc[2][p] = r2;
      }
    }
    __syncwarp(gmask);
    if (gl < 3) {  // the 3x3 robot block was changed by the column pass: refresh the row registers
      rr[0][0] = sP[0 * LD + gl];
      rr[1][0] = sP[1 * LD + gl];
      rr[2][0] = sP[2 * LD + gl];
    }
  }

  // ---- phase 2: sequential updates on the touched rows / columns only ------------------------------------
  // history in registers: hW[u][.][p] = W_u[:, t], hK[u][p] = K_u[t, :] of this lane's t = gl + GW p
  double hW0[m][NT], hW1[m][NT];
  d2 hK[m][NT];
#pragma unroll
  for (int u = 0; u < m; ++u) {
    if (!((valid >> u) & 1u)) continue;  // group-uniform
    const double z_rng = gshfl<GW>(gmask, o_rng, u), z_brg = gshfl<GW>(gmask, o_brg, u);
    const int j = 3 + 2 * sl[u];
    // (a) innovation and H5 (:443-469), group-uniform
    const Innov in = innovation(sMu[0], sMu[1], sMu[2], sMu[j], sMu[j + 1], z_rng, z_brg);
    // (b) landmark rows / columns j, j+1 from the staged P0, brought up to date with the earlier updates
    double lr[2][NT], lc[2][NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = gl + GW * p;
      const bool inld = (t < LD), real = (t < N);
      lr[0][p] = inld ? sP[j * LD + t] : 0.0;
      lr[1][p] = inld ? sP[(j + 1) * LD + t] : 0.0;
      lc[0][p] = real ? sP[t * LD + j] : 0.0;
      lc[1][p] = real ? sP[t * LD + j + 1] : 0.0;
    }
#pragma unroll
    for (int w = 0; w < m - 1; ++w) {  // static trip count keeps the register history statically indexed
      if (w >= u || !((valid >> w) & 1u)) continue;
      const d2 kj0 = lds128(sK + w * HIST + 2 * j), kj1 = lds128(sK + w * HIST + 2 * j + 2);
      const double w0j0 = gcol<GW, NT>(gmask, hW0[w], j), w1j0 = gcol<GW, NT>(gmask, hW1[w], j);
      const double w0j1 = gcol<GW, NT>(gmask, hW0[w], j + 1), w1j1 = gcol<GW, NT>(gmask, hW1[w], j + 1);
      d2 aj0, aj1;
      if (JOSEPH) {
        aj0 = lds128(sA + w * HIST + 2 * j);
        aj1 = lds128(sA + w * HIST + 2 * j + 2);
      }
#pragma unroll
      for (int p = 0; p < NT; ++p) {
        lr[0][p] = rank2(lr[0][p], kj0.x, kj0.y, hW0[w][p], hW1[w][p]);
        lr[1][p] = rank2(lr[1][p], kj1.x, kj1.y, hW0[w][p], hW1[w][p]);
        lc[0][p] = rank2(lc[0][p], hK[w][p].x, hK[w][p].y, w0j0, w1j0);
        lc[1][p] = rank2(lc[1][p], hK[w][p].x, hK[w][p].y, w0j1, w1j1);
        if (JOSEPH) {
          const int t = gl + GW * p;
          d2 aw;
          aw.x = 0.0; aw.y = 0.0;
          if (t < LD) aw = lds128(sA + w * HIST + 2 * t);
          lr[0][p] = rank2(lr[0][p], aj0.x, aj0.y, hK[w][p].x, hK[w][p].y);
          lr[1][p] = rank2(lr[1][p], aj1.x, aj1.y, hK[w][p].x, hK[w][p].y);
          lc[0][p] = rank2(lc[0][p], aw.x, aw.y, kj0.x, kj0.y);
          lc[1][p] = rank2(lc[1][p], aw.x, aw.y, kj1.x, kj1.y);
        }
      }
    }
    // (c) H P (rows) and P H^T (columns) for this lane's t, accumulated in idx5 order
    double pht0[NT], pht1[NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double vr = (k < 3) ? rr[k < 3 ? k : 0][p] : lr[k < 3 ? 0 : k - 3][p];
        const double vc = (k < 3) ? rc[k < 3 ? k : 0][p] : lc[k < 3 ? 0 : k - 3][p];
        a0 = fma(in.h[0][k], vr, a0);
        a1 = fma(in.h[1][k], vr, a1);
        b0 = fma(vc, in.h[0][k], b0);
        b1 = fma(vc, in.h[1][k], b1);
      }
      hW0[u][p] = a0; hW1[u][p] = a1; pht0[p] = b0; pht1[p] = b1;
    }
    // (d) S = (H P) H^T + Q from the five entries idx5 of H P (:473), closed-form inverse; group-uniform
    double x0[5], x1[5];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      x0[c] = gshfl<GW>(gmask, hW0[u][0], c);
      x1[c] = gshfl<GW>(gmask, hW1[u][0], c);
    }
    x0[3] = gcol<GW, NT>(gmask, hW0[u], j);     x1[3] = gcol<GW, NT>(gmask, hW1[u], j);
    x0[4] = gcol<GW, NT>(gmask, hW0[u], j + 1); x1[4] = gcol<GW, NT>(gmask, hW1[u], j + 1);
    double S[2][2], Si[2][2], Sq[2][2];
    {
      double acc00 = 0.0, acc01 = 0.0, acc10 = 0.0, acc11 = 0.0;
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        acc00 = fma(x0[c], in.h[0][c], acc00);
        acc01 = fma(x0[c], in.h[1][c], acc01);
        acc10 = fma(x1[c], in.h[0][c], acc10);
        acc11 = fma(x1[c], in.h[1][c], acc11);
      }
      Sq[0][0] = acc00; Sq[0][1] = acc01; Sq[1][0] = acc10; Sq[1][1] = acc11;
      S[0][0] = acc00 + qn; S[0][1] = acc01; S[1][0] = acc10; S[1][1] = acc11 + qn;
    }
    inv2(S, Si);
    // (e) K_u = P H^T S^-1 (:473), mu += K_u y (:476); Joseph: W = H P - Q K^T, AHt = P H^T - K (S - Q)
    double* ku = sK + u * HIST;
    d2 acur[NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = gl + GW * p;
      d2 kk;
      kk.x = dot2(pht0[p], Si[0][0], pht1[p], Si[1][0]);
      kk.y = dot2(pht0[p], Si[0][1], pht1[p], Si[1][1]);
      hK[u][p] = kk;  // zero for the pad row and beyond
      if (JOSEPH) {
        hW0[u][p] = joseph_w(hW0[u][p], qn, kk.x);
        hW1[u][p] = joseph_w(hW1[u][p], qn, kk.y);
        acur[p].x = joseph_a(pht0[p], kk.x, kk.y, Sq[0][0], Sq[1][0]);
        acur[p].y = joseph_a(pht1[p], kk.x, kk.y, Sq[0][1], Sq[1][1]);
      }
      if (t < LD) {
        *reinterpret_cast<d2*>(ku + 2 * t) = kk;
        if (JOSEPH) *reinterpret_cast<d2*>(sA + u * HIST + 2 * t) = acur[p];
        if (t < N) sMu[t] = mu_inc(sMu[t], kk.x, kk.y, in.y0, in.y1);
      }
    }
    __syncwarp(gmask);
    // (f) bring the robot rows / columns up to date for the next update
    if (u + 1 < m) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const d2 kr = lds128(ku + 2 * k);
        const double wk0 = JOSEPH ? joseph_w(x0[k], qn, kr.x) : x0[k];
        const double wk1 = JOSEPH ? joseph_w(x1[k], qn, kr.y) : x1[k];
        d2 ar;
        if (JOSEPH) ar = lds128(sA + u * HIST + 2 * k);
#pragma unroll
        for (int p = 0; p < NT; ++p) {
          rr[k][p] = rank2(rr[k][p], kr.x, kr.y, hW0[u][p], hW1[u][p]);
          rc[k][p] = rank2(rc[k][p], hK[u][p].x, hK[u][p].y, wk0, wk1);
          if (JOSEPH) {
            rr[k][p] = rank2(rr[k][p], ar.x, ar.y, hK[u][p].x, hK[u][p].y);
            rc[k][p] = rank2(rc[k][p], acur[p].x, acur[p].y, kr.x, kr.y);
          }
        }
      }
    }
  }

  // ---- phase 3: one pass over P, all updates applied in order, streamed to HBM ---------------------------
  // lane = column gl + GW p of every row of ITS filter (W_u of those columns in registers); K_u of a row is one
  // 128-bit load per group. Row chunks are compile-time bounded (immediate offsets).
  {
    constexpr int CH = (NT >= 3) ? 4 : 8;  // rows in registers per chunk
    const bool tail_ok = gl + GW * (NT - 1) < LD;
#pragma unroll
    for (int r0 = 0; r0 < N; r0 += CH) {
      double v[CH][NT];
#pragma unroll
      for (int q = 0; q < CH; ++q)
#pragma unroll
        for (int p = 0; p < NT; ++p)
          if (r0 + q < N && (p < NT - 1 || tail_ok)) v[q][p] = sP[(r0 + q) * LD + gl + GW * p];
#pragma unroll
      for (int u = 0; u < m; ++u) {
        if (!((valid >> u) & 1u)) continue;
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          if (r0 + q < N) {
            const d2 k = lds128(sK + u * HIST + 2 * (r0 + q));
#pragma unroll
            for (int p = 0; p < NT; ++p) v[q][p] = rank2(v[q][p], k.x, k.y, hW0[u][p], hW1[u][p]);
            if (JOSEPH) {
              const d2 av = lds128(sA + u * HIST + 2 * (r0 + q));
#pragma unroll
              for (int p = 0; p < NT; ++p) v[q][p] = rank2(v[q][p], av.x, av.y, hK[u][p].x, hK[u][p].y);
            }
          }
        }
      }
#pragma unroll
      for (int q = 0; q < CH; ++q)
#pragma unroll
        for (int p = 0; p < NT; ++p)
          if (r0 + q < N && (p < NT - 1 || tail_ok)) gP[(r0 + q) * LD + gl + GW * p] = v[q][p];
    }
  }
#pragma unroll
  for (int p = 0; p < NT; ++p) {
    const int t = gl + GW * p;
    if (t < LD) gMu[t] = sMu[t];
  }
  if constexpr (GEN) {
    if (a.gen_pose_out && gl < 3) a.gen_pose_out[b * 3 + gl] = sMu[gl];
  }
}

// -------------------------------------------------------------------------------------------------------
// launchers
// -------------------------------------------------------------------------------------------------------
template <int NLM, int MT, int GW, bool JOSEPH, bool GEN>
static cudaError_t launch_group_g(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  constexpr int fpw = 32 / GW;
  const size_t smem = (size_t)group_slice_doubles(NLM, MT, GW, JOSEPH) * sizeof(double) * fpw;
  auto kernel = batch_step_group_kernel<NLM, MT, GW, JOSEPH, GEN>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  // L2-prefetch distance in CTAs: 3/8 of the CTAs resident on the device (EKF_BATCH_PREFETCH overrides, 0 = off)
  static const int pf = [&] {
    if (const char* v = getenv("EKF_BATCH_PREFETCH")) return atoi(v);
    int dev = 0, sms = 0, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 32, smem) != cudaSuccess)
      return 0;
    return sms * occ * 3 / 8;
  }();
  kernel<<<(unsigned)((batch + fpw - 1) / fpw), 32, smem, st>>>(a, batch, pf);
  return cudaGetLastError();
}

template <int NLM, int MT, int GW, bool JOSEPH>
static cudaError_t launch_group(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  if constexpr (!JOSEPH && GW == 8) {  // generated inputs: plain form, four filters per warp
    if (a.gen_record) return launch_group_g<NLM, MT, 8, false, true>(a, batch, st);
  }
  if (a.gen_record) return cudaErrorInvalidValue;
  return launch_group_g<NLM, MT, GW, JOSEPH, false>(a, batch, st);
}

bool batch_group_supported(int n_landmarks, int m) { return (n_landmarks == 4 || n_landmarks == 8) && m >= 1 && m <= 4; }

template <int NLM, int GW, bool JOSEPH>
static cudaError_t launch_group_m(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  switch (a.m) {
    case 1: return launch_group<NLM, 1, GW, JOSEPH>(a, batch, st);
    case 2: return launch_group<NLM, 2, GW, JOSEPH>(a, batch, st);
    case 3: return launch_group<NLM, 3, GW, JOSEPH>(a, batch, st);
    case 4: return launch_group<NLM, 4, GW, JOSEPH>(a, batch, st);
    default: return cudaErrorInvalidValue;
  }
}

// group width 8 (four filters per warp) or 16 (two per warp); the Joseph form is built at width 8 only
cudaError_t batch_launch_step_group(int n_landmarks, int group_width, const BatchStepArgs& a, int64_t batch, bool joseph,
                                    cudaStream_t st) {
  if (joseph) {
    if (n_landmarks == 4) return launch_group_m<4, 8, true>(a, batch, st);
    if (n_landmarks == 8) return launch_group_m<8, 8, true>(a, batch, st);
    return cudaErrorInvalidValue;
  }
  if (a.gen_record) group_width = 8;  // generated inputs are built at width 8
  if (n_landmarks == 4)
    return group_width == 16 ? launch_group_m<4, 16, false>(a, batch, st) : launch_group_m<4, 8, false>(a, batch, st);
  if (n_landmarks == 8)
    return group_width == 16 ? launch_group_m<8, 16, false>(a, batch, st) : launch_group_m<8, 8, false>(a, batch, st);
  return cudaErrorInvalidValue;
}

}  // namespace ekf
