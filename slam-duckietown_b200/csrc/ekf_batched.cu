// ekf_batched.cu — fused EKF-SLAM filter-step for B independent filters, ONE WARP per filter (sm_100a, fp64).
//
// Reference semantics: one EKF_pose_estimation call (src/replay_no_ros.py:341-480) on an already associated
// window: augment new slots from the PRE-predict pose (:355-360), odometry predict (:368-430), then the
// observations as SEQUENTIAL range-bearing updates in the given order (:436-480).
//
// HBM layout (structure of arrays, one contiguous, 32-byte aligned record per filter and field):
//   mu [B][LD]        LD = n rounded up to 4 doubles, pad zero
//   P  [B][N*LD]      row-major, leading dimension LD, pad column(s) zero
// so a filter's covariance is ONE contiguous block: it is brought into shared memory with a single TMA
// bulk copy (cp.async.bulk + mbarrier) and leaves through coalesced 128-bit stores. Algorithmic traffic per
// filter-step: one read + one write of mu and P regardless of the number of updates.
//
// Why a warp per filter (round-1 ncu, profiles/): the step is a chain of short dependent phases (motion
// model -> [innovation -> S^-1 -> K -> mu] x m). With a CTA per filter 46 % of all stall samples were CTA
// barriers around single-thread sections; a warp needs only __syncwarp, and the ~11 resident warps of an SM
// (20 KB of shared memory each) are independent filters whose chains overlap.
//
// Algorithm (bitwise equal to applying the updates one after the other):
//   phase 0  TMA bulk load of P, mu; per-filter inputs into registers
//   phase 1  augment + motion model (computed warp-uniformly), predict of the robot rows/columns of P in smem
//   phase 2  for each update u: the 5 rows / 5 columns it touches are rebuilt from the staged P0 and the
//            earlier (K_v, W_v), giving H5, W_u = H5 P[idx5,:] (from rows), S = W_u[:,idx5] H5^T + Q, S^-1,
//            K_u (from columns), mu += K_u y_u. O(n) work per update, P0 is not modified.
//   phase 3  one pass over P: P[i,c] = P0[i,c] - sum_u K_u[i,:] W_u[:,c] (plain; rank-4 per update for Joseph),
//            a lane keeping W_u of its 4 columns in registers, streamed to HBM with coalesced 128-bit stores.
// Rows and columns are both gathered for real — (P H^T)^T is never used for H P.
#include "ekf_batched.h"
#include "ekf_batched_dev.cuh"

namespace ekf {

// MT = number of observation slots per filter-step when known at compile time (the update loops unroll and the
// per-update validity tests become predicates), 0 = read it from the arguments.
template <int NLM, bool JOSEPH, int MT, bool GEN>
__global__ void __launch_bounds__(32) batch_step_kernel(BatchStepArgs a) {
  constexpr int N = 3 + 2 * NLM;
  constexpr int LD = (N + 3) & ~3;
  constexpr int G = LD / 4;          // column groups of 4 doubles
  constexpr int HIST = 2 * LD;       // doubles of one update's K (or W, or A) record: [2][LD]
  constexpr int NT = (LD + 31) / 32; // lane t handles indices t, t + 32, ...

  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sP = reinterpret_cast<double*>(smem_raw);  // [N][LD]
  double* sMu = sP + N * LD;                         // [LD]
  double* sK = sMu + LD;                             // [m][LD][2]  (K_u[i,0], K_u[i,1]) pairs
  const int m = MT ? MT : a.m;
  double* sW = sK + m * HIST;                        // [m][2][LD]  W_u[0,:] then W_u[1,:]
  double* sA = sW + m * HIST;                        // [m][LD][2]  AHt_u pairs (Joseph only)
  uint64_t* bar = reinterpret_cast<uint64_t*>(JOSEPH ? sA + m * HIST : sA);

  const int lane = threadIdx.x;
  const int64_t b = blockIdx.x;
  double* __restrict__ gP = a.P + b * (int64_t)(N * LD);
  double* __restrict__ gMu = a.mu + b * (int64_t)LD;

  // ---- phase 0: TMA bulk load + per-filter inputs --------------------------------------------------
  if (lane == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)((N * LD + LD) * sizeof(double)));
    tma_bulk_g2s(sP, gP, (uint32_t)(N * LD * sizeof(double)), bar);
    tma_bulk_g2s(sMu, gMu, (uint32_t)(LD * sizeof(double)), bar);
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
  // observation u lives in lane u and is broadcast with a shuffle when its turn comes
  double o_rng = 0.0, o_brg = 0.0;
  int o_slot = -1;
  if constexpr (GEN) {  // this filter's Monte Carlo inputs, drawn here while its state is still in flight
    double z0 = 0.0, z1 = 0.0;
    if (lane <= m)
      normal_pair(a.gen_seed, a.gen_first + (unsigned long long)b, a.gen_step,
                  lane < m ? kStreamObservation0 + lane : kStreamOdometry, &z0, &z1);
    if (lane < m) {
      o_slot = reinterpret_cast<const int32_t*>(a.gen_record + 2 + 2 * m)[lane];
      o_rng = __dadd_rn(a.gen_record[2 + lane], __dmul_rn(a.gen_s_rng, z0));
      o_brg = __dadd_rn(a.gen_record[2 + m + lane], __dmul_rn(a.gen_s_brg, z1));
    }
    ang = __shfl_sync(kFull, __dadd_rn(a.gen_record[0], __dmul_rn(a.gen_s_ang, z0)), m);
    lin = __shfl_sync(kFull, __dadd_rn(a.gen_record[1], __dmul_rn(a.gen_s_lin, z1)), m);
  } else if (lane < m) {
    o_slot = a.slot[b * m + lane];
    o_rng = a.rng[b * m + lane];
    o_brg = a.brg[b * m + lane];
  }
  __syncwarp();
  mbar_wait(bar, 0);

  // ---- phase 1: augment (:355-360) + motion model (:368-417), warp-uniform ---------------------------
  unsigned valid = 0;  // bit u: update u is applied
  {
    const double px = sMu[0], py = sMu[1], pth = sMu[2];
    // even lanes evaluate sincos(theta), odd lanes sincos(theta + ang): one call instead of two
    double sn, cs;
    fm_sincos((lane & 1) ? pth + ang : pth, &sn, &cs);
    const double s1 = __shfl_sync(kFull, sn, 0), c1 = __shfl_sync(kFull, cs, 0);
    const double s2 = __shfl_sync(kFull, sn, 1), c2 = __shfl_sync(kFull, cs, 1);
    int fault = 0;
    for (int u = 0; u < m; ++u) {
      const int s = __shfl_sync(kFull, o_slot, u);
      if (s < 0) continue;
      if (s >= NLM || s > nk) {  // beyond capacity, or a slot that skips first-seen order (:359 KeyError)
        fault = 1;
        continue;
      }
      valid |= 1u << u;
      if (s == nk) {  // first sighting: initialise from the PRE-predict pose (:331-332, :359-360)
        const int j = 3 + 2 * s;
        const double r = __shfl_sync(kFull, o_rng, u), be = __shfl_sync(kFull, o_brg, u);
        double sb, cb;
        fm_sincos(be + pth, &sb, &cb);
        if (lane == 0) {
          sMu[j] = __dadd_rn(px, __dmul_rn(r, cb));
          sMu[j + 1] = __dadd_rn(py, __dmul_rn(r, sb));
          sP[j * LD + j] = kNewLandmarkVar;
          sP[(j + 1) * LD + j + 1] = kNewLandmarkVar;
        }
        ++nk;
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
    // predict covariance in shared memory (:420-430): rows 0-1 += g * row 2 over ALL columns, then columns
    // 0-1 += g * column 2 over ALL rows (using the updated rows 0-1), then the three noise terms.
    const double g0 = mo.g0, g1 = mo.g1;
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      if (t < N) {
        const double p2 = sP[2 * LD + t];
        sP[0 * LD + t] = fma(g0, p2, sP[0 * LD + t]);
        sP[1 * LD + t] = fma(g1, p2, sP[1 * LD + t]);
      }
    }
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int i = lane + 32 * p;
      if (i < N) {
        double* row = sP + i * LD;
        const double p2 = row[2];
        double r0 = fma(g0, p2, row[0]), r1 = fma(g1, p2, row[1]);
        if (i == 0) r0 = __dadd_rn(r0, __dmul_rn(mv, mv));           // F^T R F (:421-430)
        if (i == 1) r1 = __dadd_rn(r1, __dmul_rn(mv, mv));
        row[0] = r0;
        row[1] = r1;
        if (i == 2) row[2] = __dadd_rn(p2, __dmul_rn(mv / 2.0, mv / 2.0));
      }
    }
    __syncwarp();
  }

  // ---- phase 2: sequential updates on the touched rows / columns only --------------------------------
  // Lane t owns column t of the gathered ROWS (-> W_u[:, t]) and row t of the gathered COLUMNS (-> K_u[t, :])
  // for t = lane and t = lane + 32.
  for (int u = 0; u < m; ++u) {
    if (!((valid >> u) & 1u)) continue;  // warp-uniform
    const int s = __shfl_sync(kFull, o_slot, u);
    const double z_rng = __shfl_sync(kFull, o_rng, u), z_brg = __shfl_sync(kFull, o_brg, u);
    const int j = 3 + 2 * s;
    // (a) innovation and H5 (:443-469), warp-uniform
    const Innov in = innovation(sMu[0], sMu[1], sMu[2], sMu[j], sMu[j + 1], z_rng, z_brg);
    // (b) gather the 5 rows and 5 columns of the staged P0 ...
    double vr[NT][5], vc[NT][5];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      const bool inld = (t < LD), real = (t < N);
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int rk = (k < 3) ? k : j + (k - 3);
        vr[p][k] = inld ? sP[rk * LD + t] : 0.0;   // row rk, column t (pad column reads the zero pad)
        vc[p][k] = real ? sP[t * LD + rk] : 0.0;   // row t, column rk (pad row: K stays zero)
      }
    }
    // ... and bring them up to date with the earlier updates of this step (rank-2 / rank-4 each)
    for (int w = 0; w < u; ++w) {
      if (!((valid >> w) & 1u)) continue;
      const double* kw = sK + w * HIST;   // [LD][2]
      const double* ww = sW + w * HIST;   // [2][LD]
      const double* aw = sA + w * HIST;   // [LD][2]
      d2 kr[5], ar[5];                    // warp-uniform: entries idx5 of K_w, A_w, W_w
      double wr0[5], wr1[5];
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int rk = (k < 3) ? k : j + (k - 3);
        kr[k] = lds128(kw + 2 * rk);
        wr0[k] = ww[rk];
        wr1[k] = ww[LD + rk];
        if (JOSEPH) ar[k] = lds128(aw + 2 * rk);
      }
#pragma unroll
      for (int p = 0; p < NT; ++p) {
        const int t = lane + 32 * p;
        if (t < LD) {
          const double wt0 = ww[t], wt1 = ww[LD + t];
          const d2 kt = lds128(kw + 2 * t);
          d2 at = {0.0, 0.0};
          if (JOSEPH) at = lds128(aw + 2 * t);
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            vr[p][k] = rank2(vr[p][k], kr[k].x, kr[k].y, wt0, wt1);
            vc[p][k] = rank2(vc[p][k], kt.x, kt.y, wr0[k], wr1[k]);
            if (JOSEPH) {
              vr[p][k] = rank2(vr[p][k], ar[k].x, ar[k].y, kt.x, kt.y);
              vc[p][k] = rank2(vc[p][k], at.x, at.y, kr[k].x, kr[k].y);
            }
          }
        }
      }
    }
    // (c) H P (rows) and P H^T (columns) for this lane's t
    double hp0[NT], hp1[NT], pht0[NT], pht1[NT];
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
      hp0[p] = a0; hp1[p] = a1; pht0[p] = b0; pht1[p] = b1;
      const int t = lane + 32 * p;
      if (t < LD) {
        wu[t] = a0;
        wu[LD + t] = a1;
      }
    }
    __syncwarp();
    // (d) S = (H P) H^T + Q from the five entries idx5 of H P (:473), closed-form inverse; warp-uniform
    double S[2][2], Si[2][2], Sq[2][2];
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
      Sq[0][0] = acc00; Sq[0][1] = acc01; Sq[1][0] = acc10; Sq[1][1] = acc11;
      S[0][0] = acc00 + qn; S[0][1] = acc01; S[1][0] = acc10; S[1][1] = acc11 + qn;
    }
    inv2(S, Si);
    if (JOSEPH) __syncwarp();  // every lane has read the raw H P before W is overwritten below
    // (e) K_u = P H^T S^-1 (:473), mu += K_u y (:476); Joseph: W = H P - Q K^T, AHt = P H^T - K (S - Q)
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      if (t < LD) {
        d2 kk;
        kk.x = dot2(pht0[p], Si[0][0], pht1[p], Si[1][0]);
        kk.y = dot2(pht0[p], Si[0][1], pht1[p], Si[1][1]);
        *reinterpret_cast<d2*>(ku + 2 * t) = kk;  // zero for the pad row
        if (JOSEPH) {
          wu[t] = joseph_w(hp0[p], qn, kk.x);
          wu[LD + t] = joseph_w(hp1[p], qn, kk.y);
          d2 av;
          av.x = joseph_a(pht0[p], kk.x, kk.y, Sq[0][0], Sq[1][0]);
          av.y = joseph_a(pht1[p], kk.x, kk.y, Sq[0][1], Sq[1][1]);
          *reinterpret_cast<d2*>(sA + u * HIST + 2 * t) = av;
        }
        if (t < N) sMu[t] = mu_inc(sMu[t], kk.x, kk.y, in.y0, in.y1);
      }
    }
    __syncwarp();
  }

  // ---- phase 3: one pass over P, all updates applied in order, streamed to HBM -------------------------
  // Lane (ph, g): columns {2g, 2g+1} and {LD/2 + 2g, LD/2 + 2g+1} (W_u of those columns in registers) and rows
  // ph, ph + RP, ...; consecutive lanes are 16 bytes apart, which is what makes the 128-bit shared-memory loads
  // conflict-free (round-1 ncu: 256-bit accesses are issued as two LDS.128 with lanes 32 bytes apart = 2-way
  // conflicts, and a (row block, column group) mapping spent 1100 wavefronts per filter on redundant K/W loads).
  // The RP rows of one iteration are adjacent in memory. K_u of a row is one 128-bit load shared by its G lanes.
  {
    constexpr int RP = 32 / G;              // rows per iteration
    constexpr int NIT = (N + RP - 1) / RP;  // iterations
    constexpr int CH = 8;                   // rows of P in registers per lane between W reloads
    constexpr int HALF = LD / 2;
    const int ph = lane / G, g = lane - ph * G;
    if (ph < RP) {
      const int ca = 2 * g, cb = HALF + 2 * g;
#pragma unroll
      for (int it0 = 0; it0 < NIT; it0 += CH) {
        d2 pa[CH], pb[CH];
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          const int r = (it0 + q) * RP + ph;
          if (it0 + q < NIT && r < N) {
            pa[q] = lds128(sP + r * LD + ca);
            pb[q] = lds128(sP + r * LD + cb);
          }
        }
        for (int u = 0; u < m; ++u) {
          if (!((valid >> u) & 1u)) continue;
          const double* ku = sK + u * HIST;
          const double* wu = sW + u * HIST;
          const d2 w0a = lds128(wu + ca), w0b = lds128(wu + cb);
          const d2 w1a = lds128(wu + LD + ca), w1b = lds128(wu + LD + cb);
          d2 kc[4];  // Joseph: K_u pairs of this lane's 4 columns
          if (JOSEPH) {
            kc[0] = lds128(ku + 2 * ca);
            kc[1] = lds128(ku + 2 * ca + 2);
            kc[2] = lds128(ku + 2 * cb);
            kc[3] = lds128(ku + 2 * cb + 2);
          }
#pragma unroll
          for (int q = 0; q < CH; ++q) {
            const int r = (it0 + q) * RP + ph;
            if (it0 + q < NIT && r < N) {
              const d2 k = lds128(ku + 2 * r);
              pa[q].x = rank2(pa[q].x, k.x, k.y, w0a.x, w1a.x);
              pa[q].y = rank2(pa[q].y, k.x, k.y, w0a.y, w1a.y);
              pb[q].x = rank2(pb[q].x, k.x, k.y, w0b.x, w1b.x);
              pb[q].y = rank2(pb[q].y, k.x, k.y, w0b.y, w1b.y);
              if (JOSEPH) {
                const d2 av = lds128(sA + u * HIST + 2 * r);
                pa[q].x = rank2(pa[q].x, av.x, av.y, kc[0].x, kc[0].y);
                pa[q].y = rank2(pa[q].y, av.x, av.y, kc[1].x, kc[1].y);
                pb[q].x = rank2(pb[q].x, av.x, av.y, kc[2].x, kc[2].y);
                pb[q].y = rank2(pb[q].y, av.x, av.y, kc[3].x, kc[3].y);
              }
            }
          }
        }
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          const int r = (it0 + q) * RP + ph;
          if (it0 + q < NIT && r < N) {
            stg128(gP + r * LD + ca, pa[q]);
            stg128(gP + r * LD + cb, pb[q]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int p = 0; p < NT; ++p) {
    const int t = lane + 32 * p;
    if (t < LD) gMu[t] = sMu[t];
  }
  if constexpr (GEN) {
    if (a.gen_pose_out && lane < 3) a.gen_pose_out[b * 3 + lane] = sMu[lane];
  }
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
  size_t d = (size_t)n * ld + ld + (size_t)m * 2 * ld * (joseph ? 3 : 2) + kTailDoubles;
  return d * sizeof(double);
}

template <int NLM, bool JOSEPH, int MT, bool GEN>
static cudaError_t launch_step_g(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  const size_t smem = batch_step_smem_bytes(NLM, a.m, JOSEPH);
  auto kernel = batch_step_kernel<NLM, JOSEPH, MT, GEN>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  // as many one-warp CTAs per SM as shared memory allows (11 at n = 43, m = 3)
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  kernel<<<(unsigned)batch, 32, smem, st>>>(a);
  return cudaGetLastError();
}

// generated inputs (BatchStepArgs::gen_record) run on the general-m instantiation of this staged kernel (both forms)
template <int NLM, bool JOSEPH, int MT>
static cudaError_t launch_step_mt(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  if (a.gen_record) return launch_step_g<NLM, JOSEPH, 0, true>(a, batch, st);
  return launch_step_g<NLM, JOSEPH, MT, false>(a, batch, st);
}

template <int NLM, bool JOSEPH>
static cudaError_t launch_step_t(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  switch (a.m) {  // the common observation counts get unrolled update loops
    case 2: return launch_step_mt<NLM, JOSEPH, 2>(a, batch, st);
    case 3: return launch_step_mt<NLM, JOSEPH, 3>(a, batch, st);
    default: return launch_step_mt<NLM, JOSEPH, 0>(a, batch, st);
  }
}

bool batch_capacity_supported(int n_landmarks) {
  switch (n_landmarks) {
    case 4: case 8: case 12: case 16: case 20: case 24: case 28: return true;
    case 32: case 40: case 48: case 56: return true;  // staged kernel only (general m); beyond: LargeMapEKF per filter
    default: return false;
  }
}

#define EKF_STEP_CASE(L) \
  case L: return joseph ? launch_step_t<L, true>(a, batch, st) : launch_step_t<L, false>(a, batch, st)
// capacities above 28 landmarks: the general-m instantiation only (the map of the reference grows without bound,
// replay_no_ros.py:341-352; one warp still stages the whole covariance: 106 KB at 56 landmarks)
#define EKF_BIG_CASE(L) \
  case L: return joseph ? launch_step_mt<L, true, 0>(a, batch, st) : launch_step_mt<L, false, 0>(a, batch, st)
cudaError_t batch_launch_step(int n_landmarks, const BatchStepArgs& a, int64_t batch, bool joseph,
                              cudaStream_t st) {
  switch (n_landmarks) {
    EKF_STEP_CASE(4);
    EKF_STEP_CASE(8);
    EKF_STEP_CASE(12);
    EKF_STEP_CASE(16);
    EKF_STEP_CASE(20);
    EKF_STEP_CASE(24);
    EKF_STEP_CASE(28);
    EKF_BIG_CASE(32);
    EKF_BIG_CASE(40);
    EKF_BIG_CASE(48);
    EKF_BIG_CASE(56);
    default: return cudaErrorInvalidValue;
  }
}
#undef EKF_STEP_CASE
#undef EKF_BIG_CASE

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
