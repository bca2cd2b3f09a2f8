// ekf_batched_regw.cu — fused EKF-SLAM filter-step, one warp per filter, W-IN-REGISTERS variant (sm_100a, fp64).
//
// Same reference semantics (src/replay_no_ros.py:341-480) and the same per-element operation sequence as
// batch_step_kernel (ekf_batched.cu), so results are bitwise equal; what changes is where the step's small
// operands live. Round-1 ncu of the staged kernel (profiles/r1f_batch.*): shared-memory pipe 76 % busy, issue
// slots 41 %, 11 resident warps per SM limited by 19.7 KB of shared memory per filter. Here
//   * lane t owns COLUMN t (and t + 32) of P in every phase, so W_u[:, t] = (H P)[:, t] and K_u[t, :] are
//     produced in the registers of the lane that consumes them: the W history never touches shared memory
//     (2.1 KB less per filter -> 12 resident warps), only K_u is stored (phase 3 broadcasts it per row);
//   * the robot rows/columns {0,1,2}, which every update touches, stay in registers across the predict and
//     all updates and are brought up to date ONCE per update instead of being re-gathered (8-way bank
//     conflicted column reads) and rebuilt from P0 every time;
//   * the uniform operands an update needs from other lanes (W_w[:, j], the five S entries) travel by
//     warp shuffle;
//   * phase 3 packs the 12-column tail of TWO rows into one instruction (lanes 0-11 / 12-23).
// Plain covariance form with a compile-time observation count only; everything else runs on batch_step_kernel.
#include "ekf_batched.h"
#include "ekf_batched_dev.cuh"

#include <cstdlib>

namespace ekf {

namespace {

// value of column c (warp-uniform) from the lane that owns it: x[0] covers columns 0-31, x[1] columns 32-63
template <int NT>
__device__ __forceinline__ double col_bcast(const double (&x)[NT], int c) {
  // c is warp-uniform: every lane offers its value of the half that holds column c, lane c % 32 is read
  const double mine = (NT > 1 && c >= 32) ? x[NT > 1 ? 1 : 0] : x[0];
  return __shfl_sync(kFull, mine, c & 31);
}

// Phase-3 worker: NPQ row pairs starting at row r0 (sp = sP + r0 LD, sk = sK + 2 r0, gp = gP + r0 LD). Every
// bound is a compile-time constant, so loads/stores carry immediate offsets and no predicates besides the tail's.
// Main part: lane = column (W of it in mw0/mw1); tail: columns 32.. of row (2q + half), W of them in tw0/tw1.
// ODD: the last pair has only its first row. PACK: both rows' tails in one instruction (lanes [0,T) / [T,2T)).
template <int LD, int HIST, int M, int NPQ, bool ODD, bool PACK, bool JOSEPH>
__device__ __forceinline__ void regw_stream_pairs(const double* __restrict__ sp, const double* __restrict__ sk,
                                                  const double* __restrict__ sa, double* __restrict__ gp,
                                                  const double (&mw0)[M], const double (&mw1)[M],
                                                  const double (&tw0)[M], const double (&tw1)[M], const d2 (&mk)[M],
                                                  const d2 (&tk)[M], unsigned valid, int lane, int half, int tcol,
                                                  bool tact) {
  double va[NPQ], vb[NPQ], vt[NPQ], vs[PACK ? 1 : NPQ];
  const int toff = half * LD + tcol;  // tail element of this lane within a row pair
#pragma unroll
  for (int q = 0; q < NPQ; ++q) {
    const bool two = !(ODD && q == NPQ - 1);
    va[q] = sp[2 * q * LD + lane];
    if (two) vb[q] = sp[(2 * q + 1) * LD + lane];
    if (tact && (two || !half)) vt[q] = sp[2 * q * LD + toff];
    if (!PACK && two && tact) vs[q] = sp[(2 * q + 1) * LD + tcol];
  }
#pragma unroll
  for (int u = 0; u < M; ++u) {
    if (!((valid >> u) & 1u)) continue;
    const double* ku = sk + u * HIST;
    const double* au = sa + u * HIST;
#pragma unroll
    for (int q = 0; q < NPQ; ++q) {
      const bool two = !(ODD && q == NPQ - 1);
      const d2 ka = lds128(ku + 4 * q);
      const d2 kt = PACK ? lds128(ku + 4 * q + 2 * half) : ka;  // K of the row this lane's tail element is in
      va[q] = rank2(va[q], ka.x, ka.y, mw0[u], mw1[u]);
      vt[q] = rank2(vt[q], kt.x, kt.y, tw0[u], tw1[u]);
      if (JOSEPH) {  // second rank-2 term of the one-pass Joseph form: AHt of the row, K of the column
        const d2 aa = lds128(au + 4 * q);
        const d2 at = PACK ? lds128(au + 4 * q + 2 * half) : aa;
        va[q] = rank2(va[q], aa.x, aa.y, mk[u].x, mk[u].y);
        vt[q] = rank2(vt[q], at.x, at.y, tk[u].x, tk[u].y);
      }
      if (two) {
        const d2 kb = lds128(ku + 4 * q + 2);
        vb[q] = rank2(vb[q], kb.x, kb.y, mw0[u], mw1[u]);
        if (!PACK) vs[q] = rank2(vs[q], kb.x, kb.y, tw0[u], tw1[u]);
        if (JOSEPH) {
          const d2 ab = lds128(au + 4 * q + 2);
          vb[q] = rank2(vb[q], ab.x, ab.y, mk[u].x, mk[u].y);
          if (!PACK) vs[q] = rank2(vs[q], ab.x, ab.y, tk[u].x, tk[u].y);
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < NPQ; ++q) {
    const bool two = !(ODD && q == NPQ - 1);
    gp[2 * q * LD + lane] = va[q];
    if (two) gp[(2 * q + 1) * LD + lane] = vb[q];
    if (tact && (two || !half)) gp[2 * q * LD + toff] = vt[q];
    if (!PACK && two && tact) gp[(2 * q + 1) * LD + tcol] = vs[q];
  }
}

}  // namespace

// WPC = warps per CTA; only WPC == 1 is built: one filter per one-warp CTA, the hardware CTA scheduler balances the
// load. (The WPC > 1 code path — ONE persistent CTA per SM whose warps pull filters from a device-side work counter,
// which pays the 1 KB system reservation of shared memory once per SM and so fits a 13th filter at n = 43 — is kept in
// the kernel body but measured slower, see the launcher.)
#ifndef EKF_REGW_MAX_BLOCKS
#define EKF_REGW_MAX_BLOCKS 24
#endif
// resident one-warp CTAs per SM the register allocation should allow: what shared memory admits, at most
// EKF_REGW_MAX_BLOCKS (small filters are otherwise register-limited to 16 warps: ncu, profiles/r1h_batchM.*)
__host__ __device__ constexpr int regw_min_blocks(int nlm, int m, bool joseph) {
  const int n = 3 + 2 * nlm, ld = (n + 3) & ~3;
  const int fit = 233472 / ((n * ld + ld * (1 + 2 * m * (joseph ? 2 : 1)) + 2) * 8 + 1024);
  // m > 4 (the decoded-bag-shaped steps with up to 6 updates): the W history alone is 4 m registers per lane, so the
  // register allocation is given room (<= 10 resident warps) instead of being squeezed into spills
  // capacities >= 12 landmarks at m <= 4: at most 16 resident warps, i.e. 128 registers (17+ forced spills)
  const int cap = (m > 4) ? 10 : (nlm >= 12 ? 16 : EKF_REGW_MAX_BLOCKS);
  return fit < cap ? fit : cap;
}

// JOSEPH: the covariance update is the one-pass rank-4 Joseph form — per update a second K-like
// record AHt_u lives in shared memory, W_u becomes H P - Q K^T, and every element takes a second rank-2 term
// (AHt of its row, K of its column), in the order of the staged kernel: bitwise equal to it.
// GEN: the filter's Monte Carlo inputs are drawn inside the kernel (BatchStepArgs::gen_record); a separate instantiation,
// so that the kernels fed from arrays stay instruction for instruction what they were
template <int NLM, int MT, int WPC, bool JOSEPH, bool GEN>
__global__ void __launch_bounds__(32 * WPC, regw_min_blocks(NLM, MT, JOSEPH)) batch_step_regw_kernel(BatchStepArgs a, int64_t batch, int pf) {
  constexpr int N = 3 + 2 * NLM;
  constexpr int LD = (N + 3) & ~3;
  constexpr int HIST = 2 * LD;        // doubles of one update's K record: [LD][2]
  constexpr int NT = (LD + 31) / 32;  // lane t handles indices t, t + 32
  constexpr int m = MT;
  constexpr int SLICE = N * LD + LD + m * HIST * (JOSEPH ? 2 : 1) + 2;  // doubles of shared memory per warp
  // K history of this lane's rows: in registers for m <= 4; for longer steps it is re-read from the K records in
  // shared memory (one LDS.128 per earlier update and half), which frees 4 m registers per lane
  constexpr bool KREG = (MT <= 4);
  static_assert(NT <= 2, "lane owns at most two columns");
  static_assert(MT >= 1 && MT <= BATCH_MAX_M, "observation count");
  static_assert(SLICE % 2 == 0, "16-byte aligned slices");

  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  double* sP = reinterpret_cast<double*>(smem_raw) + warp * SLICE;  // [N][LD]
  double* sMu = sP + N * LD;                                        // [LD]
  double* sK = sMu + LD;                                            // [m][LD][2]  (K_u[i,0], K_u[i,1]) pairs
  double* sA = sK + m * HIST;                                       // [m][LD][2]  AHt_u pairs (Joseph only)
  uint64_t* bar = reinterpret_cast<uint64_t*>(sA + (JOSEPH ? m * HIST : 0));  // [2]: mu, P

  if (lane == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const int64_t total_warps = (int64_t)gridDim.x * WPC;
  unsigned long long* work = a.faults + 1;  // [1] filters handed out beyond the static first round, [2] warps done
  uint32_t parity = 0;
  for (int64_t b = (int64_t)blockIdx.x * WPC + warp; b < batch; parity ^= 1u) {
  double* __restrict__ gP = a.P + b * (int64_t)(N * LD);
  double* __restrict__ gMu = a.mu + b * (int64_t)LD;

  // ---- phase 0: TMA bulk load + per-filter inputs --------------------------------------------------
  // mu arrives on its own mbarrier: the scalar part of phase 1 (sincos, augment, motion model) runs while the
  // 15 KB of P are still in flight
  int64_t b_next = batch;
  if (lane == 0) {
    if (WPC > 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic writes vs TMA
    mbar_expect_tx(bar, (uint32_t)(LD * sizeof(double)));
    tma_bulk_g2s(sMu, gMu, (uint32_t)(LD * sizeof(double)), bar);
    mbar_expect_tx(bar + 1, (uint32_t)(N * LD * sizeof(double)));
    tma_bulk_g2s(sP, gP, (uint32_t)(N * LD * sizeof(double)), bar + 1);
    if (WPC > 1) b_next = total_warps + (int64_t)atomicAdd(work, 1ULL);  // latency hidden behind this filter
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
  double o_rng = 0.0, o_brg = 0.0;  // observation u lives in lane u
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
  mbar_wait(bar, parity);

  // robot rows / columns of the CURRENT covariance, lane t: rr[k][p] = P[k][t], rc[k][p] = P[t][k], t = lane + 32 p
  double rr[3][NT], rc[3][NT];

  // ---- phase 1: augment (:355-360) + motion model (:368-417), warp-uniform ---------------------------
  unsigned valid = 0;   // bit u: update u is applied
  unsigned fresh = 0;   // bit u: update u is the first sighting of its landmark
  int sl[m];            // slot of observation u (warp-uniform)
  {
    const double px = sMu[0], py = sMu[1], pth = sMu[2];
    double sn, cs;  // even lanes sincos(theta), odd lanes sincos(theta + ang): one call instead of two
    fm_sincos((lane & 1) ? pth + ang : pth, &sn, &cs);
    const double s1 = __shfl_sync(kFull, sn, 0), c1 = __shfl_sync(kFull, cs, 0);
    const double s2 = __shfl_sync(kFull, sn, 1), c2 = __shfl_sync(kFull, cs, 1);
    int fault = 0;
#pragma unroll
    for (int u = 0; u < m; ++u) {
      const int s = __shfl_sync(kFull, o_slot, u);
      sl[u] = s;
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
        }
        fresh |= 1u << u;
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
    mbar_wait(bar + 1, parity);
    if (fresh && lane == 0) {  // new landmarks' diagonals (:356-357)
#pragma unroll
      for (int u = 0; u < m; ++u)
        if ((fresh >> u) & 1u) {
          const int j = 3 + 2 * sl[u];
          sP[j * LD + j] = kNewLandmarkVar;
          sP[(j + 1) * LD + j + 1] = kNewLandmarkVar;
        }
    }
    __syncwarp();
    // predict covariance (:420-430): rows 0-1 += g * row 2 over ALL columns, then columns 0-1 += g * column 2
    // over ALL rows (using the updated rows 0-1), then the three noise terms. Shared memory keeps the result
    // (phase 3 reads it); the robot rows/columns also stay in registers.
    const double g0 = mo.g0, g1 = mo.g1;
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
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
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int i = lane + 32 * p;
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
    __syncwarp();
    if (lane < 3) {  // the 3x3 robot block was changed by the column pass: refresh the row registers
      rr[0][0] = sP[0 * LD + lane];
      rr[1][0] = sP[1 * LD + lane];
      rr[2][0] = sP[2 * LD + lane];
    }
  }

  // ---- phase 2: sequential updates on the touched rows / columns only --------------------------------
  // history in registers: hW[u][.][p] = W_u[:, t], hK[u][p] = K_u[t, :] of this lane's t = lane + 32 p
  double hW0[m][NT], hW1[m][NT];
  d2 hK[KREG ? m : 1][NT];
  auto k_hist = [&](int w, int p) -> d2 {
    if constexpr (KREG) {
      return hK[w][p];
    } else {
      const int t = lane + 32 * p;
      d2 z;
      z.x = 0.0; z.y = 0.0;
      return (t < LD) ? lds128(sK + w * HIST + 2 * t) : z;
    }
  };
  auto a_hist = [&](int w, int p) -> d2 {  // AHt_w of this lane's row t (Joseph)
    const int t = lane + 32 * p;
    d2 z;
    z.x = 0.0; z.y = 0.0;
    return (t < LD) ? lds128(sA + w * HIST + 2 * t) : z;
  };
#pragma unroll
  for (int u = 0; u < m; ++u) {
    if (!((valid >> u) & 1u)) continue;  // warp-uniform
    const double z_rng = __shfl_sync(kFull, o_rng, u), z_brg = __shfl_sync(kFull, o_brg, u);
    const int j = 3 + 2 * sl[u];
    // (a) innovation and H5 (:443-469), warp-uniform
    const Innov in = innovation(sMu[0], sMu[1], sMu[2], sMu[j], sMu[j + 1], z_rng, z_brg);
    // (b) landmark rows / columns j, j+1 from the staged P0, brought up to date with the earlier updates
    double lr[2][NT], lc[2][NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
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
      const double w0j0 = col_bcast<NT>(hW0[w], j), w1j0 = col_bcast<NT>(hW1[w], j);
      const double w0j1 = col_bcast<NT>(hW0[w], j + 1), w1j1 = col_bcast<NT>(hW1[w], j + 1);
      d2 aj0, aj1;
      if (JOSEPH) {
        aj0 = lds128(sA + w * HIST + 2 * j);
        aj1 = lds128(sA + w * HIST + 2 * j + 2);
      }
#pragma unroll
      for (int p = 0; p < NT; ++p) {
        const d2 kw = k_hist(w, p);
        lr[0][p] = rank2(lr[0][p], kj0.x, kj0.y, hW0[w][p], hW1[w][p]);
        lr[1][p] = rank2(lr[1][p], kj1.x, kj1.y, hW0[w][p], hW1[w][p]);
        lc[0][p] = rank2(lc[0][p], kw.x, kw.y, w0j0, w1j0);
        lc[1][p] = rank2(lc[1][p], kw.x, kw.y, w0j1, w1j1);
        if (JOSEPH) {
          const d2 aw = a_hist(w, p);
          lr[0][p] = rank2(lr[0][p], aj0.x, aj0.y, kw.x, kw.y);
          lr[1][p] = rank2(lr[1][p], aj1.x, aj1.y, kw.x, kw.y);
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
    // (d) S = (H P) H^T + Q from the five entries idx5 of H P (:473), closed-form inverse; warp-uniform
    double x0[5], x1[5];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      x0[c] = __shfl_sync(kFull, hW0[u][0], c);
      x1[c] = __shfl_sync(kFull, hW1[u][0], c);
    }
    x0[3] = col_bcast<NT>(hW0[u], j);     x1[3] = col_bcast<NT>(hW1[u], j);
    x0[4] = col_bcast<NT>(hW0[u], j + 1); x1[4] = col_bcast<NT>(hW1[u], j + 1);
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
    d2 kcur[NT], acur[NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) {
      const int t = lane + 32 * p;
      d2 kk;
      kk.x = dot2(pht0[p], Si[0][0], pht1[p], Si[1][0]);
      kk.y = dot2(pht0[p], Si[0][1], pht1[p], Si[1][1]);
      if constexpr (KREG) hK[u][p] = kk;  // zero for the pad row and beyond
      kcur[p] = kk;
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
    __syncwarp();
    // (f) bring the robot rows / columns up to date for the next update
    if (u + 1 < m) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const d2 kr = lds128(ku + 2 * k);
        // W_u at column k as the lane owning that column holds it (Joseph: H P - Q K^T, same fma as there)
        const double wk0 = JOSEPH ? joseph_w(x0[k], qn, kr.x) : x0[k];
        const double wk1 = JOSEPH ? joseph_w(x1[k], qn, kr.y) : x1[k];
        d2 ar;
        if (JOSEPH) ar = lds128(sA + u * HIST + 2 * k);
#pragma unroll
        for (int p = 0; p < NT; ++p) {
          rr[k][p] = rank2(rr[k][p], kr.x, kr.y, hW0[u][p], hW1[u][p]);
          rc[k][p] = rank2(rc[k][p], kcur[p].x, kcur[p].y, wk0, wk1);
          if (JOSEPH) {
            rr[k][p] = rank2(rr[k][p], ar.x, ar.y, kcur[p].x, kcur[p].y);
            rc[k][p] = rank2(rc[k][p], acur[p].x, acur[p].y, kr.x, kr.y);
          }
        }
      }
    }
  }

  // ---- phase 3: one pass over P, all updates applied in order, streamed to HBM -------------------------
  // Lane t: column t of every row (W_u[:, t] in registers); the tail columns 32..LD-1 of rows r and r+1 are
  // handled together by lanes [0, TAIL) and [TAIL, 2 TAIL). K_u of a row is one broadcast 128-bit load.
  if constexpr (NT == 2) {
    constexpr int TAIL = LD - 32;
    constexpr bool PACK = (2 * TAIL <= 32);             // both tails of a row pair in one instruction
    const int half = (PACK && lane >= TAIL) ? 1 : 0;   // which row of the pair this lane's tail element is in
    const int tcol = 32 + lane - half * TAIL;          // its column
    const bool tact = lane < (PACK ? 2 * TAIL : TAIL);
    double tW0[m], tW1[m];
#pragma unroll
    for (int u = 0; u < m; ++u) {
      tW0[u] = __shfl_sync(kFull, hW0[u][1], lane - half * TAIL);
      tW1[u] = __shfl_sync(kFull, hW1[u][1], lane - half * TAIL);
    }
    double mW0[m], mW1[m];
    d2 mK[m], tK[m];  // Joseph: K_u of this lane's main and tail column
#pragma unroll
    for (int u = 0; u < m; ++u) {
      mW0[u] = hW0[u][0];
      mW1[u] = hW1[u][0];
      mK[u].x = mK[u].y = tK[u].x = tK[u].y = 0.0;
      if (JOSEPH && ((valid >> u) & 1u)) {
        mK[u] = lds128(sK + u * HIST + 2 * lane);
        if (tact) tK[u] = lds128(sK + u * HIST + 2 * tcol);
      }
    }
    constexpr int CH = 4;  // row PAIRS in registers per lane
    constexpr int NPAIR = (N + 1) / 2;  // N is odd: the last pair holds one row
    constexpr int REM = (NPAIR % CH == 0) ? CH : NPAIR % CH;
    constexpr int MAINCH = (NPAIR - REM) / CH;
    const double* sp = sP;
    const double* sk = sK;
    const double* sa = sA;
    double* gp = gP;
#pragma unroll 1
    for (int c = 0; c < MAINCH; ++c) {
      regw_stream_pairs<LD, HIST, m, CH, false, PACK, JOSEPH>(sp, sk, sa, gp, mW0, mW1, tW0, tW1, mK, tK, valid, lane,
                                                              half, tcol, tact);
      sp += 2 * CH * LD;
      sk += 4 * CH;
      sa += 4 * CH;
      gp += 2 * CH * LD;
    }
    regw_stream_pairs<LD, HIST, m, REM, true, PACK, JOSEPH>(sp, sk, sa, gp, mW0, mW1, tW0, tW1, mK, tK, valid, lane, half,
                                                            tcol, tact);
  } else {
    constexpr int CH = 8;
#pragma unroll 1
    for (int r0 = 0; r0 < N; r0 += CH) {
      double v[CH];
#pragma unroll
      for (int q = 0; q < CH; ++q)
        if (r0 + q < N && lane < LD) v[q] = sP[(r0 + q) * LD + lane];
#pragma unroll
      for (int u = 0; u < m; ++u) {
        if (!((valid >> u) & 1u)) continue;
        const double* ku = sK + u * HIST;
#pragma unroll
        for (int q = 0; q < CH; ++q) {
          if (r0 + q < N) {
            const d2 k = lds128(ku + 2 * (r0 + q));
            v[q] = rank2(v[q], k.x, k.y, hW0[u][0], hW1[u][0]);
            if (JOSEPH) {
              const d2 av = lds128(sA + u * HIST + 2 * (r0 + q));
              const d2 kc = (lane < LD) ? lds128(ku + 2 * lane) : k;
              v[q] = rank2(v[q], av.x, av.y, kc.x, kc.y);
            }
          }
        }
      }
#pragma unroll
      for (int q = 0; q < CH; ++q)
        if (r0 + q < N && lane < LD) gP[(r0 + q) * LD + lane] = v[q];
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
  __syncwarp();  // every lane is done with this filter's shared memory before the next TMA lands
  b = __shfl_sync(kFull, b_next, 0);
  }  // filter loop
  if (WPC > 1 && lane == 0) {  // the last warp out re-arms the work counter for the next launch
    __threadfence();
    if (atomicAdd(work + 1, 1ULL) == (unsigned long long)(total_warps - 1)) {
      work[0] = 0ULL;
      work[1] = 0ULL;
    }
  }
}

// -------------------------------------------------------------------------------------------------------
// launchers
// -------------------------------------------------------------------------------------------------------
size_t batch_regw_smem_bytes(int n_landmarks, int m, bool joseph) {
  const int n = 3 + 2 * n_landmarks, ld = (n + 3) & ~3;
  return ((size_t)n * ld + ld + (size_t)m * 2 * ld * (joseph ? 2 : 1) + 2) * sizeof(double);
}

// plain form: 1-6 observation slots; Joseph form: 1-4 (the K history stays in registers there)
bool batch_regw_supported(int m, bool joseph, bool generated) {
  return m >= 1 && m <= ((joseph || generated) ? 4 : 6) && !(joseph && generated);
}

template <int NLM, int MT, bool JOSEPH, bool GEN>
static cudaError_t launch_regw_g(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  const size_t smem = batch_regw_smem_bytes(NLM, MT, JOSEPH);
  auto kernel = batch_step_regw_kernel<NLM, MT, 1, JOSEPH, GEN>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  // One filter per one-warp CTA: the hardware CTA scheduler balances the load. (A persistent CTA per SM pulling
  // filters from a work counter was measured at 0.433 ms against 0.327 ms per 65 536-filter step in round 1 —
  // phase-locked warps, bursty HBM demand — and an L2 prefetch of the successor's data costs this kernel 5-7 % at every
  // distance: profiles/r1r_regw_prefetch_inputs_only_rejected.txt. Both code paths were removed in round 2.)
  kernel<<<(unsigned)batch, 32, smem, st>>>(a, batch, 0);
  return cudaGetLastError();
}

template <int NLM, int MT, bool JOSEPH>
static cudaError_t launch_regw_mt(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  if constexpr (!JOSEPH && MT <= 4) {  // generated inputs: plain form, up to 4 observation slots
    if (a.gen_record) return launch_regw_g<NLM, MT, false, true>(a, batch, st);
  }
  if (a.gen_record) return cudaErrorInvalidValue;
  return launch_regw_g<NLM, MT, JOSEPH, false>(a, batch, st);
}

template <int NLM, bool JOSEPH>
static cudaError_t launch_regw_t(const BatchStepArgs& a, int64_t batch, cudaStream_t st) {
  switch (a.m) {
    case 1: return launch_regw_mt<NLM, 1, JOSEPH>(a, batch, st);
    case 2: return launch_regw_mt<NLM, 2, JOSEPH>(a, batch, st);
    case 3: return launch_regw_mt<NLM, 3, JOSEPH>(a, batch, st);
    case 4: return launch_regw_mt<NLM, 4, JOSEPH>(a, batch, st);
    case 5: if constexpr (!JOSEPH) return launch_regw_mt<NLM, 5, false>(a, batch, st); else return cudaErrorInvalidValue;
    case 6: if constexpr (!JOSEPH) return launch_regw_mt<NLM, 6, false>(a, batch, st); else return cudaErrorInvalidValue;
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t batch_launch_step_regw(int n_landmarks, const BatchStepArgs& a, int64_t batch, bool joseph,
                                   cudaStream_t st) {
  switch (n_landmarks) {
    case 4: return joseph ? launch_regw_t<4, true>(a, batch, st) : launch_regw_t<4, false>(a, batch, st);
    case 8: return joseph ? launch_regw_t<8, true>(a, batch, st) : launch_regw_t<8, false>(a, batch, st);
    case 12: return joseph ? launch_regw_t<12, true>(a, batch, st) : launch_regw_t<12, false>(a, batch, st);
    case 16: return joseph ? launch_regw_t<16, true>(a, batch, st) : launch_regw_t<16, false>(a, batch, st);
    case 20: return joseph ? launch_regw_t<20, true>(a, batch, st) : launch_regw_t<20, false>(a, batch, st);
    case 24: return joseph ? launch_regw_t<24, true>(a, batch, st) : launch_regw_t<24, false>(a, batch, st);
    case 28: return joseph ? launch_regw_t<28, true>(a, batch, st) : launch_regw_t<28, false>(a, batch, st);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace ekf
