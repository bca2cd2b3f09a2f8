// ekf3.cu — batched 3-state localisation EKF (src/EKF-SLAM.py:29-84), one thread per filter (sm_100a, fp64).
// SoA layout: state [3][B], cov [9][B] (row-major 3x3 index outermost) so every access is coalesced.
// Noise constants: motion diag(0.1, 0.1, radians(5)) (:12), observation diag(0.5, 0.5) (:13).
#include "ekf_common.cuh"
#include "ekf3.h"

namespace ekf {

__global__ void __launch_bounds__(256) ekf3_predict_kernel(int64_t B, double* __restrict__ st,
                                                           double* __restrict__ cv, const double* __restrict__ v_,
                                                           const double* __restrict__ om_,
                                                           const double* __restrict__ dt_) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double v = v_[b], om = om_[b], dt = dt_[b];
  const double x = st[b], y = st[B + b], th = st[2 * B + b];
  double s, c;
  sincos(th, &s, &c);
  double dx, dy, dth;
  if (fabs(om) > 1e-6) {  // :34-37
    double s2, c2;
    sincos(th + om * dt, &s2, &c2);
    dx = -v / om * s + v / om * s2;
    dy = v / om * c - v / om * c2;
    dth = om * dt;
  } else {  // :38-41
    dx = v * c * dt;
    dy = v * s * dt;
    dth = 0.0;
  }
  const double t2 = th + dth;
  double s3, c3;
  sincos(t2, &s3, &c3);
  st[b] = x + dx;
  st[B + b] = y + dy;
  st[2 * B + b] = atan2(s3, c3);  // :45
  // F = [[1,0,f0],[0,1,f1],[0,0,1]], straight-line Jacobian at the PRE-update heading (:48-52)
  const double f0 = -v * dt * s, f1 = v * dt * c;
  double p[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) p[i][k] = cv[(3 * i + k) * B + b];
  // F P
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    p[0][k] += f0 * p[2][k];
    p[1][k] += f1 * p[2][k];
  }
  // (F P) F^T
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i][0] += f0 * p[i][2];
    p[i][1] += f1 * p[i][2];
  }
  p[0][0] += 0.1;
  p[1][1] += 0.1;
  p[2][2] += 5.0 * (kPi / 180.0);  // np.radians(5)
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) cv[(3 * i + k) * B + b] = p[i][k];
}

__global__ void __launch_bounds__(256) ekf3_update_kernel(int64_t B, double* __restrict__ st,
                                                          double* __restrict__ cv, const double* __restrict__ rng,
                                                          const double* __restrict__ brg,
                                                          const double* __restrict__ lx,
                                                          const double* __restrict__ ly) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double x = st[b], y = st[B + b], th = st[2 * B + b];
  const double dx = lx[b] - x, dy = ly[b] - y;
  const double q = dx * dx + dy * dy, sq = sqrt(q);
  const double h[2][3] = {{-dx / sq, -dy / sq, 0.0}, {dy / q, -dx / q, -1.0}};  // :67-70
  double p[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) p[i][k] = cv[(3 * i + k) * B + b];
  double hp[2][3], pht[3][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int k = 0; k < 3; ++k) hp[a][k] = h[a][0] * p[0][k] + h[a][1] * p[1][k] + h[a][2] * p[2][k];  // rows
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int a = 0; a < 2; ++a) pht[i][a] = p[i][0] * h[a][0] + p[i][1] * h[a][1] + p[i][2] * h[a][2];  // cols
  double S[2][2], Si[2][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int e = 0; e < 2; ++e) S[a][e] = hp[a][0] * h[e][0] + hp[a][1] * h[e][1] + hp[a][2] * h[e][2];
  S[0][0] += 0.5;
  S[1][1] += 0.5;
  inv2(S, Si);
  double K[3][2];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    K[i][0] = pht[i][0] * Si[0][0] + pht[i][1] * Si[1][0];
    K[i][1] = pht[i][0] * Si[0][1] + pht[i][1] * Si[1][1];
  }
  const double y0 = rng[b] - sq;
  const double e1 = brg[b] - (atan2(dy, dx) - th);
  double se, ce;
  sincos(e1, &se, &ce);
  const double y1 = atan2(se, ce);  // :79
  st[b] = x + (K[0][0] * y0 + K[0][1] * y1);
  st[B + b] = y + (K[1][0] * y0 + K[1][1] * y1);
  st[2 * B + b] = th + (K[2][0] * y0 + K[2][1] * y1);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) cv[(3 * i + k) * B + b] = p[i][k] - (K[i][0] * hp[0][k] + K[i][1] * hp[1][k]);
}

cudaError_t ekf3_launch_predict(int64_t B, double* st, double* cv, const double* v, const double* om,
                                const double* dt, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  ekf3_predict_kernel<<<(unsigned)((B + 255) / 256), 256, 0, s>>>(B, st, cv, v, om, dt);
  return cudaGetLastError();
}
cudaError_t ekf3_launch_update(int64_t B, double* st, double* cv, const double* rng, const double* brg,
                               const double* lx, const double* ly, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  ekf3_update_kernel<<<(unsigned)((B + 255) / 256), 256, 0, s>>>(B, st, cv, rng, brg, lx, ly);
  return cudaGetLastError();
}

}  // namespace ekf
