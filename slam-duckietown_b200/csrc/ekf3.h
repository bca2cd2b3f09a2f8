// ekf3.h — launchers of the batched 3-state EKF kernels (internal).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
namespace ekf {
cudaError_t ekf3_launch_predict(int64_t B, double* st, double* cv, const double* v, const double* om,
                                const double* dt, cudaStream_t s);
cudaError_t ekf3_launch_update(int64_t B, double* st, double* cv, const double* rng, const double* brg,
                               const double* lx, const double* ly, cudaStream_t s);
}  // namespace ekf
