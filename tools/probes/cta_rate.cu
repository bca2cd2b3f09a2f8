// CTA dispatch-rate probe: how long do N one-warp CTAs with S bytes of dynamic shared memory take when each does
// (almost) nothing, or spins for a fixed number of cycles? Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 cta_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(32, 16) k(double* out, long long spin) {
  extern __shared__ double sm[];
  sm[threadIdx.x] = threadIdx.x;
  if (spin > 0) {
    const long long t0 = clock64();
    while (clock64() - t0 < spin) {}
  }
  if (sm[31 - threadIdx.x] < 0) out[blockIdx.x] = 1.0;
}
int main() {
  double* out; cudaMalloc(&out, 8 << 20);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int grids[] = {65536, 262144};
  const int smems[] = {1024, 15616, 17408};
  const long long spins[] = {0, 2000, 10000, 20000};
  for (int s : smems) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    for (int g : grids) for (long long sp : spins) {
      for (int w = 0; w < 3; ++w) k<<<g, 32, s>>>(out, sp);
      cudaEventRecord(a);
      for (int i = 0; i < 10; ++i) k<<<g, 32, s>>>(out, sp);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
      int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, 32, s);
      printf("smem %6d B  grid %7d  spin %6lld cyc  occ %2d/SM : %8.4f ms  = %6.2f CTAs/us/SM\n", s, g, sp, occ, ms,
             g / (ms * 1e3) / 148.0);
    }
  }
  return 0;
}
