// Warp-uniform, runtime-indexed reads of 34 doubles per "cell": __constant__ (LDC / ULDC) against shared-memory broadcast
// (LDS.128).  16 warps per SM, every SM; cycles per warp-cell.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/ubench_const tools/ubench_const.cu
#include <cstdio>
#include <cuda_runtime.h>
__constant__ double cc[4096];
template <bool CONST>
__global__ void __launch_bounds__(512) k(const double* g, double* out, long long* cyc, int n, int off) {
  __shared__ __align__(16) double sh[64];
  if (threadIdx.x < 64) sh[threadIdx.x] = g[off + threadIdx.x];
  __syncthreads();
  double a = threadIdx.x, l0 = 0, l1 = 0;
  long long t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      double v0, v1;
      if (CONST) { v0 = cc[off + 2 * i]; v1 = cc[off + 2 * i + 1]; }
      else { const double2 v = *reinterpret_cast<const double2*>(sh + 2 * i); v0 = v.x; v1 = v.y; }
      l0 = __dadd_rn(l0, __dmul_rn(a, v0));
      l1 = __dadd_rn(l1, __dmul_rn(a, v1));
    }
    a += 1.0;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = l0 + l1;
}
int main() {
  double *g, *out; long long* cyc;
  cudaMalloc(&g, 4096 * 8); cudaMalloc(&out, 148 * 512 * 8); cudaMalloc(&cyc, 8);
  double h[4096]; for (int i = 0; i < 4096; i++) h[i] = 1.0 + i * 1e-3;
  cudaMemcpy(g, h, sizeof(h), cudaMemcpyHostToDevice); cudaMemcpyToSymbol(cc, h, sizeof(h));
  const int n = 2000;
  for (int r = 0; r < 2; r++) {
    long long c0, c1;
    k<false><<<148, 512>>>(g, out, cyc, n, 340); cudaMemcpy(&c0, cyc, 8, cudaMemcpyDeviceToHost);
    k<true><<<148, 512>>>(g, out, cyc, n, 340); cudaMemcpy(&c1, cyc, 8, cudaMemcpyDeviceToHost);
    printf("16 warps/SM, per warp-iteration (32 doubles + 64 FP64 ops): shared broadcast %.1f cycles, constant %.1f cycles\n", c0 / (double)n, c1 / (double)n);
  }
  return 0;
}
