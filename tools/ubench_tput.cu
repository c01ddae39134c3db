// Issue throughput of one warp (and of 4 warps on one SM sub-partition set) for the Viterbi instruction mix:
// independent DADD, DSETP+2xFSEL, and the full per-position body.  cycles per warp-instruction.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/ubench_tput tools/ubench_tput.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int K = 8;
__global__ void k(double* out, long long* cyc, int n, double a, double b) {
  double x[K], m[K];
#pragma unroll
  for (int i = 0; i < K; i++) { x[i] = a + i + threadIdx.x; m[i] = b * i; }
  long long t0, t1;
  __syncthreads();
  t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < K; i++) x[i] = __dadd_rn(x[i], b);
  }
  t1 = clock64(); if (threadIdx.x % 32 == 0) cyc[(threadIdx.x / 32) * 4 + 0] = t1 - t0;
  __syncthreads();
  t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < K; i++) { const bool f = x[(i + 1) % K] < m[i]; m[i] = f ? x[(i + 1) % K] : m[i]; }
#pragma unroll
    for (int i = 0; i < K; i++) x[i] = __longlong_as_double(__double_as_longlong(x[i]) ^ (it & 1));
  }
  t1 = clock64(); if (threadIdx.x % 32 == 0) cyc[(threadIdx.x / 32) * 4 + 1] = t1 - t0;
  __syncthreads();
  // the per-position body: 4 DADD, 2 DSETP, 4 FSEL (no shuffles): K positions per frame, frame-to-frame dependent
  t0 = clock64();
  for (int it = 0; it < n; it++) {
    double nq[K];
#pragma unroll
    for (int i = 0; i < K; i++) {
      const double p1 = x[(i + K - 1) % K], p2 = x[(i + K - 2) % K];
      const double lv = __dadd_rn(x[i], a), fv = __dadd_rn(p1, b), kv = __dadd_rn(p2, a);
      const bool f1 = fv < lv; const double mn = f1 ? fv : lv; const bool f2 = kv < mn;
      nq[i] = __dadd_rn(m[i], f2 ? kv : mn);
    }
#pragma unroll
    for (int i = 0; i < K; i++) x[i] = nq[i];
  }
  t1 = clock64(); if (threadIdx.x % 32 == 0) cyc[(threadIdx.x / 32) * 4 + 2] = t1 - t0;
  double s = 0;
#pragma unroll
  for (int i = 0; i < K; i++) s += x[i] + m[i];
  out[threadIdx.x] = s;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 8 * 1024); cudaMalloc(&cyc, 32 * 4 * 8);
  const int n = 2048;
  for (int warps : {1, 4, 8, 16}) {
    for (int r = 0; r < 2; r++) k<<<1, 32 * warps>>>(out, cyc, n, 1.5, 0.25);
    long long h[4]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%2d warps/SM: DADD %.2f cyc/inst   DSETP+2FSEL %.2f cyc/triple   position body (4 DADD 2 DSETP 4 FSEL) %.1f cyc/position\n", warps,
           h[0] / (double)(n * K), h[1] / (double)(n * K), h[2] / (double)(n * K));
  }
  return 0;
}
