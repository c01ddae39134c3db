// Dependent-chain latencies on one warp (cycles per link): DADD, DSETP+FSEL (f64 min), 64-bit SHFL.UP, LDS.64, FSEL, ISETP.64 pair.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/ubench_lat tools/ubench_lat.cu ; run on the GPU box
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, int n, double a, double b) {
  __shared__ double sh[64];
  sh[threadIdx.x] = a + threadIdx.x; sh[threadIdx.x + 32] = b;
  __syncwarp();
  double x = a + threadIdx.x, y = b;
  long long t0, t1;
  // DADD chain
  t0 = clock64();
  for (int i = 0; i < n; i++) { x = __dadd_rn(x, y); x = __dadd_rn(x, y); x = __dadd_rn(x, y); x = __dadd_rn(x, y); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
  // DSETP + select chain (strict-less min against a moving candidate)
  double m = x, c = y;
  t0 = clock64();
  for (int i = 0; i < n; i++) {
    m = c < m ? c : m; c = __longlong_as_double(__double_as_longlong(c) ^ 1);   // keep c opaque
    m = c < m ? c : m; m = c < m ? c : m; m = c < m ? c : m;
  }
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
  // 64-bit shuffle chain
  double s = m;
  t0 = clock64();
  for (int i = 0; i < n; i++) { s = __shfl_up_sync(0xffffffffu, s, 1); s = __shfl_up_sync(0xffffffffu, s, 1); s = __shfl_up_sync(0xffffffffu, s, 1); s = __shfl_up_sync(0xffffffffu, s, 1); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
  // LDS.64 pointer-chase
  int idx = threadIdx.x;
  double l = 0;
  t0 = clock64();
  for (int i = 0; i < n; i++) {
    l = sh[idx]; idx = (int)(__double_as_longlong(l) & 31); l = sh[idx]; idx = (int)(__double_as_longlong(l) & 31);
    l = sh[idx]; idx = (int)(__double_as_longlong(l) & 31); l = sh[idx]; idx = (int)(__double_as_longlong(l) & 31);
  }
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
  // the Viterbi frame chain: shfl -> dadd -> (dsetp, sel) x2 -> dadd
  double q = x;
  t0 = clock64();
  for (int i = 0; i < n; i++) {
    const double p1 = __shfl_up_sync(0xffffffffu, q, 1);
    const double lv = __dadd_rn(q, y), fv = __dadd_rn(p1, b);
    const double m1 = fv < lv ? fv : lv;
    const double kv = __dadd_rn(p1, a);
    const double best = kv < m1 ? kv : m1;
    q = __dadd_rn(c, best);
  }
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
  // integer 64-bit compare + select chain
  unsigned long long um = (unsigned long long)__double_as_longlong(q), uc = (unsigned long long)__double_as_longlong(c);
  t0 = clock64();
  for (int i = 0; i < n; i++) { um = uc < um ? uc : um; uc ^= 1; um = uc < um ? uc : um; um = uc < um ? uc : um; um = uc < um ? uc : um; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[5] = t1 - t0;
  out[threadIdx.x] = x + m + s + l + q + (double)um;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 256); cudaMalloc(&cyc, 64);
  const int n = 4096;
  for (int r = 0; r < 2; r++) k<<<1, 32>>>(out, cyc, n, 1.5, 0.25);
  long long h[6]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("cycles per link: DADD %.1f  DSETP+sel(f64) %.1f  SHFL64 %.1f  LDS64+use %.1f  viterbi-frame-chain %.1f per frame  ISETP64+sel %.1f\n",
         h[0] / (4.0 * n), h[1] / (4.0 * n), h[2] / (4.0 * n), h[3] / (4.0 * n), h[4] / (1.0 * n), h[5] / (4.0 * n));
  return 0;
}
