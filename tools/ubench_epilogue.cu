// Micro-benchmarks behind the K1 epilogue design: TMEM read bandwidth per SM for tcgen05.ld
// shapes, and the issue cost of the min/max/select instructions the top-2 reduction is made of.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench tools/ubench_epilogue.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int X>
__device__ __forceinline__ float ld_cols(uint32_t taddr);
template <>
__device__ __forceinline__ float ld_cols<16>(uint32_t taddr) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) x ^= r[i];
  return __uint_as_float(x);
}
template <>
__device__ __forceinline__ float ld_cols<32>(uint32_t taddr) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 32; i++) x ^= r[i];
  return __uint_as_float(x);
}

// nwarps warps (4 or 8) each sweep the 512 TMEM columns of their lane quarter `iters` times
template <int X>
__global__ void k_tmem(int iters, long long* cycles, float* sink) {
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tmem_ptr + (((warp & 3) * 32) << 16);
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++)
    for (int c = 0; c < 512; c += X) acc += ld_cols<X>(base + c);
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_ptr), "r"(512));
}

// independent chains of one instruction kind: MODE 0 FMNMX (min), 1 FMNMX3 (min3), 2 FSETP+FSEL, 3 FFMA, 4 LOP3, 5 FADD.SAT-ish
template <int MODE>
__global__ void k_alu(int iters, long long* cycles, float* sink, float seed) {
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = seed + i + threadIdx.x;
  float b = seed * 0.5f, c = seed * 0.25f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) a[i] = fminf(a[i], b + it);
      if (MODE == 1) a[i] = fminf(fminf(a[i], b), c + it);
      if (MODE == 2) a[i] = (b + it < a[i]) ? c : a[i];
      if (MODE == 3) a[i] = fmaf(a[i], b, c);
      if (MODE == 4) a[i] = __uint_as_float((__float_as_uint(a[i]) & 0xffffff80u) | (uint32_t)(it & 127));
      if (MODE == 5) a[i] = a[i] + __saturatef(fmaf(a[i], -b, c));
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++) s += a[i];
  if (s == 123.456f) sink[0] = s;
}


// ---- tcgen05.mma issue rate: `n` back-to-back MMAs (M = 128, N given, one k-step each) from zeroed shared memory
__device__ __forceinline__ uint64_t desc_sw32(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)6 << 61);
}
template <int TF32>
__global__ void k_mma(int n, uint32_t N, long long* cycles, uint32_t M = 128) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_ptr;
  __shared__ uint64_t bar;
  for (int i = threadIdx.x; i < 2 * 8192 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (threadIdx.x == 0) {
    const uint32_t idesc = TF32 ? ((1u << 4) | (2u << 7) | (2u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24))
                                : ((1u << 4) | (0u << 7) | (0u << 10) | ((N >> 3) << 17) | ((128u >> 4) << 24));   // f16 x f16 -> f32
    const uint64_t ad = desc_sw32(smem_u32(sm)), bd = desc_sw32(smem_u32(sm) + 4096);
    const long long t0 = clock64();
    for (int i = 0; i < n; i++) {
      if (TF32)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(
                         tmem_ptr + (i & 1) * 256), "l"(ad), "l"(bd), "r"(idesc), "r"(i > 1 ? 1 : 0) : "memory");
      else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(
                         tmem_ptr + (i & 1) * 256), "l"(ad), "l"(bd), "r"(idesc), "r"(i > 1 ? 1 : 0) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(
                     smem_u32(&bar)), "r"(0) : "memory");
    const long long t1 = clock64();
    cycles[0] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_ptr), "r"(512));
}

// ---- the GEMM kernel's issue pattern on every SM: units of `ks` k-steps x 3 passes into rotating accumulators.
// MODE 0: all three passes A-from-shared (SS); 1: all A-from-TMEM (TS); 2: two TS passes + one SS pass (the kernel's mix)
template <int MODE>
__global__ void k_mma_mix(int units, int ks, uint32_t seed, long long* cycles, unsigned long long* ns, int commits, uint32_t a_col = 0, uint32_t d_col = 128, uint32_t N = 128) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_ptr;
  __shared__ uint64_t bar, dummy[2];
  uint32_t x = seed + threadIdx.x * 2654435761u + blockIdx.x;
  for (int i = threadIdx.x; i < 3 * 8 * 8192 / 4; i += blockDim.x) {
    x = x * 1664525u + 1013904223u;
    reinterpret_cast<float*>(sm)[i] = seed ? (float)(int)(x >> 20) * 1e-3f : 0.f;
  }
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&dummy[0])), "r"(1 << 20));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&dummy[1])), "r"(1 << 20));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 0) {   // converged warp, one elected lane issues (the kernel's pattern)
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((N >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t alo = desc_sw32(smem_u32(sm)), bhi = desc_sw32(smem_u32(sm) + 8 * 4096), blo = desc_sw32(smem_u32(sm) + 8 * 4096 + 8 * 8192);
    unsigned long long g0, g1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    const long long t0 = clock64();
    for (int u = 0; u < units; u++) {
      const uint32_t d = tmem_ptr + d_col + (N == 256 ? 0u : (u % 3) * 128);
      for (int pass = 0; pass < 3; pass++) {
        const bool ts = MODE == 1 || (MODE == 2 && pass < 2);
        const uint64_t b = pass == 1 ? blo : bhi;
        for (int kk = 0; kk < ks; kk++) {
          const uint32_t acc = (pass | kk) ? 1 : 0;
          if (ts)
            asm volatile("{\n\t.reg .pred p, pe;\n\telect.sync _|pe, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                         "r"(tmem_ptr + a_col + kk * 8), "l"(b + kk * 256), "r"(idesc), "r"(acc) : "memory");
          else
            asm volatile("{\n\t.reg .pred p, pe;\n\telect.sync _|pe, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                         "l"(alo + kk * 256), "l"(b + kk * 256), "r"(idesc), "r"(acc) : "memory");
        }
      }
      for (int cm = 0; cm < commits; cm++)
        asm volatile("{\n\t.reg .pred pe;\n\telect.sync _|pe, 0xffffffff;\n\t@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(&dummy[cm & 1])) : "memory");
    }
    asm volatile("{\n\t.reg .pred pe;\n\telect.sync _|pe, 0xffffffff;\n\t@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("{\n\t.reg .pred p;\n\tW2:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D2;\n\tbra W2;\n\tD2:\n\t}" ::"r"(
                     smem_u32(&bar)), "r"(0) : "memory");
    const long long t1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    if (threadIdx.x == 0) { cycles[blockIdx.x] = t1 - t0; ns[blockIdx.x] = g1 - g0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_ptr), "r"(512));
}

int main() {
  long long* d_cyc; float* d_sink;
  cudaMalloc(&d_cyc, 1024 * sizeof(long long)); cudaMalloc(&d_sink, 4);
  long long h;
  const int iters = 200;
  for (int nw : {4, 8, 16}) {
    k_tmem<16><<<1, nw * 32>>>(iters, d_cyc, d_sink); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("tmem x16 warps=%2d: %.1f B/cycle/SM (%.0f cycles per 128x128 fp32 read)\n", nw, (double)nw * 32 * 512 * 4 * iters / h,
           h / (double)iters / (nw / 4.0) / 4.0);
    k_tmem<32><<<1, nw * 32>>>(iters, d_cyc, d_sink); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("tmem x32 warps=%2d: %.1f B/cycle/SM\n", nw, (double)nw * 32 * 512 * 4 * iters / h);
  }
  printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
  for (uint32_t N : {64u, 128u, 256u}) {
    const int n = 4096;
    k_mma<1><<<1, 128, 16384>>>(n, N, d_cyc); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("mma tf32 128x%ux8 : %.1f cycles per MMA\n", N, h / (double)n);
    k_mma<0><<<1, 128, 16384>>>(n, N, d_cyc); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("mma f16  128x%ux16: %.1f cycles per MMA\n", N, h / (double)n);
  }
  for (uint32_t N : {32u, 64u, 128u, 256u}) {
    const int n = 4096;
    k_mma<1><<<1, 128, 16384>>>(n, N, d_cyc, 64u); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("mma tf32 M=64 64x%ux8 : %.1f cycles per MMA\n", N, h / (double)n);
  }
  for (uint32_t N : {16u, 32u}) {
    const int n = 4096;
    k_mma<1><<<1, 128, 16384>>>(n, N, d_cyc, 128u); cudaDeviceSynchronize();
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("mma tf32 M=128 128x%ux8 : %.1f cycles per MMA\n", N, h / (double)n);
  }
  printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
  const char* names[] = {"FMNMX", "FMNMX3", "FSETP+FSEL", "FFMA", "LOP3", "FFMA+SAT+FADD"};
  for (int nw : {4, 8}) {
    for (int mode = 0; mode < 6; mode++) {
      const int it2 = 4096;
      switch (mode) {
        case 0: k_alu<0><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
        case 1: k_alu<1><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
        case 2: k_alu<2><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
        case 3: k_alu<3><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
        case 4: k_alu<4><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
        case 5: k_alu<5><<<1, nw * 32>>>(it2, d_cyc, d_sink, 3.f); break;
      }
      cudaDeviceSynchronize();
      cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
      printf("alu %-14s warps/SMSP=%d: %.2f cycles per (warp-level) loop element per SMSP\n", names[mode], nw / 4,
             h / (double)(it2 * 8) / 1.0);
    }
  }
  printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
  {
    unsigned long long* d_ns; cudaMalloc(&d_ns, 1024 * 8);
    static long long hc[1024]; static unsigned long long hn[1024];
    const char* mn[] = {"SS,SS,SS", "TS,TS,TS", "TS,TS,SS"};
    const size_t smem = 3 * 8 * 8192 + 1024;
    cudaFuncSetAttribute(k_mma_mix<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_mma_mix<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_mma_mix<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int grid : {148})
      for (int commits : {0, 1, 2})
        for (int mode = 0; mode < 3; mode++) {
          const int units = 4000, ks = 5;
          const uint32_t seed = 7u;
          if (mode == 0) k_mma_mix<0><<<grid, 128, smem>>>(units, ks, seed, d_cyc, d_ns, commits);
          if (mode == 1) k_mma_mix<1><<<grid, 128, smem>>>(units, ks, seed, d_cyc, d_ns, commits);
          if (mode == 2) k_mma_mix<2><<<grid, 128, smem>>>(units, ks, seed, d_cyc, d_ns, commits);
          cudaDeviceSynchronize();
          cudaMemcpy(hc, d_cyc, grid * 8, cudaMemcpyDeviceToHost); cudaMemcpy(hn, d_ns, grid * 8, cudaMemcpyDeviceToHost);
          long long mc = 0; unsigned long long mns = 0;
          for (int i = 0; i < grid; i++) { if (hc[i] > mc) mc = hc[i]; if (hn[i] > mns) mns = hn[i]; }
          printf("mma mix %s grid=%3d commits/unit=%d: %.1f cycles per MMA, %.2f GHz\n", mn[mode], grid, commits,
                 mc / (double)(units * 3 * ks), (double)mc / (double)mns);
        }
    printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
    // A-in-TMEM placement and N = 256
    struct { uint32_t a, d, N; } cfgs[] = {{0, 128, 128}, {448, 0, 128}, {384, 0, 128}, {0, 256, 256}, {256, 0, 256}};
    for (auto& c : cfgs)
      for (int mode = 0; mode < 2; mode++) {
        if (mode == 0) k_mma_mix<0><<<148, 128, smem>>>(4000, 5, 7u, d_cyc, d_ns, 1, c.a, c.d, c.N);
        else k_mma_mix<1><<<148, 128, smem>>>(4000, 5, 7u, d_cyc, d_ns, 1, c.a, c.d, c.N);
        cudaDeviceSynchronize();
        cudaMemcpy(hc, d_cyc, 148 * 8, cudaMemcpyDeviceToHost);
        long long mc = 0;
        for (int i = 0; i < 148; i++) if (hc[i] > mc) mc = hc[i];
        printf("mma %s A col %u, D col %u, N=%u: %.1f cycles per MMA\n", mode ? "TS" : "SS", c.a, c.d, c.N, mc / (double)(4000 * 15));
      }
    printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
