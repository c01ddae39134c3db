// K1, tensor-core form: the max-approx GMM scores as the dense contraction
//     [frames x K] . [K x densities],   K = 2d+1  (x^2 | x | 1)  or  d+1 (x | 1) when variances are pooled
// per state (the x^2 block is then common to all densities of a state and cannot change the argmin),
// on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with error-compensated 3xTF32 operands
// (A = Ahi + Alo, B = Bhi + Blo; D = Ahi.Bhi + Alo.Bhi + Ahi.Blo accumulated in fp32 in TMEM), operands staged in
// shared memory by TMA (SWIZZLE_128B, K-major), accumulators in TMEM, per-state top-2 (min, argmin,
// second min) fused into the tcgen05.ld epilogue.
//
// Like the FFMA scorer this kernel is a FILTER: its f32 top-2 go to k_refine (below), which certifies the
// argmin against a rigorous error bound and evaluates the winner's score in f64 in the reference's exact
// operation order (Mixtures.cpp:645-713) — or rescans the state in f64 when the gap is inside the bound.
// Results are therefore bit-identical to the reference; the tensor cores only decide where f64 work goes.
//
// Roles per CTA (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected
// lane), warps 2..5 = epilogue (TMEM lane quarter = warp_id % 4).  One CTA owns 128 frames (TMEM lanes) and
// walks the density blocks of its task's states; B tiles of 128 densities stream through a 2-stage
// full/empty mbarrier ring, accumulators rotate through 4 TMEM slots of 128 columns.
#include <cuda.h>

#include "sgmm_internal.cuh"

namespace sgmm_tc {

constexpr int BM = 128;        // frames per CTA (TMEM lanes)
constexpr int BN = 128;        // densities per B stage / MMA N
constexpr int kMaxStage = 4;   // B stages in the TMA ring (the launcher picks what shared memory allows)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 3-D tile {8 floats, 128 rows, ksteps} -> shared [kstep][row][8 floats] (one 4 KB SWIZZLE_32B block per k-step)
__device__ __forceinline__ void tma_load_rows(void* dst, const CUtensorMap* map, int row0, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(0), "r"(row0), "r"(0)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]   (kind::tf32, fp32 accumulate)
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// 16 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_32B.  One k-step (8 tf32 = 32 bytes) of a 128-row operand
// is its own 4 KB block [128 rows][32 B]: 8-row groups are 256 B apart (SBO), LBO unused.  K = 2d+1 or d+1 is
// short and odd-sized here (34 -> 5 k-steps at d = 33), so 32-byte atoms waste no shared memory or L2
// bandwidth on K padding, which 128-byte atoms (K padded to 64) did.
__device__ __forceinline__ uint64_t make_desc_sw32(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // start address
  d |= (uint64_t)1 << 16;                           // leading byte offset (ignored for swizzled K-major)
  d |= (uint64_t)(256 >> 4) << 32;                  // stride byte offset
  d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
  d |= (uint64_t)6 << 61;                           // layout type: SWIZZLE_32B
  return d;
}
// Instruction descriptor: D = f32, A = B = tf32, both K-major, M x N.
__host__ __device__ inline uint32_t make_idesc_tf32(uint32_t M, uint32_t N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

__device__ __forceinline__ float tf32_rna(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// A operand builder: row f = [x^2 (if full) | x | 1 | 0...] split into tf32 hi / lo; also the per-frame norms
// the error bound needs.  One warp per frame.
__global__ void k_tc_prep(const float* __restrict__ x, uint64_t n_frames, uint32_t dim, int full, uint32_t KP,
                          float* __restrict__ Ahi, float* __restrict__ Alo, float2* __restrict__ xnorm) {
  const uint64_t f = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (f >= n_frames) return;
  const float* xf = x + f * dim;
  const uint32_t nsq = full ? dim : 0;
  float s2 = 0.f, s4 = 0.f;
  for (uint32_t k = lane; k < KP; k += 32) {
    float a = 0.f;
    if (k < nsq) { const float v = xf[k]; a = v * v; }
    else if (k < nsq + dim) a = xf[k - nsq];
    else if (k == nsq + dim) a = 1.f;
    if (k >= nsq && k < nsq + dim) { s2 += a * a; s4 += a * a * a * a; }
    const float hi = tf32_rna(a);
    Ahi[f * KP + k] = hi;
    Alo[f * KP + k] = tf32_rna(a - hi);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s2 += __shfl_xor_sync(0xffffffffu, s2, o); s4 += __shfl_xor_sync(0xffffffffu, s4, o); }
  if (lane == 0) xnorm[f] = make_float2(sqrtf(s2) * 1.0001f, sqrtf(s4) * 1.0001f);   // ||x||, ||x^2|| (rounded up)
}


struct TcParams {
  uint32_t ksteps;             // K steps of 8 (KP = 8 * ksteps)
  uint32_t nstage;             // B stages in flight (<= kMaxStage)
  const uint32_t* col_off;     // [n_states + 1] first B row (padded density layout) of every state
  uint64_t n_tasks;
  // gathered mode: task = {utt, t0 (multiple of 128)}
  const uint4* tasks;
  const uint64_t* frame_off;
  const uint32_t* uniq;
  const uint32_t* uniq_off;
  const uint32_t* nuniq_pad;
  const uint64_t* E_off;
  // dense mode
  int dense;
  uint64_t f0, nframes;
  uint32_t n_states;
  // candidates, indexed like the emission matrix (gathered) / [frame - f0][state] (dense)
  float* cb1;
  float* cb2;
  uint32_t* ci1;
};

// NHALF frame halves of 128 rows (TMEM lanes) per CTA share every B tile: NHALF = 2 halves the L2 -> shared
// traffic per flop and gives the tensor pipe two accumulators per B stage.  Warps: 0 = TMA producer, 1 = TMEM
// allocator + MMA issuer (one lane), 2 .. 2 + 4*NHALF - 1 = epilogue (TMEM lane quarter = warp_id % 4, frame
// half = (warp_id - 2) / 4).  TMEM: NSLOT accumulator slots x NHALF halves x BN columns = 512 columns.
template <int NHALF>
__global__ void __launch_bounds__(64 + 128 * NHALF, 1) k_tc_gemm(const __grid_constant__ CUtensorMap mapAhi,
                                                                  const __grid_constant__ CUtensorMap mapAlo,
                                                                  const __grid_constant__ CUtensorMap mapBhi,
                                                                  const __grid_constant__ CUtensorMap mapBlo, TcParams p) {
  extern __shared__ unsigned char smem_raw[];
  constexpr uint32_t NSLOT = 4 / NHALF;                // accumulator slots (of NHALF x BN columns)
  constexpr uint32_t KSTEP_BYTES = 128 * 32;           // one k-step of a 128-row operand
  const uint32_t op_bytes = p.ksteps * KSTEP_BYTES;    // one 128-row operand (hi or lo)
  const uint32_t a_bytes = 2 * NHALF * op_bytes;       // hi + lo of every half
  const uint32_t b_stage = 2 * op_bytes;               // hi + lo
  const uint32_t nstage = p.nstage;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* smem = smem_raw + (base - smem_u32(smem_raw));
  unsigned char* sA = smem;                            // [half][hi, lo][kstep][128 rows][32 B]
  unsigned char* sB = smem + a_bytes;                  // [stage][hi, lo][kstep][128 rows][32 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + nstage * b_stage);
  uint64_t* a_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + kMaxStage;
  uint64_t* tfull = empty + kMaxStage;
  uint64_t* tempty = tfull + NSLOT;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + NSLOT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = blockIdx.x;
  constexpr uint32_t ROWS = BM * NHALF;

  // ---- task decode (uniform over the CTA)
  uint64_t row0;        // first A row (absolute frame index)
  uint32_t nvalid, nu, out_stride;
  const uint32_t* ulist = nullptr;
  uint64_t out_base;
  if (p.dense) {
    const uint64_t r0 = task * ROWS;
    row0 = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)ROWS, p.nframes - r0);
    nu = p.n_states;
    out_stride = p.n_states;
    out_base = r0 * p.n_states;
  } else {
    const uint4 t = p.tasks[task];
    const uint64_t uf0 = p.frame_off[t.x];
    const uint32_t T = (uint32_t)(p.frame_off[t.x + 1] - uf0);
    row0 = uf0 + t.y;
    nvalid = min(ROWS, T - t.y);
    ulist = p.uniq + p.uniq_off[t.x];
    nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
    out_stride = p.nuniq_pad[t.x];
    out_base = p.E_off[t.x] + (uint64_t)t.y * out_stride;
  }
  const uint32_t nh = (NHALF == 2 && nvalid > (uint32_t)BM) ? 2u : 1u;   // halves that hold valid frames

  // ---- one-time setup
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAhi));
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBhi));
    mbar_init(a_full, 1);
    for (uint32_t i = 0; i < nstage; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (uint32_t i = 0; i < NSLOT; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4 * nh); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_expect_tx(a_full, 2 * nh * op_bytes);
      for (uint32_t h = 0; h < nh; h++) {
        tma_load_rows(sA + (2 * h) * op_bytes, &mapAhi, (int)(row0 + h * BM), a_full);
        tma_load_rows(sA + (2 * h + 1) * op_bytes, &mapAlo, (int)(row0 + h * BM), a_full);
      }
      uint32_t it = 0;
      for (uint32_t j = 0; j < nu; j++) {
        const uint32_t s = ulist ? ulist[j] : j;
        const uint32_t c0 = p.col_off[s], nc = p.col_off[s + 1] - c0;
        for (uint32_t n0 = 0; n0 < nc; n0 += BN, it++) {
          const uint32_t stage = it % nstage;
          mbar_wait(&empty[stage], ((it / nstage) & 1) ^ 1);
          mbar_expect_tx(&full[stage], b_stage);
          unsigned char* dst = sB + stage * b_stage;
          tma_load_rows(dst, &mapBhi, (int)(c0 + n0), &full[stage]);
          tma_load_rows(dst + op_bytes, &mapBlo, (int)(c0 + n0), &full[stage]);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      mbar_wait(a_full, 0);
      uint32_t it = 0;
      for (uint32_t j = 0; j < nu; j++) {
        const uint32_t s = ulist ? ulist[j] : j;
        const uint32_t nc = p.col_off[s + 1] - p.col_off[s];
        for (uint32_t n0 = 0; n0 < nc; n0 += BN, it++) {
          const uint32_t N = min((uint32_t)BN, nc - n0);
          const uint32_t stage = it % nstage, slot = it % NSLOT;
          mbar_wait(&tempty[slot], ((it / NSLOT) & 1) ^ 1);
          mbar_wait(&full[stage], (it / nstage) & 1);
          tc_fence_after();
          const uint32_t idesc = make_idesc_tf32(BM, N);
          const uint32_t bhi = smem_u32(sB + stage * b_stage), blo = bhi + op_bytes;
          for (uint32_t h = 0; h < nh; h++) {
            const uint32_t d = tmem_base + (slot * NHALF + h) * BN;
            const uint32_t ahi = smem_u32(sA + (2 * h) * op_bytes), alo = ahi + op_bytes;
            uint32_t acc = 0;
#pragma unroll 1
            for (int pass = 0; pass < 3; pass++) {
              const uint32_t a0 = pass == 1 ? alo : ahi;
              const uint32_t b0 = pass == 2 ? blo : bhi;
              for (uint32_t kk = 0; kk < p.ksteps; kk++) {
                tc_mma_tf32(d, make_desc_sw32(a0 + kk * KSTEP_BYTES), make_desc_sw32(b0 + kk * KSTEP_BYTES), idesc, acc);
                acc = 1;
              }
            }
          }
          tc_commit(&empty[stage]);   // smem stage reusable once these MMAs have read it
          tc_commit(&tfull[slot]);    // accumulator slot complete (all halves)
        }
      }
    }
  } else {
    // ===== epilogue: per-state top-2 over the accumulator columns =====
    const uint32_t quarter = warp & 3;
    const uint32_t half = (uint32_t)(warp - 2) >> 2;
    if (half < nh) {
      const uint32_t row = half * BM + quarter * 32 + lane;
      const uint32_t tlane = tmem_base + ((quarter * 32) << 16);
      uint32_t it = 0;
      for (uint32_t j = 0; j < nu; j++) {
        const uint32_t s = ulist ? ulist[j] : j;
        const uint32_t nc = p.col_off[s + 1] - p.col_off[s];
        float b1 = INFINITY, b2 = INFINITY;
        uint32_t i1 = 0;
        for (uint32_t n0 = 0; n0 < nc; n0 += BN, it++) {
          const uint32_t N = min((uint32_t)BN, nc - n0);
          const uint32_t slot = it % NSLOT;
          mbar_wait(&tfull[slot], (it / NSLOT) & 1);
          tc_fence_after();
          const uint32_t tcol = tlane + (slot * NHALF + half) * BN;
          for (uint32_t cc = 0; cc < N; cc += 16) {
            float v[16];
            tc_ld16(tcol + cc, v);
#pragma unroll
            for (int e = 0; e < 16; e++) {
              b2 = fminf(b2, fmaxf(b1, v[e]));
              if (v[e] < b1) { b1 = v[e]; i1 = n0 + cc + e; }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[slot]);
        }
        if (row < nvalid) {
          const uint64_t o = out_base + (uint64_t)row * out_stride + j;
          p.cb1[o] = b1;
          p.cb2[o] = b2;
          p.ci1[o] = i1;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}


// ---- refinement: certify the f32 argmin, evaluate the winner in f64, queue near-ties for a full f64 rescan
struct RefineParams {
  ExactTables m;
  const uint32_t* state_off;
  const float* x;
  const float* cb1;
  const float* cb2;
  const uint32_t* ci1;
  const float2* xnorm;
  const float4* sguard;        // per state {Gmax, Hmax, cabs, -}
  float eps;                   // relative error of one 3xTF32 dot product term + fp32 accumulation (see host)
  uint64_t n_cells;
  int dense;
  uint64_t f0;
  uint32_t n_states;
  uint32_t n_utts;
  const uint64_t* frame_off;
  const uint64_t* E_off;
  const uint32_t* uniq;
  const uint32_t* uniq_off;
  const uint32_t* nuniq_pad;
  double* out;                 // emission matrix / dense scores
  uint32_t* argmin;            // dense only (may be NULL)
  uint16_t* amin16;            // gathered only (may be NULL)
  unsigned long long* rescan;  // [0] count, [1..cap] cell ids
  unsigned long long cap;
  unsigned long long* fallback_count;
};

// cell -> (frame, state); returns false for the padding column of a gathered row
__device__ __forceinline__ bool decode_cell(const RefineParams& p, uint64_t e, uint64_t& frame, uint32_t& s) {
  if (p.dense) {
    frame = p.f0 + e / p.n_states;
    s = (uint32_t)(e % p.n_states);
    return true;
  }
  uint32_t lo = 0, hi = p.n_utts;          // largest u with E_off[u] <= e
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (p.E_off[mid] <= e) lo = mid; else hi = mid;
  }
  const uint64_t local = e - p.E_off[lo];
  const uint32_t ncol = p.nuniq_pad[lo];
  const uint32_t j = (uint32_t)(local % ncol);
  const uint32_t nu = p.uniq_off[lo + 1] - p.uniq_off[lo];
  if (j >= nu) return false;
  frame = p.frame_off[lo] + local / ncol;
  s = p.uniq[p.uniq_off[lo] + j];
  return true;
}

__global__ void __launch_bounds__(256) k_tc_refine(RefineParams p) {
  const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.n_cells) return;
  uint64_t frame;
  uint32_t s;
  if (!decode_cell(p, e, frame, s)) return;
  const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
  double best = 1e10;
  uint32_t bk = 0;
  if (k1 > k0) {
    const float b1 = p.cb1[e], b2 = p.cb2[e];
    const uint32_t i1 = p.ci1[e];
    const float2 xn = p.xnorm[frame];
    const float4 g = p.sguard[s];
    const float guard = 2.f * p.eps * (xn.x * g.x + xn.y * g.y + g.z);
    const bool certain = (b2 - b1) > guard && i1 < k1 - k0;
    if (!certain) {
      const unsigned long long slot = atomicAdd(p.rescan, 1ull);
      if (slot < p.cap) { p.rescan[1 + slot] = e; return; }
      // list full: rescan here
      const float* xf = p.x + frame * p.m.dim;
      for (uint32_t k = k0; k < k1; k++) {
        const double sc = sgmm_exact_score(p.m, xf, k);
        if (sc < best) { best = sc; bk = k - k0; }
      }
    } else {
      const double sc = sgmm_exact_score_rm(p.m, p.x + frame * p.m.dim, k0 + i1);
      if (sc < best) { best = sc; bk = i1; }
    }
  }
  p.out[e] = best;
  if (p.argmin) p.argmin[e] = bk;
  if (p.amin16) p.amin16[e] = (uint16_t)bk;
}

// one warp per queued near-tie cell: exact f64 rescan of the whole state, first minimum wins
__global__ void __launch_bounds__(256) k_tc_rescan(RefineParams p) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const unsigned long long count = min(p.rescan[0], p.cap);
  for (uint64_t i = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < count; i += nwarps) {
    const uint64_t e = p.rescan[1 + i];
    uint64_t frame;
    uint32_t s;
    decode_cell(p, e, frame, s);
    const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
    const float* xf = p.x + frame * p.m.dim;
    double best = 1e10;
    uint32_t bk = 0xffffffffu;
    for (uint32_t k = k0 + lane; k < k1; k += 32) {
      const double sc = sgmm_exact_score(p.m, xf, k);
      if (sc < best) { best = sc; bk = k - k0; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const uint32_t ok = __shfl_xor_sync(0xffffffffu, bk, o);
      if (ob < best || (ob == best && ok < bk)) { best = ob; bk = ok; }
    }
    if (lane == 0) {
      p.out[e] = best;
      if (p.argmin) p.argmin[e] = bk == 0xffffffffu ? 0 : bk;
      if (p.amin16) p.amin16[e] = bk == 0xffffffffu ? 0 : (uint16_t)bk;
      if (p.fallback_count) atomicAdd(p.fallback_count, 1ull);
    }
  }
}

}  // namespace sgmm_tc

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// f32 matrix [rows][KP] (KP = 8 * ksteps innermost) seen as the 3-D tensor {8, rows, ksteps}: one box =
// 128 rows x all k-steps, delivered as [kstep][row][8 floats] with the 32-byte swizzle the MMA descriptors name.
int make_map(sgmm_ctx* c, CUtensorMap* map, const float* base, uint64_t rows, uint32_t ksteps) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) SGMM_FAIL(c, "cuTensorMapEncodeTiled is not available from the driver");
  const uint32_t KP = 8 * ksteps;
  cuuint64_t dims[3] = {8, rows, ksteps};
  cuuint64_t strides[2] = {(cuuint64_t)KP * sizeof(float), 8 * sizeof(float)};
  cuuint32_t box[3] = {8, 128, ksteps};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) SGMM_FAIL(c, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return 0;
}

// shared-memory plan: A = 2 * NHALF operands, each B stage = 2 operands of ksteps * 4 KB
constexpr size_t kSmemLimit = 227 * 1024;
size_t tc_smem(uint32_t ksteps, int nhalf, int nstage) {
  return 1024 + (size_t)(2 * nhalf + 2 * nstage) * ksteps * 4096 + 256;
}

}  // namespace

int sgmm_launch_emission_tc(sgmm_ctx* c, bool dense, uint64_t dense_f0, uint64_t dense_nframes, double* d_scores,
                            uint32_t* d_argmin) {
  using namespace sgmm_tc;
  if (!c->tc_ok) SGMM_FAIL(c, "tensor-core scorer not available for this model shape");
  const uint32_t KP = c->tc_KP, ks = c->tc_ksteps;
  // two frame halves per CTA when at least two B stages still fit next to them
  int nhalf = (c->tc_halves == 1 || tc_smem(ks, 2, 2) > kSmemLimit) ? 1 : 2;
  int nstage = 1;
  while (nstage < sgmm_tc::kMaxStage && tc_smem(ks, nhalf, nstage + 1) <= kSmemLimit) nstage++;
  const uint32_t rows_per_task = 128u * nhalf;
  // A operand (once per feature upload / model shape)
  if (!c->tc_A_valid) {
    SGMM_CHECK(c, c->d_Ahi.reserve((size_t)c->n_frames * KP));
    SGMM_CHECK(c, c->d_Alo.reserve((size_t)c->n_frames * KP));
    SGMM_CHECK(c, c->d_xnorm.reserve(c->n_frames));
    const unsigned grid = (unsigned)((c->n_frames + 7) / 8);
    k_tc_prep<<<grid, 256, 0, c->stream>>>(c->d_x.p, c->n_frames, c->dim, c->tc_full, KP, c->d_Ahi.p, c->d_Alo.p,
                                           c->d_xnorm.p);
    SGMM_LAUNCH_CHECK(c);
    c->tc_A_valid = true;
  }
  const uint64_t n_cells = dense ? dense_nframes * c->n_states : c->E_elems;
  const uint64_t n_tasks = dense ? (dense_nframes + rows_per_task - 1) / rows_per_task
                                 : (nhalf == 2 ? c->n_tasks256 : c->n_tasks128);
  if (n_tasks == 0) return 0;
  SGMM_CHECK(c, c->d_cb1.reserve(n_cells)); SGMM_CHECK(c, c->d_cb2.reserve(n_cells));
  SGMM_CHECK(c, c->d_ci1.reserve(n_cells));
  const size_t cap = (size_t)std::max<uint64_t>(1024, n_cells / 8);
  SGMM_CHECK(c, c->d_rescan.reserve(cap + 1));
  c->rescan_cap = cap;
  SGMM_CHECK(c, cudaMemsetAsync(c->d_rescan.p, 0, sizeof(unsigned long long), c->stream));

  CUtensorMap mAhi, mAlo, mBhi, mBlo;
  if (make_map(c, &mAhi, c->d_Ahi.p, c->n_frames, ks) || make_map(c, &mAlo, c->d_Alo.p, c->n_frames, ks) ||
      make_map(c, &mBhi, c->d_Bhi.p, c->tc_Dpad, ks) || make_map(c, &mBlo, c->d_Blo.p, c->tc_Dpad, ks))
    return 1;

  TcParams p{};
  p.ksteps = ks; p.nstage = (uint32_t)nstage; p.col_off = c->d_col_off.p; p.n_tasks = n_tasks;
  p.tasks = nhalf == 2 ? c->d_tasks256.p : c->d_tasks128.p; p.frame_off = c->d_frame_off.p; p.uniq = c->d_uniq.p; p.uniq_off = c->d_uniq_off.p;
  p.nuniq_pad = c->d_nuniq_pad.p; p.E_off = c->d_E_off.p;
  p.dense = dense ? 1 : 0; p.f0 = dense_f0; p.nframes = dense_nframes; p.n_states = c->n_states;
  p.cb1 = c->d_cb1.p; p.cb2 = c->d_cb2.p; p.ci1 = c->d_ci1.p;
  const size_t smem = tc_smem(ks, nhalf, nstage);
  cudaEventRecord(c->ev[2 * SGMM_T_SCORE_FILTER], c->stream);
  if (nhalf == 2) {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_tc_gemm<2><<<(unsigned)n_tasks, 320, smem, c->stream>>>(mAhi, mAlo, mBhi, mBlo, p);
  } else {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_tc_gemm<1><<<(unsigned)n_tasks, 192, smem, c->stream>>>(mAhi, mAlo, mBhi, mBlo, p);
  }
  SGMM_LAUNCH_CHECK(c);
  cudaEventRecord(c->ev[2 * SGMM_T_SCORE_FILTER + 1], c->stream);
  c->stage_ms[SGMM_T_SCORE_FILTER] = -1.f;

  RefineParams r{};
  r.m.meanT = c->d_meanT.p; r.m.invT = c->d_invT.p; r.m.mean_tail = c->d_mean_tail.p; r.m.inv_tail = c->d_inv_tail.p;
  r.m.norm = c->d_norm.p; r.m.logw = c->d_logw.p; r.m.dim = c->dim; r.m.npair = c->npair; r.m.Dtot = c->Dtot;
  r.m.meanR = c->d_meanR.p; r.m.invU = c->d_invU.p; r.m.var_of = c->d_var_of.p; r.m.dimP = c->dimP;
  r.state_off = c->d_state_off.p; r.x = c->d_x.p; r.cb1 = c->d_cb1.p; r.cb2 = c->d_cb2.p; r.ci1 = c->d_ci1.p;
  r.xnorm = c->d_xnorm.p; r.sguard = c->d_sguard_tc.p; r.eps = c->tc_eps; r.n_cells = n_cells;
  r.dense = dense ? 1 : 0; r.f0 = dense_f0; r.n_states = c->n_states; r.n_utts = c->n_utts;
  r.frame_off = c->d_frame_off.p; r.E_off = c->d_E_off.p; r.uniq = c->d_uniq.p; r.uniq_off = c->d_uniq_off.p;
  r.nuniq_pad = c->d_nuniq_pad.p;
  r.out = dense ? d_scores : c->d_E.p; r.argmin = dense ? d_argmin : nullptr;
  r.amin16 = dense ? nullptr : c->d_Amin.p;
  r.rescan = c->d_rescan.p; r.cap = cap; r.fallback_count = c->d_fallbacks.p;
  cudaEventRecord(c->ev[2 * SGMM_T_SCORE_REFINE], c->stream);
  k_tc_refine<<<(unsigned)((n_cells + 255) / 256), 256, 0, c->stream>>>(r);
  SGMM_LAUNCH_CHECK(c);
  k_tc_rescan<<<148 * 4, 256, 0, c->stream>>>(r);
  SGMM_LAUNCH_CHECK(c);
  cudaEventRecord(c->ev[2 * SGMM_T_SCORE_REFINE + 1], c->stream);
  c->stage_ms[SGMM_T_SCORE_REFINE] = -1.f;
  return 0;
}
