// K1, tensor-core form: the max-approx GMM scores as the dense contraction
//     [frames x K] . [K x densities],   K = 2d+1  (x^2 | x | 1)  or  d+1 (x | 1) when variances are pooled
// per state (the x^2 block is then common to all densities of a state and cannot change the argmin),
// on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with error-compensated 3xTF32 operands
// (A = Ahi + Alo, B = Bhi + Blo; D = Ahi.Bhi + Alo.Bhi + Ahi.Blo accumulated in fp32 in TMEM).
//
// What shapes the kernel is that K is SHORT (5 k-steps of 8 at d = 33 pooled): an MMA of 128 x 128 x 8 runs 64
// cycles and, with both operands in shared memory, reads 8 KB for it = the whole 128 B/cycle of the SM's shared
// memory, leaving nothing for the TMA writes that refill the B ring (measured: loads and MMAs serialised).  So
//   * A (the CTA's frames, reused by every tile) lives in TENSOR MEMORY: written once with tcgen05.st, read by
//     tcgen05.mma in its A-from-TMEM form; shared memory only carries B (64 B/cycle of reads);
//   * one CTA owns 256 frames = two halves of 128 TMEM lanes that share every B tile (halves the L2 -> shared
//     traffic per flop); the two halves' accumulators are each other's double buffer: while the tensor pipe
//     works on one half, the epilogue warps of the other half reduce theirs;
//   * B tiles are pre-tiled and pre-swizzled on the host (sgmm_set_model) so a ring stage is one contiguous
//     cp.async.bulk;
//   * the epilogue is two passes over the accumulator tile kept in TMEM: pass 1 the plain minimum (FMNMX, ALU
//     pipe), pass 2 counts and locates every column within the certified error bound of it with saturating FMAs
//     (FMA pipe): c = sat((min + G - v) * 2^100) in {0, 1}, packed sums of c * (1 + block / 4096) (TcCount).  count == 1
//     exactly <=> the minimum is the only candidate below min + G, and the fraction is then its column.
//     FSETP/FSEL-based top-2 tracking measured 2x slower.
//
// Like the FFMA scorer this kernel is a FILTER: k_tc_refine evaluates the certified winner's score in f64 in
// the reference's exact operation order (Mixtures.cpp:645-713) and queues uncertified cells for a full f64
// rescan of the state (k_tc_rescan).  Results are therefore bit-identical to the reference; the tensor cores
// only decide where f64 work goes.
#include <cuda.h>
#include <algorithm>
#include <vector>

#include <cuda_fp16.h>

#include <cstring>

#include <mutex>

#include "sgmm_internal.cuh"

namespace sgmm_tc {

constexpr int BM = 128;        // frames per CTA (TMEM lanes)
constexpr int BN = 128;        // densities per B stage / MMA N
constexpr int kMaxStage = 6;   // B stages in the TMA ring (the launcher picks what shared memory allows)

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
#ifndef SGMM_TC_TWOPASS_REF
#define SGMM_TC_TWOPASS_REF 0
#endif
constexpr bool kTwoPassRef = SGMM_TC_TWOPASS_REF != 0;   // keep the round-1 two-pass epilogue as an A/B reference (tc_debug & 128)
#ifndef SGMM_TC_PROF
#define SGMM_TC_PROF 0
#endif
constexpr bool kProf = SGMM_TC_PROF != 0;   // k_tc_gemm256: in-kernel cycle counters (tc_debug & 32) and the timing experiments (tc_debug & 1);
                                            // build with -DSGMM_TC_PROF=1 for tools/tc_prof.py — off by default for the same reason
constexpr uint32_t kUncertain = 0x80000000u;   // verdict: no certified winner -> rescan
constexpr uint32_t kPair = 0x40000000u;        // verdict: kPair | column a | column b << 15 — the argmin is one of the two
// suspend-time hint: a waiting warp sleeps in hardware until the phase completes instead of re-polling (polling loops
// were ~23 % of all issued instructions and compete with the MMA-issuing warp for its scheduler)
constexpr uint32_t kWaitHintNs = 20000;
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity), "r"(kWaitHintNs) : "memory");
}
// busy poll (no suspend): lower wake-up latency for the MMA <-> epilogue hand-offs on the critical path
__device__ __forceinline__ void mbar_spin(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "SPIN_%=:\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra SDONE_%=;\n\t"
      "bra SPIN_%=;\n\t"
      "SDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared (UBLKCP), completion on the mbarrier
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem desc]   (kind::tf32, fp32 accumulate; A: lane = row, one 32-bit column per k)
// Called by the WHOLE (converged) MMA warp with warp-uniform operands; one elected lane issues.  Issuing from
// inside an `if (lane == 0)` region instead makes the compiler wrap every MMA in a uniform-register "waterfall"
// loop (ELECT / R2UR / BRA.U.ANY, ~15 dependent instructions): measured 158 cycles per 64-cycle MMA.
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred pe, pa;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, pa;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// same, both operands from shared memory (used for the Alo pass only)
__device__ __forceinline__ void tc_mma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred pe, pa;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, pa;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// One lane of the converged warp (elect.sync): the caller issues a whole unit's MMAs and commits inside
// `if (tc_elect_one()) { ... }`, so the election and the uniform-register moves are paid once per unit instead of
// once per MMA (the per-MMA form compiled to ~33 SASS instructions per MMA and left the tensor pipe ~50 % idle).
__device__ __forceinline__ uint32_t tc_elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, pe;\n\t}" : "=r"(pred));
  return pred;
}
__device__ __forceinline__ void tc_mma_ts_1(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred pa;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, pa;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tc_mma_ss_1(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred pa;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, pa;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// kind::f16 (A, B fp16, K = 16 per instruction = one 32-byte atom row, fp32 accumulate)
__device__ __forceinline__ void tc_mma_ss_f16_1(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred pa;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, pa;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tc_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
// 8 consecutive columns of this thread's TMEM lane <- registers
__device__ __forceinline__ void tc_st8(uint32_t taddr, const float4& a, const float4& b) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(__float_as_uint(a.x)),
               "r"(__float_as_uint(a.y)), "r"(__float_as_uint(a.z)), "r"(__float_as_uint(a.w)), "r"(__float_as_uint(b.x)),
               "r"(__float_as_uint(b.y)), "r"(__float_as_uint(b.z)), "r"(__float_as_uint(b.w))
               : "memory");
}
// 16 consecutive accumulator columns of this thread's TMEM lane (no wait: the caller overlaps the latency)
__device__ __forceinline__ void tc_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
// wait for the outstanding tcgen05.ld's; the registers about to be consumed are operands so that the compiler
// cannot move their uses above the wait
__device__ __forceinline__ void tc_ld_wait(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}

// Pass 2 of the epilogue: count and locate the accumulator columns below thr, 16 columns per call.
// c = sat((thr - v) * 2^100) is EXACTLY 1 for v < thr and EXACTLY 0 for v >= thr (the products are exact
// power-of-two scalings; for 1e-20 < |thr| < 1e8 two different floats around thr differ by >= 2^-91, so a positive
// difference saturates; the caller guards that range).  Column e = 8 h + 2 q + l of block b adds c * W[h] to lane l
// of the packed accumulator A[q], W[h] = 1 + (2 b + h) / 4096, with one FFMA2 (fma.rn.f32x2) per two columns:
// 1.5 FMA-pipe instructions per column.  All sums are exact (count <= 128, 12 fraction bits).  At the end
// sum A = count + ..., and when count == 1 the fraction names (b, h) and the one non-zero accumulator names (q, l).
struct TcCount {
  unsigned long long A[4], W[2];
  __device__ __forceinline__ static unsigned long long pack(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
  }
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int q = 0; q < 4; q++) A[q] = 0ull;
    W[0] = pack(1.0f, 1.0f);
    W[1] = pack(1.0f + 1.0f / 4096.0f, 1.0f + 1.0f / 4096.0f);
  }
  __device__ __forceinline__ void block16(const uint32_t (&r)[16], float nbig, float T) {
#pragma unroll
    for (int pq = 0; pq < 8; pq++) {
      const float c0 = __saturatef(fmaf(__uint_as_float(r[2 * pq]), nbig, T));
      const float c1 = __saturatef(fmaf(__uint_as_float(r[2 * pq + 1]), nbig, T));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(A[pq & 3]) : "l"(pack(c0, c1)), "l"(W[pq >> 2]));
    }
    const unsigned long long step = pack(2.0f / 4096.0f, 2.0f / 4096.0f);
    asm("add.rn.f32x2 %0, %0, %1;" : "+l"(W[0]) : "l"(step));
    asm("add.rn.f32x2 %0, %0, %1;" : "+l"(W[1]) : "l"(step));
  }
  // acc = number of columns below thr; idx = the column (relative to the first block) when acc == 1
  __device__ __forceinline__ void finish(float& acc, float& idx) const {
    float a[8];
#pragma unroll
    for (int q = 0; q < 4; q++) asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * q]), "=f"(a[2 * q + 1]) : "l"(A[q]));
    const float total = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    float ql = a[1];
#pragma unroll
    for (int e = 2; e < 8; e++) ql = fmaf(a[e], (float)e, ql);
    acc = floorf(total);
    idx = fmaf((total - acc) * 4096.0f, 8.0f, rintf(ql));
  }
};

// compiler-only fence for registers filled by an earlier tcgen05.ld whose wait was issued on another block: pins their
// uses behind that wait (asm volatile statements keep their order) at zero instructions
__device__ __forceinline__ void tc_ld_pin(uint32_t (&r)[16]) {
  asm volatile(""
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}

// ONE pass over tensor memory for accumulator columns [c_lo, c_hi) of one TMEM lane (= frame row).  The columns are taken
// in sub-parts of NB 16-column blocks that are loaded into registers together; the minimum (FMNMX) and then the count /
// locate step (TcCount) run on those registers, and the sub-parts merge like tiles: the result is certain (acc == 1)
// iff exactly one column of the whole range lies below min + G — the same condition as the two-pass scan, with half
// its tcgen05.ld traffic.  The accumulator is handed back (release_bar) as soon as the LAST load has landed, before the
// arithmetic on it.
// One sub-part of the one-pass scan: NB 16-column blocks already in registers -> its minimum ms, the certified bound
// Gsub = Gs + 4e-6 |ms|, and (a_s, i_s) = number of columns below ms + Gsub and, when that is 1, the column (relative to
// the sub-part).  Pipe budget (B300_MICROARCH.md: FMA / ALU pipes issue one warp instruction per 2 cycles, the immediate
// form of FFMA one per cycle): per column 0.5 FMNMX3 on the ALU pipe and TWO immediate-form FFMAs on the FMA pipe —
//   c = sat(v * -2^100 + thr * 2^100)   exactly 1 below thr, exactly 0 otherwise (see TcCount), and
//   a[b] += c * 2^e                      column e of block b sets bit e of an exact integer < 2^16,
// so the candidates of the sub-part end up as a bit mask: count = popc, column = ffs.  (The packed FFMA2 form this
// replaces kept the FMA pipe as busy as the tensor pipe: the epilogue, not the MMAs, set the kernel's pace.)
template <int NB, bool PAIR = false>   // PAIR: also the packed candidate code of TcBest
__device__ __forceinline__ void tc_subpart(const uint32_t (&v)[NB][16], float Gs, float& ms, float& Gsub, float& a_s,
                                           float& i_s, uint32_t* packed = nullptr) {
  float mm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
#pragma unroll
  for (int b = 0; b < NB; b++) {
#pragma unroll
    for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(v[b][e]));
  }
  ms = fminf(fminf(mm[0], mm[1]), fminf(mm[2], mm[3]));
  Gsub = Gs + 4e-6f * fabsf(ms);
  const float thr = ms + Gsub, nbig = -0x1p100f, T = thr * 0x1p100f;
  const bool thr_ok = fabsf(thr) > 1e-20f && fabsf(thr) < 1e8f;   // else c may be fractional: uncertain
  float a[NB];
#pragma unroll
  for (int b = 0; b < NB; b++) a[b] = 0.f;
#pragma unroll
  for (int e = 0; e < 16; e++) {
#pragma unroll
    for (int b = 0; b < NB; b++) {
      const float c = __saturatef(fmaf(__uint_as_float(v[b][e]), nbig, T));
      a[b] = fmaf(c, (float)(1u << e), a[b]);
    }
  }
  // a[b] is an integer in [0, 65535]: a[b] + 2^23 carries it in the low mantissa bits
  uint32_t mk[4] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int b = 0; b < NB; b++) mk[b] = __float_as_uint(a[b] + 8388608.f);
  const uint32_t lo = __byte_perm(mk[0], mk[1], 0x5410), hi = NB > 2 ? __byte_perm(mk[2], mk[3], 0x5410) : 0u;
  const int n = __popc(lo) + __popc(hi);
  const int pos = lo ? __ffs(lo) - 1 : 31 + __ffs(hi);
  a_s = thr_ok ? (float)n : 2.f;
  i_s = (float)pos;
  if (PAIR) {   // packed for TcBest: count code << 30 | (last - first candidate column) << 16 | first candidate column
    uint32_t code = (thr_ok && n == 1) ? (1u << 30) | (uint32_t)pos : (3u << 30) | (uint32_t)pos;
    if (__any_sync(0xffffffffu, thr_ok && n == 2)) {   // rare, warp-uniform branch: the second candidate's column
      const int last = hi ? 63 - __clz(hi) : 31 - __clz(lo | 1u);
      if (thr_ok && n == 2) code = (2u << 30) | ((uint32_t)(last - pos) << 16) | (uint32_t)pos;
    }
    *packed = code;
  }
}

// ---- candidates of a (frame, state) cell across sub-parts, column parts and tiles (default epilogue of k_tc_gemm256).
// Every sub-result is {minimum m, bound G = Gs + 4e-6 |m|, acc = columns below m + G, their first (idx) and last (idxP)
// column}.  A column of ANOTHER sub-result S can lie below the best one's threshold a + G_a only if it is in S's own
// count (a <= m_S and |G_a - G_S| <= 4e-6 |a - m_S| give a + G_a <= m_S + G_S), so the cell's candidate set is the union
// of the counts of the sub-results whose minimum is <= a + G_a.  Kept: the best sub-result, the runner-up (with its
// count and column) and the third-smallest minimum.  Verdict:
//   certified      runner-up > a + G_a and acc == 1
//   pair           runner-up > a + G_a and acc == 2 (both in one sub-result), or
//                  third > a + G_a >= runner-up and both counts are 1 (one candidate each)       [round 2]
//   uncertain      otherwise
// A pair verdict names the ONLY two columns that can be the f64 argmin: the refinement evaluates both exactly (first
// minimum wins), and the cell never reaches the rescan queue.
struct TcBest {
  // 5 registers (as many as the certified-only bookkeeping it replaces; anything more spills inside the column loops and
  // costs the filter 50 %, measured): three minima and two packed candidate codes
  //   code = count << 30 | (last - first candidate column) << 16 | first candidate column      count: 1, 2, 3 = more / unknown
  float m, m2, m3;
  uint32_t c1, c2;
  __device__ __forceinline__ void init() { m = m2 = m3 = INFINITY; c1 = c2 = 0xc0000000u; }
  // (selects only: a data-dependent branch here diverges the 32 rows of an epilogue warp — measured +0.5 ms on the filter)
  __device__ __forceinline__ void ins(float nm, uint32_t nc) {
    const bool b1 = nm < m, b2 = nm < m2;                     // m <= m2, so b1 implies b2
    const float om = m, om2 = m2;
    const uint32_t oc1 = c1;
    m3 = b2 ? om2 : fminf(m3, nm);
    m2 = b1 ? om : (b2 ? nm : om2);
    c2 = b1 ? oc1 : (b2 ? nc : c2);
    m = b1 ? nm : om;
    c1 = b1 ? nc : oc1;
  }
  // all of another TcBest whose columns start at column off (off + column < 65536): both triples are sorted, so the
  // merged one is  winner's m | min(winner's m2, loser's m) | min(max(those two), winner's m3, loser's m2)
  __device__ __forceinline__ void merge(const TcBest& o, uint32_t off) {
    const bool ob = o.m < m;
    const float wm2 = ob ? o.m2 : m2, wm3 = ob ? o.m3 : m3, lm = ob ? m : o.m, lm2 = ob ? m2 : o.m2;
    const uint32_t wc1 = ob ? o.c1 + off : c1, wc2 = ob ? o.c2 + off : c2, lc1 = ob ? c1 : o.c1 + off;
    const bool s2 = lm < wm2;
    m = ob ? o.m : m;
    c1 = wc1;
    m2 = s2 ? lm : wm2;
    c2 = s2 ? lc1 : wc2;
    m3 = fminf(fminf(s2 ? wm2 : lm, wm3), lm2);
  }
  // called by a CONVERGED warp: the pair logic sits behind a warp-uniform vote (all 32 cells certified: the common case)
  __device__ __forceinline__ uint32_t verdict(float Gs) const {
    const float thr = m + (Gs + 4e-6f * fabsf(m));          // the best sub-result's own bound (tc_subpart)
    const uint32_t n1 = c1 >> 30, a = c1 & 0xffffu;
    const bool alone = m2 > thr, cert = alone && n1 == 1u;
    uint32_t vd = cert ? a : kUncertain;
    if (!__all_sync(0xffffffffu, cert)) {
      const bool pa = alone && n1 == 2u, pb = !alone && m3 > thr && n1 == 1u && (c2 >> 30) == 1u;
      const uint32_t b = pa ? a + ((c1 >> 16) & 0x3fffu) : (c2 & 0xffffu);
      const bool okp = (pa || pb) && a != b && a < 32768u && b < 32768u;
      vd = cert ? a : (okp ? (kPair | min(a, b) | (max(a, b) << 15)) : kUncertain);
    }
    return vd;
  }
};

template <int NB, bool SINGLE = false>   // SINGLE: the caller guarantees c_hi - c_lo <= 16 * NB (one sub-part, no loop)
__device__ __forceinline__ void tc_scan_onepass(uint32_t tcol, uint32_t c_lo, uint32_t c_hi, float Gs, uint64_t* release_bar,
                                                int lane, uint32_t dbg, float& m, float& acc, float& idx, float& G) {
  // dbg & 1 (timing experiment, results invalid): no loads, no arithmetic
  m = INFINITY; acc = 0.f; idx = 0.f; G = 0.f;
  bool released = false;
  if (!(dbg & 1)) {
    for (uint32_t s_lo = c_lo; s_lo < c_hi; s_lo += 16 * NB) {
      uint32_t v[NB][16];
      const bool last = SINGLE || s_lo + 16 * NB >= c_hi;          // this warp's last read of the accumulator
      if (s_lo + 16 * NB <= c_hi) {                                 // full sub-part (the common case): no padding
#pragma unroll
        for (int b = 0; b < NB; b++) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
      } else {
#pragma unroll
        for (int b = 0; b < NB; b++) {
          if (s_lo + 16 * b < c_hi) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
          else {
#pragma unroll
            for (int e = 0; e < 16; e++) v[b][e] = 0x7f800000u;   // +inf: never the minimum, indicator 0
          }
        }
      }
      tc_ld_wait(v[0]);
#pragma unroll
      for (int b = 1; b < NB; b++) tc_ld_pin(v[b]);
      if (last) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(release_bar);
        released = true;
      }
      float ms, Gsub, a_s, i_s;
      tc_subpart<NB>(v, Gs, ms, Gsub, a_s, i_s);
      i_s += (float)s_lo;
      if (SINGLE) { acc = a_s; idx = i_s; G = Gsub; m = ms; break; }
      // merge: every other sub-part whose minimum lies below the best one's threshold is one more candidate.  (A sub-part
      // that was "close" to an earlier best and is forgotten when a better one arrives cannot matter: the earlier best is
      // at least as close to the new one, and is counted.)
      if (s_lo == c_lo || ms < m) {
        acc = a_s + ((s_lo != c_lo && m <= ms + Gsub) ? 1.f : 0.f);
        idx = i_s; G = Gsub; m = ms;
      } else if (ms <= m + G) acc += 1.f;
    }
  }
  if (!released) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(release_bar);
  }
}

// ---- sum scores on the tensor path (option sum_tc): instead of a verdict, the epilogue writes for every
// 64-column sub-part of a cell {candidate mask (64 bits), the state's minimum gm, the width W}: bit e is set iff
// accumulator column e lies below gm + W + 4e-6 |gm|, W = 45 (in accumulator units) + twice the certified error bound.
// By the bound the mask is a SUPERSET of the densities within 45 of the best exact score —
// the ones whose exp(-score) can change the f64 sum (k_sum_pruned's criterion); k_tc_refine_sm<.., SUM> evaluates them.
// A threshold outside the exact-indicator range (NaN / inf rows) selects every column.
template <int NB>
__device__ __forceinline__ void tc_subpart_mask(const uint32_t (&v)[NB][16], float W, float ms, uint32_t& lo, uint32_t& hi) {
  // ms: the minimum the threshold hangs on (the whole state's, exchanged between the column parts before this pass)
  const float thr = ms + (W + 4e-6f * fabsf(ms)), nbig = -0x1p100f, T = thr * 0x1p100f;
  const bool thr_ok = fabsf(thr) > 1e-20f && fabsf(thr) < 1e8f;
  float a[NB];
#pragma unroll
  for (int b = 0; b < NB; b++) a[b] = 0.f;
#pragma unroll
  for (int e = 0; e < 16; e++) {
#pragma unroll
    for (int b = 0; b < NB; b++) {
      const float c = __saturatef(fmaf(__uint_as_float(v[b][e]), nbig, T));
      a[b] = fmaf(c, (float)(1u << e), a[b]);
    }
  }
  uint32_t mk[4] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int b = 0; b < NB; b++) mk[b] = __float_as_uint(a[b] + 8388608.f);
  lo = __byte_perm(mk[0], mk[1], 0x5410);
  hi = NB > 2 ? __byte_perm(mk[2], mk[3], 0x5410) : 0u;
  if (!thr_ok) { lo = 0xffffffffu; hi = 0xffffffffu; }
}

// mask mode, pass 1: the minimum of accumulator columns [c_lo, c_hi) of this thread's row
template <int NB>
__device__ __forceinline__ float tc_part_min(uint32_t tcol, uint32_t c_lo, uint32_t c_hi) {
  float mm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
  for (uint32_t s_lo = c_lo; s_lo < c_hi; s_lo += 16 * NB) {
    uint32_t v[NB][16];
#pragma unroll
    for (int b = 0; b < NB; b++) {
      if (s_lo + 16 * b < c_hi) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
      else {
#pragma unroll
        for (int e = 0; e < 16; e++) v[b][e] = 0x7f800000u;
      }
    }
    tc_ld_wait(v[0]);
#pragma unroll
    for (int b = 1; b < NB; b++) tc_ld_pin(v[b]);
#pragma unroll
    for (int b = 0; b < NB; b++) {
#pragma unroll
      for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(v[b][e]));
    }
  }
  return fminf(fminf(mm[0], mm[1]), fminf(mm[2], mm[3]));
}

// mask mode, pass 2: columns [c_lo, c_hi) of the state's ONLY tile against the state's minimum gm; out = the cell's four
// entries (sub-part = columns / 64), or NULL for rows / columns that are not stored
template <int NB>
__device__ __forceinline__ void tc_scan_masks(uint32_t tcol, uint32_t c_lo, uint32_t c_hi, float W, float gm,
                                              uint64_t* release_bar, int lane, uint4* out) {
  bool released = false;
  for (uint32_t s_lo = c_lo; s_lo < c_hi; s_lo += 16 * NB) {
    uint32_t v[NB][16];
    const bool last = s_lo + 16 * NB >= c_hi;
#pragma unroll
    for (int b = 0; b < NB; b++) {
      if (s_lo + 16 * b < c_hi) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
      else {
#pragma unroll
        for (int e = 0; e < 16; e++) v[b][e] = 0x7f800000u;   // +inf: never below a threshold
      }
    }
    tc_ld_wait(v[0]);
#pragma unroll
    for (int b = 1; b < NB; b++) tc_ld_pin(v[b]);
    if (last) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(release_bar);
      released = true;
    }
    uint32_t lo, hi;
    tc_subpart_mask<NB>(v, W, gm, lo, hi);
    if (out) out[s_lo / (16 * NB)] = make_uint4(lo, hi, __float_as_uint(gm), __float_as_uint(W));
  }
  if (!released) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(release_bar);
  }
}

// tc_scan_onepass with the candidate bookkeeping of TcBest (columns relative to the tile)
template <int NB>
__device__ __forceinline__ void tc_scan_onepass2(uint32_t tcol, uint32_t c_lo, uint32_t c_hi, float Gs, uint64_t* release_bar,
                                                 int lane, uint32_t dbg, TcBest& st) {
  st.init();
  bool released = false;
  if (!(dbg & 1)) {
    for (uint32_t s_lo = c_lo; s_lo < c_hi; s_lo += 16 * NB) {
      uint32_t v[NB][16];
      const bool last = s_lo + 16 * NB >= c_hi;                     // this warp's last read of the accumulator
      if (s_lo + 16 * NB <= c_hi) {                                 // full sub-part (the common case): no padding
#pragma unroll
        for (int b = 0; b < NB; b++) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
      } else {
#pragma unroll
        for (int b = 0; b < NB; b++) {
          if (s_lo + 16 * b < c_hi) tc_ld16_nowait(tcol + s_lo + 16 * b, v[b]);
          else {
#pragma unroll
            for (int e = 0; e < 16; e++) v[b][e] = 0x7f800000u;   // +inf: never the minimum, indicator 0
          }
        }
      }
      tc_ld_wait(v[0]);
#pragma unroll
      for (int b = 1; b < NB; b++) tc_ld_pin(v[b]);
      if (last) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(release_bar);
        released = true;
      }
      float ms, Gsub, a_s, i_s;
      uint32_t code;
      tc_subpart<NB, true>(v, Gs, ms, Gsub, a_s, i_s, &code);
      if (s_lo == c_lo) { st.m = ms; st.c1 = code + s_lo; }      // (st.init() left the rest empty)
      else st.ins(ms, code + s_lo);
    }
  }
  if (!released) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(release_bar);
  }
}

// The two passes of the epilogue over accumulator columns [c_lo, c_hi) of one TMEM lane (= frame row): the minimum
// m, the certified bound G = Gs + 4e-6 |m|, and (acc, idx) = number of columns below m + G and — when that is 1 —
// the column.  Loads run one 16-column block ahead.
__device__ __forceinline__ void tc_scan_columns(uint32_t tcol, uint32_t c_lo, uint32_t c_hi, float Gs, float& m,
                                                float& acc, float& idx, float& G) {
  uint32_t ra[16], rb[16];
#pragma unroll
  for (int e = 0; e < 16; e++) rb[e] = 0;
  float mm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
  tc_ld16_nowait(tcol + c_lo, ra);
  tc_ld_wait(ra);
  for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
    if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
#pragma unroll
    for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(ra[e]));
    tc_ld_wait(rb);
    if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
    if (cc + 16 < c_hi) {
#pragma unroll
      for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(rb[e]));
    }
    tc_ld_wait(ra);
  }
  m = fminf(fminf(mm[0], mm[1]), fminf(mm[2], mm[3]));
  G = Gs + 4e-6f * fabsf(m);
  const float thr = m + G, nbig = -0x1p100f, T = thr * 0x1p100f;
  const bool thr_ok = fabsf(thr) > 1e-20f && fabsf(thr) < 1e8f;   // else c may be fractional: uncertain
  TcCount cnt;
  cnt.init();
  tc_ld16_nowait(tcol + c_lo, ra);
  tc_ld_wait(ra);
  for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
    if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
    cnt.block16(ra, nbig, T);
    tc_ld_wait(rb);
    if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
    if (cc + 16 < c_hi) cnt.block16(rb, nbig, T);
    tc_ld_wait(ra);
  }
  cnt.finish(acc, idx);
  idx += (float)c_lo;
  if (!thr_ok) acc = 2.f;
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

// D = f32, A = B = f16, both K-major
__host__ __device__ inline uint32_t make_idesc_f16(uint32_t M, uint32_t N) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
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

// A operand builder of the fp16 form (pooled models): row f = s_f * [x | 1 | 0...], s_f = the power of two that
// brings max(|x|, 1) of the row into [2^14, 2^15), each entry split into fp16 hi + fp16 lo (11 + 11 bits; the lo
// part of an entry below 2^-3 is subnormal: exact to 2^-25 if the tensor core keeps subnormals, and what would be
// lost if it flushed them is added to the row's norm, so the certified bound holds either way).  The row scale is
// common to all columns of the row and the state's column scale (sgmm_set_model) to all candidates of a (frame,
// state) cell, so neither changes which column is the minimum; the epilogue's bound is linear in both.
// xnorm[f] = {||s_f x|| (rounded up) + flush / (2 eps), s_f}.  A row with a non-finite feature becomes NaN: every
// cell of that frame is then "uncertain" and rescanned in f64.
__global__ void k_tc_prep16(const float* __restrict__ x, uint64_t n_frames, uint32_t dim, uint32_t KH, uint32_t rem16,
                            float inv2eps, __half* __restrict__ Ahi, __half* __restrict__ Alo, float2* __restrict__ xnorm) {
  const uint64_t f = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (f >= n_frames) return;
  const float* xf = x + f * dim;
  float mx = 1.f;
  bool bad = false;
  for (uint32_t k = lane; k < dim; k += 32) {
    const float v = fabsf(xf[k]);
    bad |= !(v < 1048576.f);                       // (and the row scale stays a normal fp16 number)
    mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  bad = __any_sync(0xffffffffu, bad);
  int e;
  frexpf(mx, &e);                                  // mx = m * 2^e, m in [0.5, 1)  ->  mx * 2^(15 - e) in [2^14, 2^15)
  const float sf = ldexpf(1.f, 15 - e);
  // Destination position p of the row <- term k of [s x | s | 0 ..], hi or lo part.  Unpacked: p = k in both rows.  Packed
  // remainder (rem16 = r > 0, K = 16 q + r): the hi row's last k-step holds [hi r | lo r | hi r] of terms 16 q .. K - 1
  // (against [Bhi r | Bhi r | Blo r], sgmm_internal.cuh) and the lo row's last k-step is unused.
  const uint32_t K = dim + 1, kfull = K - rem16;
  float s2 = 0.f, fl = 0.f;
  for (uint32_t pos = lane; pos < KH; pos += 32) {
    uint32_t k = pos, sel_hi = 0;                  // sel_hi: which part lands in the hi row at pos (0 hi, 1 lo)
    bool primary = true, lo_used = true;           // primary: the term's one position that feeds the norms
    if (rem16 && pos >= kfull) {
      const uint32_t j = pos - kfull, seg = j / rem16;
      k = seg < 3 ? kfull + (j - seg * rem16) : K;         // K: a zero term
      sel_hi = seg == 1 ? 1u : 0u;
      primary = seg == 0;
      lo_used = false;
    }
    float a = 0.f;
    if (k < dim) a = xf[k] * sf;                   // exact (power of two; |a| < 2^15)
    else if (k == dim) a = sf;
    __half hi = __float2half_rn(a);
    const float r = a - __half2float(hi);          // exact
    __half lo = __float2half_rn(r);
    if (primary) {
      if (k < dim) s2 += a * a;
      // subnormal parts (below 2^-14): rounded to 2^-25 if kept, lost entirely if the tensor core flushes them
      if (r != 0.f && fabsf(r) < 6.103515625e-05f) fl += fmaxf(fabsf(r), 2.98023223876953125e-08f);
      if (a != 0.f && fabsf(a) < 6.103515625e-05f) fl += fabsf(a);
    }
    if (bad) { hi = __float2half_rn(nanf("")); lo = hi; }
    Ahi[f * KH + pos] = sel_hi ? lo : hi;
    Alo[f * KH + pos] = lo_used ? lo : __float2half_rn(0.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s2 += __shfl_xor_sync(0xffffffffu, s2, o); fl += __shfl_xor_sync(0xffffffffu, fl, o); }
  if (lane == 0) xnorm[f] = make_float2(sqrtf(s2) * 1.0001f + fl * inv2eps, sf);
}


struct TcParams {
  uint32_t ksteps;             // K steps of 8 (KP = 8 * ksteps)
  uint32_t nstage;             // B stages in the ring (<= kMaxStage)
  uint32_t stage_bytes;        // bytes of one ring stage (a full 128-density tile: 2 * ksteps * 128 * 32)
  uint32_t nhalf;              // frame halves (of 128) per task: 2 when 32 * ksteps <= 256 TMEM columns, else 1
  uint32_t a_smem;             // where the A parts live: 0 hi TMEM / lo smem, 1 both smem, 2 hi smem / lo TMEM
  uint32_t debug;              // timing experiments only (results invalid): 1 = no epilogue math, 2 = no MMAs
  uint32_t krem;               // fp16 form: the last k-step is the packed remainder (one MMA, hi slot only)
  const float* Ahi;            // [n_frames][8 * ksteps] tf32 hi / lo of the A rows (k_tc_prep)
  const float* Alo;
  uint64_t n_frames;
  const float* Bimg;           // pre-tiled, pre-swizzled B (sgmm_set_model)
  const uint32_t* tile_first;  // [n_states + 1]
  const uint2* tile_info;      // {offset / 16 bytes, N}
  const float2* xnorm;         // per frame {||x||, ||x^2||}, rounded up
  const float4* sguard;        // per state {max ||g||, max ||h||, max |c|, -}
  float eps;                   // relative error of one 3xTF32 dot-product term + fp32 accumulation (see host)
  uint64_t n_tasks;
  // gathered mode: task = {utt, t0}
  const uint4* tasks;
  const uint64_t* frame_off;
  const uint32_t* uniq;
  const uint32_t* uniq_off;
  const uint32_t* nuniq_pad;
  const uint64_t* E_off;
  // packed tail tasks (task.x & 0x80000000): rows from several utterances against the union of their state lists
  const uint4* pack_desc;            // {first entry in pack_uniq, states in the union, valid rows, 0}
  const uint32_t* pack_uniq;
  const sgmm_pack_row* pack_rows;    // [pack][128]
  const uint16_t* pack_colmap;
  // dense mode
  int dense;
  uint64_t f0, nframes;
  uint32_t n_states;
  // per cell, indexed like the emission matrix (gathered) / [frame - f0][state] (dense):
  // column of the certified minimum within the state's (padded) density list, or kUncertain
  uint32_t* cand;
  unsigned long long* prof;   // cycle profile (debug & 32), else NULL
  uint4* summask;             // mask mode (sum scores on the tensor path): [cell][4] entries, see tc_subpart_mask; else NULL
  int pairs;                  // default epilogue of k_tc_gemm256: emit pair verdicts (option tc_pairs; 0: such cells are uncertain)
  int tol;                    // tolerance mode (fp16 form, k_tc_gemm256): besides the verdict, e32[cell] = min accumulator / (s_f t_s)
  float* e32;                 // = the state's best score minus the frame's x^2 term (k_tol_finish adds it)
};
constexpr uint32_t kAccCol = 128;    // TMEM: columns [0, 128) hold Ahi of both halves, then three 128-column accumulators
constexpr uint32_t kNumAcc = 3;

// Verdict cell of (row, column): the tiled layout of sgmm_cell_rel — row_cell = first cell of the row's 16-frame tile +
// (t & 15), so that the 32 rows of a TMEM lane quarter (consecutive frames) store two full 64-byte runs per column instead
// of 32 sectors.  The dense pass uses it for its verdicts too (stride = all states, tiles counted from the pass's first
// frame); only its SCORES are row-major, as sgmm_score_all returns them.
__device__ __forceinline__ uint64_t tc_row_cell(uint64_t out_base, uint32_t row, uint32_t out_stride, bool tiled) {
  return tiled ? out_base + (uint64_t)(row >> 4) * (16u * out_stride) + (row & 15u) : out_base + (uint64_t)row * out_stride;
}
__device__ __forceinline__ uint64_t tc_cell(uint64_t row_cell, uint32_t col, bool tiled) {
  return tiled ? (row_cell & ~15ull) + (col << 4) + (((uint32_t)row_cell ^ col) & 15u) : row_cell + col;
}

// Warps: 0 = bulk-copy producer, 1 = TMEM allocator + MMA issuer, 2..17 = epilogue: warp 2 + 8 * part + 4 * half + i
// reduces columns [64 * part, +64) of frame half `half` (TMEM lane quarter = warp_id % 4).  With two valid halves, group g owns frame half g and accumulator g on
// every tile.  With one valid half (short tail of an utterance, or large K) group 0 works alone and the two
// accumulators alternate by tile.
template <int KS>   // KS > 0: k-steps known at compile time (fully unrolled MMA issue); 0: p.ksteps
__global__ void __launch_bounds__(576, 1) k_tc_gemm(TcParams p) {
  const uint32_t dbg = kProf ? p.debug : 0u;   // timing experiments (tc_debug): compiled out by default
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* sAlo = smem_raw + (base - smem_u32(smem_raw));    // [half][kstep][128 rows][32 B], 32-byte swizzle
  unsigned char* sAhi = sAlo + p.nhalf * (p.stage_bytes / 2);      // same layout (a_smem only)
  unsigned char* sB = sAhi + (p.a_smem == 1 ? p.nhalf * (p.stage_bytes / 2) : 0u);   // [stage][hi, lo][kstep][128 rows][32 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + p.nstage * p.stage_bytes);
  uint64_t* a_full = bars;                 // A rows are in TMEM (8 epilogue warps arrive)
  uint64_t* full = bars + 1;
  uint64_t* empty = full + kMaxStage;
  uint64_t* tfull = empty + kMaxStage;     // [3] accumulator complete
  uint64_t* tempty = tfull + kNumAcc;      // [3] accumulator drained
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + kNumAcc);
  float* xchg = reinterpret_cast<float*>(tmem_ptr + 4);   // [2 slots][2 halves][128 rows][5]: column-part merge

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = blockIdx.x;
  const uint32_t ROWS = BM * p.nhalf;
  const uint32_t nstage = p.nstage, ks = KS > 0 ? (uint32_t)KS : p.ksteps, KP = 8 * ks;
  constexpr uint32_t kstride = 4096 >> 4;   // descriptor units between the 128-row k-step blocks of a B operand

  // ---- task decode (uniform over the CTA)
  uint64_t row0;        // first A row (absolute frame index)
  uint32_t nvalid, nu, out_stride;
  const uint32_t* ulist = nullptr;
  const sgmm_pack_row* prow = nullptr;   // packed task: per-row frame / output cell / column map
  uint64_t out_base;
  if (p.dense) {
    const uint64_t r0 = task * ROWS;
    row0 = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)ROWS, p.nframes - r0);
    nu = p.n_states;
    out_stride = p.n_states;
    out_base = r0 * p.n_states;
  } else if (p.tasks[task].x & 0x80000000u) {
    const uint32_t id = p.tasks[task].x & 0x7fffffffu;
    const uint4 d = p.pack_desc[id];
    prow = p.pack_rows + (size_t)id * 128;
    row0 = 0; nvalid = d.z; nu = d.y; out_stride = 0; out_base = 0;
    ulist = p.pack_uniq + d.x;
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
  const uint32_t nh = nvalid > (uint32_t)BM ? 2u : 1u;   // halves that hold valid frames

  // ---- one-time setup
  if (warp == 0 && lane == 0) {
    mbar_init(a_full, 4 * nh);
    for (uint32_t i = 0; i < nstage; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (uint32_t i = 0; i < kNumAcc; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 8); }
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
    // ===== producer: one bulk copy per B tile =====
    if (lane == 0) {
      uint32_t it = 0;
      for (uint32_t j = 0; j < nu; j++) {
        const uint32_t s = ulist ? ulist[j] : j;
        for (uint32_t t = p.tile_first[s]; t < p.tile_first[s + 1]; t++, it++) {
          const uint32_t stage = it % nstage, bytes = p.stage_bytes;
          if (dbg & 16) continue;
          mbar_wait(&empty[stage], ((it / nstage) & 1) ^ 1);
          mbar_expect_tx(&full[stage], bytes);
          bulk_load(sB + stage * p.stage_bytes, p.Bimg + (size_t)t * (p.stage_bytes / 4), bytes, &full[stage]);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop, one elected lane issues =====
    const bool prof = kProf && p.prof != nullptr;
    long long w_a = 0, w_full = 0, w_tempty = 0, c0 = 0, c1 = 0;
    unsigned long long g0 = 0;
    if (prof) { c0 = clock64(); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0)); }
    mbar_wait(a_full, 0);
    if (prof) { c1 = clock64(); w_a = c1 - c0; }
    tc_fence_after();
    uint32_t it = 0;
    for (uint32_t j = 0; j < nu; j++) {
      const uint32_t s = ulist ? ulist[j] : j;
      for (uint32_t t = p.tile_first[s]; t < p.tile_first[s + 1]; t++, it++) {
        const uint32_t N = p.tile_info[t].y;
        const uint32_t stage = it % nstage;
        if (prof) c1 = clock64();
        if (!(dbg & 16)) mbar_wait(&full[stage], (it / nstage) & 1);
        if (prof) w_full += clock64() - c1;
        const uint32_t idesc = make_idesc_tf32(BM, N);
        const uint64_t bhi = make_desc_sw32(smem_u32(sB + stage * p.stage_bytes));
        const uint64_t blo = bhi + (uint64_t)ks * kstride;   // every k-step block is 128 rows (4 KB) whatever N
        for (uint32_t h = 0; h < nh; h++) {
          // (tile, half) units rotate through the three accumulators: the tensor pipe may run a unit ahead of the
          // epilogue that drains the previous unit of the same half
          const uint32_t unit = it * nh + h, buf = unit % kNumAcc, use = unit / kNumAcc;
          if (prof) c1 = clock64();
          if (!(dbg & 8)) mbar_wait(&tempty[buf], (use & 1) ^ 1);
          if (prof) w_tempty += clock64() - c1;
          tc_fence_after();
          const uint32_t d = tmem_base + kAccCol + buf * BN;
          // the A part kept in TMEM (columns [h * KP, +KP)) and the part(s) kept in shared memory, by p.a_smem:
          // 0: hi in TMEM, lo in sA0;  1: lo in sA0, hi in sA1 (all three passes A-from-shared);  2: hi in sA0, lo in TMEM
          const uint32_t a_t = tmem_base + h * KP;
          const uint64_t a_s0 = make_desc_sw32(smem_u32(sAlo + h * (p.stage_bytes / 2)));
          const uint64_t a_s1 = make_desc_sw32(smem_u32(sAhi + h * (p.stage_bytes / 2)));
          const bool hi_t = p.a_smem == 0, lo_t = p.a_smem == 2;
          const uint64_t hi_s = p.a_smem == 1 ? a_s1 : a_s0, lo_s = a_s0;
          if (tc_elect_one()) {
            if (!(dbg & 2)) {
              const uint32_t nk = KS > 0 ? (uint32_t)KS : ks;
              if (hi_t) {
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ts_1(d, a_t + kk * 8, bhi + kk * kstride, idesc, kk);
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ts_1(d, a_t + kk * 8, blo + kk * kstride, idesc, 1);
              } else {
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ss_1(d, hi_s + kk * kstride, bhi + kk * kstride, idesc, kk);
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ss_1(d, hi_s + kk * kstride, blo + kk * kstride, idesc, 1);
              }
              if (lo_t) {
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ts_1(d, a_t + kk * 8, bhi + kk * kstride, idesc, 1);
              } else {
#pragma unroll
                for (uint32_t kk = 0; kk < nk; kk++) tc_mma_ss_1(d, lo_s + kk * kstride, bhi + kk * kstride, idesc, 1);
              }
            }
            if (h + 1 == nh) tc_commit(&empty[stage]);   // stage reusable once every MMA that reads it is done
            tc_commit(&tfull[buf]);
          }
          __syncwarp();
        }
      }
    }
    if (prof && lane == 0) {
      // the last accumulator's completion marks the end of the tensor work of this CTA
      const uint32_t last = it * nh - 1;
      mbar_wait(&tfull[last % kNumAcc], (last / kNumAcc) & 1);
      const long long c2 = clock64();
      unsigned long long g1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
      atomicAdd(p.prof + 1, (unsigned long long)(c2 - c0));      // issue-warp lifetime, cycles
      atomicAdd(p.prof + 2, g1 - g0);                            // the same in ns
      atomicAdd(p.prof + 3, (unsigned long long)w_a);            // waiting for the A rows
      atomicAdd(p.prof + 4, (unsigned long long)w_full);         // waiting for B stages
      atomicAdd(p.prof + 5, (unsigned long long)w_tempty);       // waiting for the epilogue to drain an accumulator
      atomicAdd(p.prof + 6, (unsigned long long)it * nh * 3 * ks);   // MMAs issued
      atomicAdd(p.prof + 7, 1ull);                               // CTAs
    }
  } else {
    // ===== epilogue warps: (half, lane quarter, column part) =====
    const uint32_t quarter = warp & 3;
    const uint32_t group = ((uint32_t)(warp - 2) >> 2) & 1u;      // frame half
    const uint32_t part = (uint32_t)(warp - 2) >> 3;              // columns [64 * part, 64 * part + 64) of every tile
    if (group < nh) {
      const uint32_t lrow = quarter * 32 + lane;
      const uint32_t row = group * BM + lrow;
      const uint32_t tlane = tmem_base + ((quarter * 32) << 16);
      // A rows of this thread's frame: hi -> TMEM, lo -> shared memory (part 0 only)
      if (part == 0) {
        const uint64_t fr = prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1);   // rows past the end are never read back
        const float4* ah = reinterpret_cast<const float4*>(p.Ahi + fr * KP);
        const float4* al = reinterpret_cast<const float4*>(p.Alo + fr * KP);
        const uint32_t acol = tlane + group * KP;
        // lo part: [kstep][row][32 B] with byte-address bit 7 (= row bit 2) XORed into bit 4
        // shared-memory parts: [kstep][row][32 B] with byte-address bit 7 (= row bit 2) XORed into bit 4
        unsigned char* s0_row = sAlo + group * (p.stage_bytes / 2) + lrow * 32;
        unsigned char* s1_row = sAhi + group * (p.stage_bytes / 2) + lrow * 32;
        const uint32_t sw = ((lrow >> 2) & 1u) << 4;
        const float4* to_t = p.a_smem == 0 ? ah : (p.a_smem == 2 ? al : nullptr);   // part kept in TMEM
        const float4* to_s0 = p.a_smem == 2 ? ah : al;
        for (uint32_t kk = 0; kk < ks; kk++) {
          if (to_t) tc_st8(acol + kk * 8, __ldg(to_t + 2 * kk), __ldg(to_t + 2 * kk + 1));
          *reinterpret_cast<float4*>(s0_row + kk * 4096 + (0 ^ sw)) = __ldg(to_s0 + 2 * kk);
          *reinterpret_cast<float4*>(s0_row + kk * 4096 + (16 ^ sw)) = __ldg(to_s0 + 2 * kk + 1);
          if (p.a_smem == 1) {
            *reinterpret_cast<float4*>(s1_row + kk * 4096 + (0 ^ sw)) = __ldg(ah + 2 * kk);
            *reinterpret_cast<float4*>(s1_row + kk * 4096 + (16 ^ sw)) = __ldg(ah + 2 * kk + 1);
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the MMA
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full);
      }
      const float2 xn = p.xnorm[prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1)];
      const uint64_t my_cell = prow ? prow[row].cell : tc_row_cell(out_base, row, out_stride, true);
      const uint16_t* my_map = prow ? p.pack_colmap + prow[row].colmap : nullptr;
      float* my_x = xchg + (group * BM + lrow) * 5;               // [slot][half][row][5]
      uint32_t it = 0;
      long long e_wait = 0;
      const long long e_start = (kProf && p.prof) ? clock64() : 0;
      for (uint32_t j = 0; j < nu; j++) {
        const uint32_t s = ulist ? ulist[j] : j;
        const float4 g = __ldg(p.sguard + s);
        // certified bound on |f32 score - exact score| differences between two candidates (+0.1 %)
        const float Gs = 2.002f * p.eps * (xn.x * g.x + xn.y * g.y + g.z);
        float Bbest = INFINITY, Bsecond = INFINITY, Bacc = 0.f, Bidx = 0.f, BG = 0.f;
        uint32_t n0 = 0;
        for (uint32_t t = p.tile_first[s]; t < p.tile_first[s + 1]; t++, it++) {
          const uint32_t N = p.tile_info[t].y;
          const uint32_t unit = it * nh + group, buf = unit % kNumAcc, use = unit / kNumAcc;
          long long e0 = 0;
          if (kProf && p.prof) e0 = clock64();
          mbar_wait(&tfull[buf], use & 1);
          if (kProf && p.prof) e_wait += clock64() - e0;
          tc_fence_after();
          const uint32_t tcol = tlane + kAccCol + buf * BN;
          const uint32_t c_lo = part * 64, c_hi = min(N, c_lo + 64);
          float m = INFINITY, acc = 0.f, idx = 0.f, G = 0.f;
          if (!(dbg & 1) && c_lo < c_hi) {
            uint32_t ra[16], rb[16];
#pragma unroll
            for (int e = 0; e < 16; e++) rb[e] = 0;
            // pass 1: minimum of this part's columns (loads run one block ahead; four independent chains)
            float mm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
            tc_ld16_nowait(tcol + c_lo, ra);
            tc_ld_wait(ra);
            for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
              if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
#pragma unroll
              for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(ra[e]));
              tc_ld_wait(rb);
              if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
              if (cc + 16 < c_hi) {
#pragma unroll
                for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(rb[e]));
              }
              tc_ld_wait(ra);
            }
            m = fminf(fminf(mm[0], mm[1]), fminf(mm[2], mm[3]));
            // pass 2: count and locate every column below min + G (the minimum itself always counts)
            G = Gs + 4e-6f * fabsf(m);
            const float thr = m + G, nbig = -0x1p100f, T = thr * 0x1p100f;
            const bool thr_ok = fabsf(thr) > 1e-20f && fabsf(thr) < 1e8f;   // else c may be fractional: uncertain
            TcCount cnt;
            cnt.init();
            tc_ld16_nowait(tcol + c_lo, ra);
            tc_ld_wait(ra);
            for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
              if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
              cnt.block16(ra, nbig, T);
              tc_ld_wait(rb);
              if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
              if (cc + 16 < c_hi) cnt.block16(rb, nbig, T);
              tc_ld_wait(ra);
            }
            cnt.finish(acc, idx);
            idx += (float)c_lo;
            if (!thr_ok) acc = 2.f;
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[buf]);
          // merge with the state's earlier tiles
          if (m < Bbest) { Bsecond = Bbest; Bbest = m; Bacc = acc; Bidx = (float)n0 + idx; BG = G; }
          else Bsecond = fminf(Bsecond, m);
          n0 += N;
        }
        // the two column parts of a (half, quarter) meet through shared memory once per state
        float* slot = my_x + (j & 1) * (2 * BM * 5);
        if (part == 1) { slot[0] = Bbest; slot[1] = Bsecond; slot[2] = Bacc; slot[3] = Bidx; slot[4] = BG; }
        asm volatile("bar.sync %0, 64;" ::"r"(1 + group * 4 + quarter) : "memory");
        if (part == 0) {
          const float ob = slot[0], os = slot[1];
          if (ob < Bbest) { Bsecond = fminf(Bbest, os); Bbest = ob; Bacc = slot[2]; Bidx = slot[3]; BG = slot[4]; }
          else Bsecond = fminf(Bsecond, ob);
          if (row < nvalid) {
            const bool certain = Bacc == 1.0f && Bsecond > Bbest + BG;
            const uint32_t col = my_map ? (uint32_t)my_map[j] : j;     // packed: this frame's utterance may not use state j
            if (col != 0xFFFFu || !my_map) p.cand[tc_cell(my_cell, col, true)] = certain ? (uint32_t)Bidx : kUncertain;
          }
        }
      }
      if (kProf && p.prof && lane == 0 && quarter == 0) {   // one lane per (half, part) warp group
        atomicAdd(p.prof + 8 + part, (unsigned long long)e_wait);                  // epilogue waiting for the tensor pipe
        atomicAdd(p.prof + 10 + part, (unsigned long long)(clock64() - e_start));  // epilogue loop lifetime
        atomicAdd(p.prof + 12, 1ull);
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


// per-state metadata of a GEMM task (k_tc_gemm256): state, its first B tile, number of tiles, width of the first tile,
// guard constants — 32 bytes, cached in shared memory for the first kMetaCache states of a task
struct __align__(16) TcMeta { uint32_t s, tf, nt, N0; float4 g; };
constexpr uint32_t kMetaCache = 256;
__device__ __forceinline__ TcMeta tc_meta_load(const TcParams& p, const uint32_t* ulist, uint32_t j) {
  TcMeta m;
  m.s = ulist ? ulist[j] : j;
  m.tf = p.tile_first[m.s];
  m.nt = p.tile_first[m.s + 1] - m.tf;
  m.N0 = m.nt ? p.tile_info[m.tf].y : 0u;
  m.g = __ldg(p.sguard + m.s);
  return m;
}

// ---- N = 256 variant.  Measured (tools/ubench_epilogue.cu): an N = 128 tf32 MMA retires every 80-106 cycles (shared-
// memory bandwidth with A from shared memory, ~101 cycles with A from TMEM), an N = 256 one at the full 128 cycles.
// Tiles here are up to 256 densities wide: two 256-column accumulators (all of TMEM), both A parts in shared memory,
// and a ring of three HALF stages — a tile's hi block [k-step][256 rows][32 B] and its lo block are separate slots, so
// that three slots (3 x 40 KB at d = 33) keep one tile in use and half of the next in flight.  Per (tile, half) unit:
// Ahi.Bhi, Alo.Bhi (hi slot), then Ahi.Blo (lo slot).  With two halves, half h always owns accumulator h.
// F16: operands are fp16 hi / lo pairs (k-step = 16 halves) and the MMAs kind::f16 — same bytes per k-step, same
// shared-memory images and descriptors, K = d + 1 = 34 takes 3 k-steps instead of 5 (see k_tc_prep16).
// E16: all sixteen epilogue warps work on every unit — warp = (lane quarter, one of four 64-column parts), each thread
// owning the same row of BOTH halves — so that an accumulator is handed back in about half the time; the parts of a
// (half, quarter) meet through shared memory once per state (three writers, one reader, 128-thread named barrier).
// MASK: the mask mode of the sum scores (tc_subpart_mask) — an instantiation of its own: compiled into the default kernel,
// the extra path cost the max-approx filter 0.45 ms (3.95 -> 4.41), measured.
template <int KS, bool F16, bool E16 = false, bool MASK = false>
__global__ void __launch_bounds__(576, 1) k_tc_gemm256(TcParams p) {   // 96 registers: 18 warps = 5 on two of the four sub-partitions (16 K registers each)
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* sA = smem_raw + (base - smem_u32(smem_raw));      // [half][hi, lo][kstep][128 rows][32 B], 32-byte swizzle
  unsigned char* sB = sA + p.nhalf * 2 * (p.stage_bytes / 2);      // [slot][kstep][256 rows][32 B]; stage_bytes = one slot
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + p.nstage * p.stage_bytes);
  uint64_t* a_full = bars;                 // A rows are in TMEM (8 epilogue warps arrive)
  uint64_t* full = bars + 1;
  uint64_t* empty = full + kMaxStage;
  uint64_t* tfull = empty + kMaxStage;     // [3] accumulator complete
  uint64_t* tempty = tfull + kNumAcc;      // accumulator drained (two of the slots used)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + kNumAcc);
  float* xchg = reinterpret_cast<float*>(tmem_ptr + 4);   // [2 slots][2 halves][128 rows][5]: column-part merge
  // per-state metadata of this task, staged once: the three roles below walk the task's states in lock step, and a chain
  // of dependent global loads per state (state list -> first tile -> tile width) in the MMA-issuing warp was ~800 cycles
  // of its critical path per unit (measured: its busy time did not change when the MMA count went from 9 to 7)
  TcMeta* meta = reinterpret_cast<TcMeta*>(
      (reinterpret_cast<uintptr_t>(xchg) + (E16 ? 3 : 1) * (2 * 2 * 128 * 5 * sizeof(float)) + 15) & ~(uintptr_t)15);

  // Warp roles.  The SM sub-partition's arbiter prefers the HIGHEST warp id (B300_MICROARCH.md): the two warps that feed
  // the tensor pipe get the highest ids of their sub-partitions, so that sixteen busy epilogue warps cannot starve them
  // (as warp 1 among warps 5, 9, 13, 17 the MMA issuer took 97 cycles per 64-cycle-equivalent MMA, measured).
  constexpr int kProdWarp = 16, kMmaWarp = 17;      // epilogue: warps 0 .. 15 (TMEM lane quarter = warp % 4)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = blockIdx.x;
  const uint32_t ROWS = BM * p.nhalf;
  const uint32_t nstage = p.nstage, ks = KS > 0 ? (uint32_t)KS : p.ksteps, KP = 8 * ks;
  constexpr uint32_t kstrideA = 4096 >> 4, kstrideB = 8192 >> 4;   // descriptor units between k-step blocks (128 / 256 rows)
  const uint32_t a_part = p.stage_bytes / 2;   // bytes of one A part of one half: ksteps * 4 KB

  // ---- task decode (uniform over the CTA)
  uint64_t row0;        // first A row (absolute frame index)
  uint32_t nvalid, nu, out_stride;
  const uint32_t* ulist = nullptr;
  const sgmm_pack_row* prow = nullptr;   // packed task: per-row frame / output cell / column map
  uint64_t out_base;
  if (p.dense) {
    const uint64_t r0 = task * ROWS;
    row0 = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)ROWS, p.nframes - r0);
    nu = p.n_states;
    out_stride = p.n_states;
    out_base = r0 * p.n_states;
  } else if (p.tasks[task].x & 0x80000000u) {
    const uint32_t id = p.tasks[task].x & 0x7fffffffu;
    const uint4 d = p.pack_desc[id];
    prow = p.pack_rows + (size_t)id * 128;
    row0 = 0; nvalid = d.z; nu = d.y; out_stride = 0; out_base = 0;
    ulist = p.pack_uniq + d.x;
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
  const uint32_t nh = nvalid > (uint32_t)BM ? 2u : 1u;   // halves that hold valid frames

  for (uint32_t j = threadIdx.x; j < min(nu, kMetaCache); j += blockDim.x) meta[j] = tc_meta_load(p, ulist, j);
  auto meta_of = [&](uint32_t j) -> TcMeta { return j < kMetaCache ? meta[j] : tc_meta_load(p, ulist, j); };
  // ---- one-time setup
  if (warp == kProdWarp && lane == 0) {
    mbar_init(a_full, E16 ? 4 : 4 * nh);
    for (uint32_t i = 0; i < nstage; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (uint32_t i = 0; i < kNumAcc; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], E16 ? 16 : 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == kProdWarp) {
    // ===== producer: two bulk copies per B tile (hi block, lo block), one ring slot each =====
    if (lane == 0) {
      uint32_t ld = 0;
      for (uint32_t j = 0; j < nu; j++) {
        const TcMeta mt = meta_of(j);
        for (uint32_t t = mt.tf; t < mt.tf + mt.nt; t++) {
          for (uint32_t op = 0; op < 2; op++, ld++) {
            // (packed remainder: the lo slot's last k-step block is unused and not fetched)
            const uint32_t slot = ld % nstage, bytes = p.stage_bytes, nb = (op && p.krem) ? bytes - 8192u : bytes;
            mbar_wait(&empty[slot], ((ld / nstage) & 1) ^ 1);
            mbar_expect_tx(&full[slot], nb);
            bulk_load(sB + slot * bytes, p.Bimg + ((size_t)t * 2 + op) * (bytes / 4), nb, &full[slot]);
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer: the whole warp walks the loop, one elected lane issues a unit's MMAs =====
    const bool prof = kProf && p.prof != nullptr;
    long long w_a = 0, w_full = 0, w_tempty = 0, c0 = 0, c1 = 0;
    unsigned long long g0 = 0;
    if (prof) { c0 = clock64(); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0)); }
    mbar_wait(a_full, 0);
    if (prof) w_a = clock64() - c0;
    tc_fence_after();
    uint32_t ld = 0, unit = 0;
    auto mma1 = [](uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
      if (F16) tc_mma_ss_f16_1(d_tmem, adesc, bdesc, idesc, acc);
      else tc_mma_ss_1(d_tmem, adesc, bdesc, idesc, acc);
    };
    for (uint32_t j = 0; j < nu; j++) {
      const TcMeta mt = meta_of(j);
      for (uint32_t t = mt.tf; t < mt.tf + mt.nt; t++, ld += 2) {
        const uint32_t N = t == mt.tf ? mt.N0 : p.tile_info[t].y;
        const uint32_t idesc = F16 ? make_idesc_f16(BM, N) : make_idesc_tf32(BM, N);
        const uint32_t slot_hi = ld % nstage, par_hi = (ld / nstage) & 1;
        const uint32_t slot_lo = (ld + 1) % nstage, par_lo = ((ld + 1) / nstage) & 1;
        const uint64_t bhi = make_desc_sw32(smem_u32(sB + slot_hi * p.stage_bytes));
        const uint64_t blo = make_desc_sw32(smem_u32(sB + slot_lo * p.stage_bytes));
        for (uint32_t h = 0; h < nh; h++, unit++) {
          const uint32_t buf = unit & 1u, use = unit >> 1;
          if (prof) c1 = clock64();
          mbar_wait(&tempty[buf], (use & 1) ^ 1);
          if (prof) { const long long c2 = clock64(); w_tempty += c2 - c1; c1 = c2; }
          if (h == 0) mbar_wait(&full[slot_hi], par_hi);
          if (prof) w_full += clock64() - c1;
          tc_fence_after();
          const uint32_t d = tmem_base + buf * 256;
          const uint64_t ahi = make_desc_sw32(smem_u32(sA + (h * 2 + 0) * a_part));
          const uint64_t alo = make_desc_sw32(smem_u32(sA + (h * 2 + 1) * a_part));
          const uint32_t nk = KS > 0 ? (uint32_t)KS : ks;
          const uint32_t nfull = p.krem ? nk - 1 : nk;     // whole k-steps; then (packed) one remainder MMA on the hi slot
          if (tc_elect_one()) {
#pragma unroll
            for (uint32_t kk = 0; kk < nk; kk++)
              if (kk < nfull || p.krem) mma1(d, ahi + kk * kstrideA, bhi + kk * kstrideB, idesc, kk);
#pragma unroll
            for (uint32_t kk = 0; kk < nk; kk++)
              if (kk < nfull) mma1(d, alo + kk * kstrideA, bhi + kk * kstrideB, idesc, 1);
            if (h + 1 == nh) tc_commit(&empty[slot_hi]);    // the hi block is free once these MMAs are done
          }
          __syncwarp();
          if (prof) c1 = clock64();
          if (h == 0) mbar_wait(&full[slot_lo], par_lo);
          if (prof) w_full += clock64() - c1;
          tc_fence_after();
          if (tc_elect_one()) {
#pragma unroll
            for (uint32_t kk = 0; kk < nk; kk++)
              if (kk < nfull) mma1(d, ahi + kk * kstrideA, blo + kk * kstrideB, idesc, 1);
            if (h + 1 == nh) tc_commit(&empty[slot_lo]);
            tc_commit(&tfull[buf]);
          }
          __syncwarp();
        }
      }
    }
    if (prof && lane == 0 && unit > 0) {
      const uint32_t last = unit - 1;
      mbar_wait(&tfull[last & 1u], (last >> 1) & 1);
      const long long c2 = clock64();
      unsigned long long g1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
      atomicAdd(p.prof + 1, (unsigned long long)(c2 - c0));      // issue-warp lifetime, cycles
      atomicAdd(p.prof + 2, g1 - g0);                            // the same in ns
      atomicAdd(p.prof + 3, (unsigned long long)w_a);            // waiting for the A rows
      atomicAdd(p.prof + 4, (unsigned long long)w_full);         // waiting for B slots
      atomicAdd(p.prof + 5, (unsigned long long)w_tempty);       // waiting for the epilogue to hand an accumulator back
      atomicAdd(p.prof + 6, (unsigned long long)unit * (p.krem ? 3 * ks - 2 : 3 * ks) * 2);   // MMAs issued, in N = 128 equivalents
      atomicAdd(p.prof + 7, 1ull);                               // CTAs
    }
  } else if (E16) {
    // ===== epilogue warps, sixteen per unit: (lane quarter, 64-column part); thread = one row of both halves =====
    const uint32_t quarter = warp & 3;
    const uint32_t part = (uint32_t)warp >> 2;                    // columns [64 * part, 64 * part + 64) of every tile
    const uint32_t lrow = quarter * 32 + lane;
    const uint32_t tlane = tmem_base + ((quarter * 32) << 16);
    if (part == 0) {   // A rows of this thread's frames (both halves): hi and lo parts into shared memory
      for (uint32_t g = 0; g < nh; g++) {
        const uint32_t row = g * BM + lrow;
        const uint64_t fr = prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1);
        const float4* ah = reinterpret_cast<const float4*>(p.Ahi + fr * KP);
        const float4* al = reinterpret_cast<const float4*>(p.Alo + fr * KP);
        unsigned char* hi_row = sA + (g * 2 + 0) * a_part + lrow * 32;
        unsigned char* lo_row = sA + (g * 2 + 1) * a_part + lrow * 32;
        const uint32_t sw = ((lrow >> 2) & 1u) << 4;
        for (uint32_t kk = 0; kk < ks; kk++) {
          *reinterpret_cast<float4*>(hi_row + kk * 4096 + (0 ^ sw)) = __ldg(ah + 2 * kk);
          *reinterpret_cast<float4*>(hi_row + kk * 4096 + (16 ^ sw)) = __ldg(ah + 2 * kk + 1);
          *reinterpret_cast<float4*>(lo_row + kk * 4096 + (0 ^ sw)) = __ldg(al + 2 * kk);
          *reinterpret_cast<float4*>(lo_row + kk * 4096 + (16 ^ sw)) = __ldg(al + 2 * kk + 1);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full);
    }
    float2 xn[2];
    uint64_t my_cell[2];
    const uint16_t* my_map[2];
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const uint32_t row = min((uint32_t)g, nh - 1) * BM + lrow;
      xn[g] = p.xnorm[prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1)];
      my_cell[g] = prow ? prow[row].cell : tc_row_cell(out_base, row, out_stride, true);
      my_map[g] = prow ? p.pack_colmap + prow[row].colmap : nullptr;
    }
    uint32_t it = 0;
    for (uint32_t j = 0; j < nu; j++) {
      const TcMeta mt = meta_of(j);
      const float4 g4 = mt.g;
      float Gs[2], Bbest[2], Bsecond[2], Bacc[2], Bidx[2], BG[2];
#pragma unroll
      for (int g = 0; g < 2; g++) {
        Gs[g] = 2.002f * p.eps * (xn[g].x * g4.x + xn[g].y * g4.y + g4.z);
        Bbest[g] = INFINITY; Bsecond[g] = INFINITY; Bacc[g] = 0.f; Bidx[g] = 0.f; BG[g] = 0.f;
      }
      uint32_t n0 = 0;
      for (uint32_t t = mt.tf; t < mt.tf + mt.nt; t++, it++) {
        const uint32_t N = t == mt.tf ? mt.N0 : p.tile_info[t].y;
        const uint32_t c_lo = part * 64, c_hi = min(N, c_lo + 64);
#pragma unroll
        for (int g = 0; g < 2; g++) {
          if ((uint32_t)g < nh) {
            const uint32_t unit = it * nh + g, buf = unit & 1u, use = unit >> 1;
            mbar_wait(&tfull[buf], use & 1);
            tc_fence_after();
            float m = INFINITY, acc = 0.f, idx = 0.f, G = 0.f;
            if (kTwoPassRef && (p.debug & 128)) {   // A/B reference: the two-pass scan, accumulator handed back after the arithmetic
              if (!(p.debug & 1) && c_lo < c_hi) tc_scan_columns(tlane + buf * 256, c_lo, c_hi, Gs[g], m, acc, idx, G);
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&tempty[buf]);
            } else {
              // the part's 64 columns go to registers in one burst and the accumulator is handed back BEFORE the
              // arithmetic: the next unit's MMAs run under this unit's epilogue math
              tc_scan_onepass<4, true>(tlane + buf * 256, c_lo, c_hi, Gs[g], &tempty[buf], lane, p.debug, m, acc, idx, G);
            }
            if (m < Bbest[g]) { Bsecond[g] = Bbest[g]; Bbest[g] = m; Bacc[g] = acc; Bidx[g] = (float)n0 + idx; BG[g] = G; }
            else Bsecond[g] = fminf(Bsecond[g], m);
          }
        }
        n0 += N;
      }
      // the four column parts of a (half, quarter) meet through shared memory once per state
      float* slots = xchg + (size_t)(j & 1) * (2 * 3 * BM * 5);    // [slot][half][part - 1][row][5]
      if (part > 0) {
#pragma unroll
        for (int g = 0; g < 2; g++) {
          if ((uint32_t)g < nh) {
            float* sl = slots + ((size_t)(g * 3 + part - 1) * BM + lrow) * 5;
            sl[0] = Bbest[g]; sl[1] = Bsecond[g]; sl[2] = Bacc[g]; sl[3] = Bidx[g]; sl[4] = BG[g];
          }
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(1 + quarter) : "memory");
      if (part == 0) {
#pragma unroll
        for (int g = 0; g < 2; g++) {
          if ((uint32_t)g < nh) {
#pragma unroll
            for (int q = 0; q < 3; q++) {
              const float* sl = slots + ((size_t)(g * 3 + q) * BM + lrow) * 5;
              const float ob = sl[0], os = sl[1];
              if (ob < Bbest[g]) { Bsecond[g] = fminf(Bbest[g], os); Bbest[g] = ob; Bacc[g] = sl[2]; Bidx[g] = sl[3]; BG[g] = sl[4]; }
              else Bsecond[g] = fminf(Bsecond[g], ob);
            }
            const uint32_t row = g * BM + lrow;
            if (row < nvalid) {
              const bool certain = Bacc[g] == 1.0f && Bsecond[g] > Bbest[g] + BG[g];
              const uint32_t col = my_map[g] ? (uint32_t)my_map[g][j] : j;
              if (col != 0xFFFFu || !my_map[g]) p.cand[tc_cell(my_cell[g], col, true)] = certain ? (uint32_t)Bidx[g] : kUncertain;
            }
          }
        }
      }
    }
  } else {
    // ===== epilogue warps: (half, lane quarter, column part) =====
    const uint32_t quarter = warp & 3;
    const uint32_t group = ((uint32_t)warp >> 2) & 1u;            // frame half
    const uint32_t part = (uint32_t)warp >> 3;              // columns [64 * part, 64 * part + 64) of every tile
    if (group < nh) {
      const uint32_t lrow = quarter * 32 + lane;
      const uint32_t row = group * BM + lrow;
      const uint32_t tlane = tmem_base + ((quarter * 32) << 16);
      // A rows of this thread's frame: hi -> TMEM, lo -> shared memory (part 0 only)
      if (part == 0) {
        const uint64_t fr = prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1);   // rows past the end are never read back
        const float4* ah = reinterpret_cast<const float4*>(p.Ahi + fr * KP);
        const float4* al = reinterpret_cast<const float4*>(p.Alo + fr * KP);
        // [kstep][row][32 B] with byte-address bit 7 (= row bit 2) XORed into bit 4
        unsigned char* hi_row = sA + (group * 2 + 0) * a_part + lrow * 32;
        unsigned char* lo_row = sA + (group * 2 + 1) * a_part + lrow * 32;
        const uint32_t sw = ((lrow >> 2) & 1u) << 4;
        for (uint32_t kk = 0; kk < ks; kk++) {
          *reinterpret_cast<float4*>(hi_row + kk * 4096 + (0 ^ sw)) = __ldg(ah + 2 * kk);
          *reinterpret_cast<float4*>(hi_row + kk * 4096 + (16 ^ sw)) = __ldg(ah + 2 * kk + 1);
          *reinterpret_cast<float4*>(lo_row + kk * 4096 + (0 ^ sw)) = __ldg(al + 2 * kk);
          *reinterpret_cast<float4*>(lo_row + kk * 4096 + (16 ^ sw)) = __ldg(al + 2 * kk + 1);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the MMA
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full);
      }
      const float2 xn = p.xnorm[prow ? (uint64_t)prow[row].frame : min(row0 + row, p.n_frames - 1)];
      const uint64_t my_cell = prow ? prow[row].cell : tc_row_cell(out_base, row, out_stride, true);
      const uint16_t* my_map = prow ? p.pack_colmap + prow[row].colmap : nullptr;
      float* my_x = xchg + (group * BM + lrow) * 5;               // [slot][half][row][5]: a part's TcBest
      uint32_t it = 0;
      long long e_wait = 0;
      const long long e_start = (kProf && p.prof) ? clock64() : 0;
      for (uint32_t j = 0; j < nu; j++) {
        const TcMeta mt = meta_of(j);
        const float4 g = mt.g;
        // certified bound on |f32 score - exact score| differences between two candidates (+0.1 %)
        const float Gs = 2.002f * p.eps * (xn.x * g.x + xn.y * g.y + g.z);
        TcBest B;                                                   // candidates of this (row, state) cell so far
        B.init();
        uint32_t n0 = 0;
        for (uint32_t t = mt.tf; t < mt.tf + mt.nt; t++, it++) {
          const uint32_t N = t == mt.tf ? mt.N0 : p.tile_info[t].y;
          const uint32_t unit = it * nh + group, buf = unit & 1u, use = unit >> 1;
          long long e0 = 0;
          if (kProf && p.prof) e0 = clock64();
          mbar_wait(&tfull[buf], use & 1);
          if (kProf && p.prof) e_wait += clock64() - e0;
          tc_fence_after();
          const uint32_t tcol = tlane + buf * 256;
          const uint32_t c_lo = part * 128, c_hi = min(N, c_lo + 128);
          float m = INFINITY, acc = 0.f, idx = 0.f, G = 0.f;
          if (MASK) {
            // mask mode: 45 in accumulator units (s_f t_s = xn.y / g.w) + twice the certified bound; single-tile states
            const uint32_t col = my_map ? (uint32_t)my_map[j] : j;
            uint4* out = (row < nvalid && (col != 0xFFFFu || !my_map)) ? p.summask + 4 * tc_cell(my_cell, col, true) : nullptr;
            // pass 1: this part's minimum; the two column parts of a (half, quarter) swap theirs through the exchange
            // slots, so that pass 2's threshold hangs on the STATE's minimum (masks relative to a sub-part's own
            // minimum held 3x the candidates: measured 10.4 ms of summation per 13 M cells)
            float* mslot = my_x + (j & 1) * (2 * BM * 5);
            mslot[part] = tc_part_min<4>(tcol, c_lo, c_hi);
            asm volatile("bar.sync %0, 64;" ::"r"(1 + group * 4 + quarter) : "memory");
            const float gm = fminf(mslot[0], mslot[1]);
            tc_scan_masks<4>(tcol, c_lo, c_hi, 45.f * (xn.y / g.w) + 2.f * Gs, gm, &tempty[buf], lane, out);
          } else if (!kTwoPassRef || !(p.debug & 128)) {
            // one pass over tensor memory (default); built with -DSGMM_TC_TWOPASS_REF=1, tc_debug & 128 selects the two-pass
            // scan below as the A/B reference (compiled out by default: dead code in this kernel is not free, see MASK)
            TcBest st;
            tc_scan_onepass2<4>(tcol, c_lo, c_hi, Gs, &tempty[buf], lane, kProf ? p.debug : 0u, st);
            if (t == mt.tf) B = st;                               // first tile of the state (n0 = 0): the common case
            else B.merge(st, n0);
          } else {
          if (!(p.debug & 1) && c_lo < c_hi) {
            uint32_t ra[16], rb[16];
#pragma unroll
            for (int e = 0; e < 16; e++) rb[e] = 0;
            // pass 1: minimum of this part's columns (loads run one block ahead; four independent chains)
            float mm[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
            tc_ld16_nowait(tcol + c_lo, ra);
            tc_ld_wait(ra);
            for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
              if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
#pragma unroll
              for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(ra[e]));
              tc_ld_wait(rb);
              if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
              if (cc + 16 < c_hi) {
#pragma unroll
                for (int e = 0; e < 16; e++) mm[e & 3] = fminf(mm[e & 3], __uint_as_float(rb[e]));
              }
              tc_ld_wait(ra);
            }
            m = fminf(fminf(mm[0], mm[1]), fminf(mm[2], mm[3]));
            // pass 2: count and locate every column below min + G (the minimum itself always counts)
            G = Gs + 4e-6f * fabsf(m);
            const float thr = m + G, nbig = -0x1p100f, T = thr * 0x1p100f;
            const bool thr_ok = fabsf(thr) > 1e-20f && fabsf(thr) < 1e8f;   // else c may be fractional: uncertain
            TcCount cnt;
            cnt.init();
            tc_ld16_nowait(tcol + c_lo, ra);
            tc_ld_wait(ra);
            for (uint32_t cc = c_lo; cc < c_hi; cc += 32) {
              if (cc + 16 < c_hi) tc_ld16_nowait(tcol + cc + 16, rb);
              cnt.block16(ra, nbig, T);
              tc_ld_wait(rb);
              if (cc + 32 < c_hi) tc_ld16_nowait(tcol + cc + 32, ra);
              if (cc + 16 < c_hi) cnt.block16(rb, nbig, T);
              tc_ld_wait(ra);
            }
            cnt.finish(acc, idx);
            idx += (float)c_lo;
            if (!thr_ok) acc = 2.f;
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[buf]);
          // (two-pass reference: a count other than 1 carries no columns)
          B.ins(m, ((acc == 1.0f ? 1u : 3u) << 30) | (n0 + (uint32_t)idx));
          }
          n0 += N;
        }
        // the two column parts of a (half, quarter) meet through shared memory once per state
        float* slot = my_x + (j & 1) * (2 * BM * 5);
        if (part == 1) {
          slot[0] = B.m; slot[1] = B.m2; slot[2] = B.m3; slot[3] = __uint_as_float(B.c1); slot[4] = __uint_as_float(B.c2);
        }
        asm volatile("bar.sync %0, 64;" ::"r"(1 + group * 4 + quarter) : "memory");
        if (part == 0) {
          TcBest o;
          o.m = slot[0]; o.m2 = slot[1]; o.m3 = slot[2]; o.c1 = __float_as_uint(slot[3]); o.c2 = __float_as_uint(slot[4]);
          B.merge(o, 0u);                                             // (its columns are already relative to the state)
          uint32_t vd = B.verdict(Gs);                                // (warp vote inside: before the rows diverge)
          if (row < nvalid) {
            const uint32_t col = my_map ? (uint32_t)my_map[j] : j;     // packed: this frame's utterance may not use state j
            if (col != 0xFFFFu || !my_map) {
              const uint64_t cell = tc_cell(my_cell, col, true);
              if (!p.pairs && (vd & kPair) && !(vd & kUncertain)) vd = kUncertain;
              if (!MASK) p.cand[cell] = vd;
              if (F16 && p.tol) p.e32[cell] = B.m * (g.w / xn.y);
            }
          }
        }
      }
      if (kProf && p.prof && lane == 0 && quarter == 0) {   // one lane per (half, part) warp group
        atomicAdd(p.prof + 8 + part, (unsigned long long)e_wait);                  // epilogue waiting for the tensor pipe
        atomicAdd(p.prof + 10 + part, (unsigned long long)(clock64() - e_start));  // epilogue loop lifetime
        atomicAdd(p.prof + 12, 1ull);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}



// ---- refinement: certify the f32 argmin, evaluate the winner in f64, queue near-ties for a full f64 rescan
struct RefineParams {
  ExactTables m;
  const uint32_t* state_off;
  const float* x;
  const uint32_t* cand;        // GEMM verdict per cell: certified argmin column, or kUncertain
  const sgmm_alignment_item* ali;   // k_sum_pruned, dense == 2: one cell per frame, its aligned state
  uint64_t n_cells;
  int dense;
  uint64_t f0;
  uint32_t n_states;
  uint32_t n_utts;
  const uint4* tasks;          // gathered: {utt, t0} granules of 128 frames
  int pooled;                  // every state's densities share one inverse-variance row
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
  // state-major kernel
  const uint4* sm_items;
  const uint4* sm_ent_a;
  const uint2* sm_ent_b;
  uint32_t sm_rw;              // doubles per table row: mean row, norm, padding (2 * odd: bank spread for LDS.128)
  uint32_t sm_maxnd;           // most densities in one state
  const uint32_t* sm_ivoff;    // per item: first double of its state's inverse-variance row in invU / c_iv (k_sm_ivoff)
  const uint4* summask;        // SUM: the filter's candidate masks, [cell][4] (tc_subpart_mask)
  uint32_t sm_dense_stride;    // 0: tiled cells (gathered pass); else the dense pass: cell = base + frame * stride + state (row-major)
  uint32_t sm_dense_f0;        // dense pass: its first frame (the verdict tiles are counted from there)
  uint32_t sm_lq;              // entries of the CTA-local near-tie list (0: none, every near-tie goes straight to the global queue)
  // two-stage rescan: f32 tables of the FFMA scorer ([D][4 * dv4] mean and h = 1 / (2 var), c = norm - log w) and
  // the per-state constants of its certified error bound {cmin, |c| max, M max}
  const float4* mu32T;         // [dv4][D]
  const float4* h32;
  const float* c32;
  const float4* sguard;
};

// The near-tie queue: rescan[0] = entries, then 16 bytes per entry {cell, frame | state << 40}.  The producer knows the
// cell's frame and state; carrying them spares the rescan a 12-step binary search over E_off per cell (dependent global
// loads: most of that kernel's time, measured).
__device__ __forceinline__ void rq_put(unsigned long long* q, unsigned long long slot, uint64_t cell, uint64_t frame, uint32_t s) {
  *reinterpret_cast<ulonglong2*>(q + 2 + 2 * slot) = make_ulonglong2(cell, frame | ((unsigned long long)s << 40));
}
__device__ __forceinline__ void rq_get(const unsigned long long* q, unsigned long long slot, uint64_t& cell, uint64_t& frame, uint32_t& s) {
  const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(q + 2 + 2 * slot);
  cell = v.x; frame = v.y & 0xffffffffffull; s = (uint32_t)(v.y >> 40);
}

// exact score from the row-major tables with NP dimension pairs known at compile time: every load of the mean
// row is in flight before the (sequential, reference-ordered) f64 chain starts; the frame's features (xs, f32
// pairs) and — for pooled models — the state's inverse variances (ivs) come from shared memory so that the
// global load queue only carries the mean row.  NP == 0: generic (everything from global memory).
template <int NP>
__device__ __forceinline__ double exact_score_unrolled(const ExactTables& m, const float* __restrict__ xf, uint32_t k,
                                                       const float* xs, const double* ivs) {
  if (NP == 0) return sgmm_exact_score_rm(m, xf, k);
  // 256-bit loads (LDG.256): one L1 tag lookup per 32-byte sector instead of two — this kernel is bound by the
  // 32-distinct-rows-per-instruction gather of the mean rows, not by bytes
  const double* mrow = m.meanR + (size_t)k * m.dimP;       // rows are 32-byte aligned (dimP % 4 == 0)
  double2 mu[NP > 0 ? NP : 1];
#pragma unroll
  for (int q = 0; q < NP / 2; q++)
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(mu[2 * q].x), "=d"(mu[2 * q].y), "=d"(mu[2 * q + 1].x), "=d"(mu[2 * q + 1].y)
                 : "l"(mrow + 4 * q));
  static_assert(NP % 2 == 0, "unrolled refinement expects an even number of dimension pairs");
  const double mt = (m.dim & 1) ? __ldg(mrow + m.dim - 1) : 0.0;
  const double* vg = m.invU + (size_t)__ldg(m.var_of + k) * m.dimP;
  double l0 = 0.0, l1 = 0.0;
#pragma unroll
  for (int ip = 0; ip < NP; ip++) {
    const double2 iv = ivs ? *reinterpret_cast<const double2*>(ivs + 2 * ip) : __ldg(reinterpret_cast<const double2*>(vg) + ip);
    const float2 xx = *reinterpret_cast<const float2*>(xs + 2 * ip);
    double a = __dsub_rn((double)xx.x, mu[ip].x);
    double b = __dsub_rn((double)xx.y, mu[ip].y);
    a = __dmul_rn(a, a); b = __dmul_rn(b, b);
    a = __dmul_rn(a, iv.x); b = __dmul_rn(b, iv.y);
    l0 = __dadd_rn(l0, a); l1 = __dadd_rn(l1, b);
  }
  double dist = __dadd_rn(l0, l1);
  if (m.dim & 1) {
    const double iv = ivs ? ivs[m.dim - 1] : __ldg(vg + m.dim - 1);
    const double c = __dsub_rn((double)xs[m.dim - 1], mt);
    dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), iv));
  }
  const double s = __dadd_rn(__ldg(m.norm + k), __dmul_rn(dist, 0.5));
  return __dsub_rn(s, __ldg(m.logw + k));
}

// One CTA per GEMM task granule (128 frames of one utterance, or 128 dense frames): cell -> (frame, state) is
// plain arithmetic, consecutive threads own consecutive cells of the emission matrix.
template <int NP>
__global__ void __launch_bounds__(256) k_tc_refine(RefineParams p) {
  uint64_t cell0, frame0;
  uint32_t nrows, ncol, nu;
  const uint32_t* ulist = nullptr;
  if (p.dense) {
    const uint64_t r0 = (uint64_t)blockIdx.x * 128;
    nrows = (uint32_t)min((uint64_t)128, p.n_cells / p.n_states - r0);
    ncol = nu = p.n_states;
    cell0 = r0 * p.n_states;
    frame0 = p.f0 + r0;
  } else {
    const uint4 t = p.tasks[blockIdx.x];
    const uint64_t uf0 = p.frame_off[t.x];
    nrows = min(128u, (uint32_t)(p.frame_off[t.x + 1] - uf0) - t.y);
    ncol = p.nuniq_pad[t.x];
    nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
    ulist = p.uniq + p.uniq_off[t.x];
    cell0 = p.E_off[t.x] + (uint64_t)t.y * ncol;
    frame0 = uf0 + t.y;
  }
  // stage the granule's feature rows (and, pooled models, the inverse-variance row of each of its states)
  extern __shared__ __align__(16) unsigned char refine_smem[];
  constexpr uint32_t XS = 2 * NP + 2;                        // floats per staged feature row
  float* xs_all = reinterpret_cast<float*>(refine_smem);
  double* iv_all = reinterpret_cast<double*>(refine_smem + 128 * XS * sizeof(float));
  if (NP > 0) {
    const uint32_t dim = p.m.dim;
    for (uint32_t i = threadIdx.x; i < nrows * dim; i += 256) {
      const uint32_t r = i / dim, q = i - r * dim;
      xs_all[r * XS + q] = __ldg(p.x + (frame0 + r) * dim + q);
    }
    if (p.pooled) {
      for (uint32_t i = threadIdx.x; i < nu * p.m.dimP; i += 256) {
        const uint32_t j = i / p.m.dimP, q = i - j * p.m.dimP;
        const uint32_t s = ulist ? ulist[j] : j;
        const uint32_t k0 = p.state_off[s];
        iv_all[i] = k0 < p.state_off[s + 1] ? __ldg(p.m.invU + (size_t)__ldg(p.m.var_of + k0) * p.m.dimP + q) : 0.0;
      }
    }
    __syncthreads();
  }
  // dense: cells are row-major.  Gathered: the granule starts on a tile boundary of the tiled layout (sgmm_cell_rel), its
  // cells are [tile][column][16 frames] and consecutive threads walk a column's 16-frame run.
  const bool tiled = !p.dense;
  const uint32_t ncell = tiled ? ((nrows + 15u) >> 4) * 16u * ncol : nrows * ncol;
  for (uint32_t i = threadIdx.x; i < ncell; i += 256) {
    uint32_t r, j;
    if (tiled) {
      const uint32_t tile = i / (16u * ncol), rem = i - tile * 16u * ncol;
      j = rem >> 4;
      r = tile * 16u + ((rem ^ j) & 15u);
      if (r >= nrows) continue;                  // padding frames of the last tile
    } else {
      r = i / ncol; j = i - r * ncol;
    }
    if (j >= nu) continue;                       // padding column of a gathered row
    const uint64_t e = cell0 + i, frame = frame0 + r;
    const uint32_t s = ulist ? ulist[j] : j;
    const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
    double best = 1e10;
    uint32_t bk = 0;
    if (k1 > k0) {
      const uint32_t i1 = p.cand[tiled ? e : cell0 + (uint64_t)(r >> 4) * (16u * ncol) + (j << 4) + ((r ^ j) & 15u)];   // verdicts are tiled in both passes
      if (i1 < k1 - k0) {                        // certified (kUncertain and padding columns fail this)
        const double sc = exact_score_unrolled<NP>(p.m, p.x + frame * p.m.dim, k0 + i1, xs_all + r * XS,
                                                   p.pooled ? iv_all + j * p.m.dimP : nullptr);
        if (sc < best) { best = sc; bk = i1; }
      } else if ((i1 & kPair) && !(i1 & kUncertain) && (i1 & 0x7fffu) < k1 - k0 && ((i1 >> 15) & 0x7fffu) < k1 - k0) {
        // pair verdict: the argmin is one of two columns (a < b) — both evaluated, the first minimum wins
        const uint32_t ka = i1 & 0x7fffu, kb = (i1 >> 15) & 0x7fffu;
        const double sa = sgmm_exact_score_rm(p.m, p.x + frame * p.m.dim, k0 + ka);
        const double sb = sgmm_exact_score_rm(p.m, p.x + frame * p.m.dim, k0 + kb);
        if (sa < best) { best = sa; bk = ka; }
        if (sb < best) { best = sb; bk = kb; }
      } else {
        // one atomic per group of lanes that arrive here together (same-address atomics serialise at the L2)
        const uint32_t am = __activemask(), ln = threadIdx.x & 31u;
        unsigned long long slot = 0ull;
        if (ln == (uint32_t)__ffs(am) - 1u) slot = atomicAdd(p.rescan, (unsigned long long)__popc(am));
        slot = __shfl_sync(am, slot, __ffs(am) - 1) + __popc(am & ((1u << ln) - 1u));
        if (slot < p.cap) { rq_put(p.rescan, slot, e, frame, s); continue; }
        const float* xf = p.x + frame * p.m.dim;   // list full: rescan here
        for (uint32_t k = k0; k < k1; k++) {
          const double sc = sgmm_exact_score(p.m, xf, k);
          if (sc < best) { best = sc; bk = k - k0; }
        }
      }
    }
    p.out[e] = best;
    if (p.argmin) p.argmin[e] = bk;
    if (p.amin16) p.amin16[e] = (uint16_t)bk;
  }
}

// exact score of one density from a shared-memory table row {mean[dim], norm}: exact_score_unrolled's operation
// order (two running sums over even / odd dimensions, then the odd tail), hence the same bits.  VEC: the feature row is
// 16-byte aligned (even dimensions) and is read as float4.
// The inverse-variance rows of the model, as many as fit (kCivDoubles), in the constant bank: every thread of a
// state-major refinement CTA reads the SAME row, and a warp-uniform LDS.128 still costs two shared-memory wavefronts
// (measured: 33 of the ~200 wavefronts per 32 cells, on the pipe that bounds the kernel); constant-bank operands come
// through the uniform datapath instead.  One copy per device and process: refine_civ_bind hands it from owner to owner.
constexpr uint32_t kCivDoubles = 5632;       // 44 KB of the 64 KB bank
__constant__ double c_iv[kCivDoubles];

// CIV: the inverse variances are c_iv[ivoff ...] instead of iv[...]
template <uint32_t DIM, bool VEC, bool CIV>
__device__ __forceinline__ double sm_exact(const double* row, double logw, const float* xr, const double* iv, uint32_t ivoff) {
  constexpr int NP = DIM / 2;
  static_assert(NP % 4 == 0, "dimension pairs are walked in blocks of four");
  double l0 = 0.0, l1 = 0.0;
  // (d >= 40: unrolled by two blocks only — fully unrolled, the compiler hoists every load of the 24 pairs and spills 400 B)
#pragma unroll(NP > 20 ? 2 : 6)
  for (int g = 0; g < NP; g += 4) {          // four 16-byte units of the mean row in flight
    double2 mu[4];
    float xv[8];
#pragma unroll
    for (int j = 0; j < 4; j++) mu[j] = *reinterpret_cast<const double2*>(row + 2 * (g + j));
    if (VEC) {
      const float4 x0 = *reinterpret_cast<const float4*>(xr + 2 * g), x1 = *reinterpret_cast<const float4*>(xr + 2 * g + 4);
      xv[0] = x0.x; xv[1] = x0.y; xv[2] = x0.z; xv[3] = x0.w; xv[4] = x1.x; xv[5] = x1.y; xv[6] = x1.z; xv[7] = x1.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; j++) xv[j] = xr[2 * g + j];
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int ip = g + j;
      double2 v;
      if (CIV) v = make_double2(c_iv[ivoff + 2 * ip], c_iv[ivoff + 2 * ip + 1]);
      else v = *reinterpret_cast<const double2*>(iv + 2 * ip);
      double a = __dsub_rn((double)xv[2 * j], mu[j].x);
      double b = __dsub_rn((double)xv[2 * j + 1], mu[j].y);
      a = __dmul_rn(a, a); b = __dmul_rn(b, b);
      a = __dmul_rn(a, v.x); b = __dmul_rn(b, v.y);
      l0 = __dadd_rn(l0, a); l1 = __dadd_rn(l1, b);
    }
  }
  double dist = __dadd_rn(l0, l1);
  if (DIM & 1) {
    const double c = __dsub_rn((double)xr[DIM - 1], row[DIM - 1]);
    dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), CIV ? c_iv[ivoff + DIM - 1] : iv[DIM - 1]));
  }
  return __dsub_rn(__dadd_rn(row[DIM], __dmul_rn(dist, 0.5)), logw);
}

// one more density of the same cell, out of line (the second column of a pair verdict: ~8 % of the warps; inlined it
// doubles the evaluation's register footprint on the main path)
template <uint32_t DIM, bool VEC>
__device__ __noinline__ double sm_exact_again(const double* row, double logw, const float* xr, const double* iv) {
  return sm_exact<DIM, VEC, false>(row, logw, xr, iv, 0u);
}

// whole-state rescan of one cell, densities first, first + step, ... (out of line: only taken when the rescan list is full)
template <uint32_t DIM, bool VEC>
__device__ __noinline__ void sm_rescan(const double* tbl, const double* lw, const float* xb, const double* iv, uint32_t nd,
                                       uint32_t RW, uint32_t first, uint32_t step, double& rb_out, uint32_t& rk_out) {
  double rb = 1e10;
  uint32_t rk = 0xffffffffu;
  for (uint32_t k = first; k < nd; k += step) {
    const double sc = sm_exact<DIM, VEC, false>(tbl + (size_t)k * RW, lw[k], xb, iv, 0u);
    if (sc < rb) { rb = sc; rk = k; }
  }
  rb_out = rb; rk_out = rk;
}

// per item of the state-major refinement: where its state's inverse-variance row starts (model-dependent, so not part of
// the corpus's item list)
__global__ void k_sm_ivoff(const uint4* __restrict__ items, uint32_t n, const uint32_t* __restrict__ state_off,
                           const uint32_t* __restrict__ var_of, uint32_t dimP, uint32_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = items[i].x, k0 = state_off[s];
  out[i] = state_off[s + 1] > k0 ? var_of[k0] * dimP : 0u;
}

// State-major refinement (gathered, pooled models).  One CTA per item = (block of utterances, state): the state's
// f64 mean rows + norms (nd x RW doubles, 72 KB at 256 densities, d = 33), its log weights and its inverse-variance
// row are loaded into shared memory ONCE and serve every frame of the block's utterances that visit the state,
// so the per-cell gather of the winner's mean row (272 B) is a shared-memory read instead of an L2 -> SM gather;
// what streams from L2 is the feature rows (132 B per cell, contiguous).
//   * A warp's unit of work is a CHUNK: 32 consecutive frames [32 c, 32 c + 32) of ONE utterance in the state's column.
//     In the tiled emission layout (sgmm_cell_rel) its verdicts are two 64-byte runs and its scores two full 128-byte
//     lines; its feature rows are one contiguous run of global memory (odd dimensions: ONE bulk copy from the
//     16-byte-aligned address below the first row, rows keep their odd stride = conflict-free thread = row reads;
//     dimensions that are multiples of four: one bulk copy per row into a 16-byte-padded stride, read as float4).
//   * Every warp owns TWO tiles: the next chunk's rows and verdicts are in flight while the current chunk is evaluated.
//   * thread = cell evaluates the certified winner with exact_score_unrolled's operation order (same bits); lanes swap
//     cells first so that the 8 lanes of a quarter-warp (one LDS.128 wavefront group) gather table rows from different
//     16-byte bank groups (a row's group is (winner + unit) mod 8).
// Near-tie cells are queued for k_tc_rescan as in the frame-major kernel.
// SUM (sum scores on the tensor path): the cell's input is not a verdict but the filter's candidate masks
// (tc_subpart_mask); every candidate density is evaluated exactly and  -log sum_k exp(-score_k)  is accumulated in density
// order (Mixtures.cpp:719-728) — k_sum_pruned's result with the f32 stage replaced by the tensor filter.
// DENSE: the dense pass (row-major cells, p.sm_dense_stride) — a template parameter because the option costs the
// gathered instantiation 0.1 ms when it is a run-time one.
template <uint32_t DIM, bool CIV, bool SUM = false, bool DENSE = false>   // CIV: the model's inverse-variance rows are in the constant bank (c_iv)
__global__ void __launch_bounds__(512, 1) k_tc_refine_sm(RefineParams p) {
  extern __shared__ __align__(16) unsigned char sm_raw[];
  constexpr uint32_t kMaxEnt = kSmMaxEntries, kMaxChunks = kSmMaxChunks;   // (run-time caps cost the d = 33 instantiation 0.09 ms, measured)
  constexpr bool kOdd = (DIM & 1u) != 0;
  static_assert(kOdd || DIM % 4 == 0, "even dimensions take one 16-byte-aligned bulk copy per row");
  constexpr uint32_t dim = DIM, kRowB = DIM * 4;
  // Even dimensions: rows travel in groups of kGrp rows per bulk copy, 16 bytes of padding behind every group.  A row of
  // d = 48 is twelve 16-byte units: with the padding behind every PAIR of rows the eight rows of a quarter-warp start in
  // eight different bank groups (12 r + (r >> 1) mod 8 = 0 4 1 5 2 6 3 7) and a tile takes 16 copies instead of 32 —
  // the copy-issue loop (one elected lane per copy) was a tenth of the kernel's samples at the stress shape.
  constexpr uint32_t kGrp = (!kOdd && (DIM / 4) % 8 == 4) ? 2u : 1u;
  // tile bytes (the launcher's sm_tile): odd — rows at their own stride + the alignment slack of the one bulk copy
  constexpr uint32_t kTileB = kOdd ? 32 * kRowB + 32 : (32 / kGrp) * (kGrp * kRowB + 16);
  auto row_off = [](uint32_t r) -> uint32_t { return (r / kGrp) * (kGrp * kRowB + 16u) + (r % kGrp) * kRowB; };
  // The constant-bank reads of the chunk loop must come out as LDCU (uniform datapath, immediate offsets): per-thread
  // LDC measured SLOWER than the shared-memory broadcast.  The compiler emits LDCU only inside warp-uniform control flow
  // and from an offset it can prove uniform — a value loaded from a uniform address (sm_ivoff[blockIdx.x]; the chain
  // state -> first density -> variance row is not followed), and a warp index passed through a lane-0 broadcast.
  const uint4 item = p.sm_items[blockIdx.x];
  const uint32_t s = item.x, n_ent = item.z, n_chunks = __shfl_sync(0xffffffffu, item.w, 0);
  const uint32_t k0 = p.state_off[s], nd = __shfl_sync(0xffffffffu, p.state_off[s + 1] - k0, 0);
  const uint32_t RW = p.sm_rw, dimP = p.m.dimP;
  const uint32_t lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const uint32_t nthr = blockDim.x, nwarp = blockDim.x >> 5;
  double* tbl = reinterpret_cast<double*>(sm_raw);           // [nd][RW]
  double* lw = tbl + (size_t)p.sm_maxnd * RW;                // [nd]
  double* iv = lw + ((p.sm_maxnd + 1u) & ~1u);               // [dimP], 16-byte aligned
  uint4* ent_a = reinterpret_cast<uint4*>(iv + dimP);        // [kMaxEnt] {E_off lo, hi, ncol | column << 16, chunks before}
  uint2* ent_b = reinterpret_cast<uint2*>(ent_a + kMaxEnt);  // [kMaxEnt] {first frame, frames}
  uint16_t* chunk_ent = reinterpret_cast<uint16_t*>(ent_b + kMaxEnt);   // [kMaxChunks] entry of a chunk
  unsigned char* tail = reinterpret_cast<unsigned char*>(chunk_ent + kMaxChunks);
  uint64_t* bar = reinterpret_cast<uint64_t*>(tail) + 2 * warp;         // [nwarp][2] one mbarrier per tile
  uint8_t* perm = tail + 16 * nwarp + warp * 32;                         // [nwarp][32] lane permutation
  uint32_t* verdict = reinterpret_cast<uint32_t*>(tail + 48 * nwarp) + warp * 64;   // [nwarp][2][32] GEMM verdicts of a chunk
  unsigned char* tiles = tail + 304 * nwarp + (size_t)warp * 2 * kTileB; // [nwarp][2][kTileB]
  // near-ties of the item: collected here and appended to the global queue as ONE run at the end, so that the rescan
  // kernel finds the cells of a state next to each other (it keeps a state's f32 tables in L1 across a run)
  unsigned long long* lq = reinterpret_cast<unsigned long long*>(tail + 304 * nwarp + (size_t)nwarp * 2 * kTileB);   // [sm_lq] {cell, frame}
  uint32_t* lq_n = reinterpret_cast<uint32_t*>(lq + 2 * p.sm_lq);            // {entries, -, first slot of the run lo, hi}
  uint32_t n_rescans = 0;
  if (p.sm_lq && threadIdx.x == 0) lq_n[0] = 0u;   // (no list: nothing is allocated behind the tiles)

  if (lane == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); }
  for (uint32_t i = threadIdx.x; i < nd * dimP; i += nthr) {
    const uint32_t k = i / dimP, q = i - k * dimP;
    if (q < dim) tbl[(size_t)k * RW + q] = __ldg(p.m.meanR + (size_t)(k0 + k) * dimP + q);
  }
  for (uint32_t k = threadIdx.x; k < nd; k += nthr) {
    tbl[(size_t)k * RW + dim] = __ldg(p.m.norm + k0 + k);
    lw[k] = __ldg(p.m.logw + k0 + k);
  }
  const uint32_t ivoff = p.sm_ivoff[blockIdx.x];   // the state's row of invU / c_iv (k_sm_ivoff); a load from a uniform address
  if (nd > 0)   // (kept in shared memory too: the in-kernel rescan of a full queue reads it there)
    for (uint32_t q = threadIdx.x; q < dimP; q += nthr) iv[q] = __ldg(p.m.invU + (size_t)ivoff + q);
  for (uint32_t i = threadIdx.x; i < n_ent; i += nthr) {
    ent_a[i] = p.sm_ent_a[item.y + i];
    ent_b[i] = p.sm_ent_b[item.y + i];
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < n_chunks; j += nthr) {
    uint32_t lo = 0, hi = n_ent;                             // largest entry with chunks-before <= j
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (ent_a[mid].w <= j) lo = mid; else hi = mid;
    }
    chunk_ent[j] = (uint16_t)lo;
  }
  __syncthreads();

  const unsigned char* xbytes = reinterpret_cast<const unsigned char*>(p.x);
  const uint32_t dstride = DENSE ? p.sm_dense_stride : 0u;
  // A chunk in flight: where its cells are, how many are valid, its first frame and the alignment of its first row.
  // Its verdicts travel by cp.async into the warp's own shared-memory slots: a register destination would have to stay
  // live across the evaluation of the previous chunk, and the compiler spilled it straight away (measured: every chunk
  // then waited a full global-memory latency on the spill store).
  struct Chunk { uint64_t tile0; uint32_t col, nvalid, f0, al; };
  auto issue = [&](uint32_t j, uint32_t b) {
    Chunk ck;
    const uint32_t e = chunk_ent[j];
    const uint4 ea = ent_a[e];
    const uint2 eb = ent_b[e];
    const uint32_t t0 = (j - ea.w) * 32u, ncol = ea.z & 0xffffu;
    ck.col = ea.z >> 16;
    ck.nvalid = min(32u, eb.y - t0);
    ck.f0 = eb.x + t0;                                       // (the launcher takes this kernel below 2^32 frames only)
    // first cell of the chunk's first 16-frame tile, in the state's column; lane l is frame t0 + l:
    //   cell = tile0 + (l >> 4) * 16 ncol + ((l ^ col) & 15)
    // (dense pass: row-major cells, cell = tile0 + l * stride with tile0 = block base + t0 * stride + state)
    ck.tile0 = dstride ? (((uint64_t)ea.y << 32) | ea.x) + (uint64_t)t0 * dstride + ck.col
                       : (((uint64_t)ea.y << 32) | ea.x) + (uint64_t)(t0 >> 4) * (16u * ncol) + (ck.col << 4);
    if (!SUM && lane < ck.nvalid) {
      // (dense: the verdicts are tiled over all states, tiles counted from the pass's first frame)
      const uint64_t cell = dstride ? (uint64_t)(((ck.f0 - p.sm_dense_f0) >> 4) + (lane >> 4)) * (16u * dstride) + (ck.col << 4) + ((lane ^ ck.col) & 15u)
                                    : ck.tile0 + (uint64_t)(lane >> 4) * (16u * ncol) + ((lane ^ ck.col) & 15u);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(verdict + b * 32 + lane)), "l"(p.cand + cell) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    ck.tile0 |= (uint64_t)ncol << 48;                       // (cells stay below 2^48)
    unsigned char* tile = tiles + b * kTileB;
    const size_t src = (size_t)ck.f0 * kRowB;
    ck.al = 0u;
    if (kOdd) {
      ck.al = (uint32_t)src & 15u;
      const uint32_t size = (ck.nvalid * kRowB + ck.al + 15u) & ~15u;
      if (lane == 0) {
        mbar_expect_tx(bar + b, size);
        bulk_load(tile, xbytes + src - ck.al, size, bar + b);
      }
    } else {
      if (lane == 0) mbar_expect_tx(bar + b, ck.nvalid * kRowB);
      __syncwarp();
      if (lane * kGrp < ck.nvalid)
        bulk_load(tile + row_off(lane * kGrp), xbytes + src + (size_t)lane * kGrp * kRowB, min(kGrp, ck.nvalid - lane * kGrp) * kRowB, bar + b);
    }
    return ck;
  };

  uint32_t phases = 0u;                                       // bit b: parity the next wait on tile b expects
  uint32_t b = 0;
  Chunk cur{};
  if (warp < n_chunks) cur = issue(warp, 0);
  for (uint32_t j = warp; j < n_chunks; j += nwarp, b ^= 1u) {
    __syncwarp();                                             // every lane is done with the other tile
    Chunk nxt{};
    const bool more = j + nwarp < n_chunks;
    if (more) nxt = issue(j + nwarp, b ^ 1u);
    const uint32_t ncol = (uint32_t)(cur.tile0 >> 48);
    const uint64_t tile0 = cur.tile0 & 0xffffffffffffull;
    if (SUM) {
      // ---- sum scores: lane = frame (no lane swap: every lane walks its own candidate list)
      const bool act = lane < cur.nvalid;
      const uint64_t my_cell = tile0 + (uint64_t)(lane >> 4) * (16u * ncol) + ((lane ^ cur.col) & 15u);
      const float* my_xr = reinterpret_cast<const float*>(tiles + b * kTileB + (kOdd ? cur.al + lane * kRowB : row_off(lane)));
      uint4 ent[4];
#pragma unroll
      for (int q = 0; q < 4; q++) ent[q] = act ? __ldg(p.summask + 4 * my_cell + q) : make_uint4(0u, 0u, 0x7f800000u, 0u);
      mbar_wait(bar + b, (phases >> b) & 1u);
      phases ^= 1u << b;
      // the lane's candidates: the four sub-parts' masks, walked in density order; a warp iteration evaluates every
      // lane's NEXT candidate (so a chunk takes as many iterations as its busiest cell has candidates)
      unsigned long long bits[4];
#pragma unroll
      for (int q = 0; q < 4; q++)
        bits[q] = (act && 64u * q < nd) ? (((unsigned long long)ent[q].y << 32) | ent[q].x) : 0ull;
      double acc = 0.0;                                       // Mixtures.cpp:721
      while (__any_sync(0xffffffffu, (bits[0] | bits[1] | bits[2] | bits[3]) != 0ull)) {
        const uint32_t q = bits[0] ? 0u : (bits[1] ? 1u : (bits[2] ? 2u : 3u));
        const unsigned long long cur_bits = q == 0u ? bits[0] : (q == 1u ? bits[1] : (q == 2u ? bits[2] : bits[3]));
        const bool has = cur_bits != 0ull;
        const uint32_t k = 64u * q + (has ? (uint32_t)__ffsll((long long)cur_bits) - 1u : 0u);
        const unsigned long long rest = cur_bits & (cur_bits - 1ull);
        bits[0] = q == 0u ? rest : bits[0]; bits[1] = q == 1u ? rest : bits[1];
        bits[2] = q == 2u ? rest : bits[2]; bits[3] = q == 3u ? rest : bits[3];
        const bool ok = has && k < nd;                        // (padding columns of the tile carry no density)
        const uint32_t ki = ok ? k : 0u;
        const double sck = sm_exact<DIM, !kOdd, CIV>(tbl + (size_t)ki * RW, lw[ki], my_xr, iv, ivoff);
        if (ok) acc = __dadd_rn(acc, exp(-sck));
      }
      if (act) {
        p.out[my_cell] = -log(acc);                           // :727
        if (p.amin16) p.amin16[my_cell] = 0;
      }
      cur = nxt;
      continue;
    }
    // this chunk's verdicts have landed (the next chunk's group may still be in flight); each lane reads its own slot
    if (more) asm volatile("cp.async.wait_group 1;" ::: "memory");
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    const uint32_t i1 = (nd && lane < cur.nvalid) ? verdict[b * 32 + lane] : 0u;
    // ---- lanes swap cells: rank the cells by winner mod 8 (ties by lane) and deal the ranks round-robin to the four
    // quarter-warps, so that a quarter's eight table rows start in different 16-byte bank groups where possible
    const bool valid = lane < cur.nvalid;
    uint32_t src;
    {
      const uint32_t key = (valid && (i1 & ~kPair) < (i1 & kPair ? 0x40000000u : nd)) ? (i1 & 7u) : (lane & 7u);   // (a pair: its first column)
      uint32_t lt = 0u, eq = 0xffffffffu;
#pragma unroll
      for (int bit = 2; bit >= 0; bit--) {
        const uint32_t m = __ballot_sync(0xffffffffu, (key >> bit) & 1u);
        if ((key >> bit) & 1u) { lt |= eq & ~m; eq &= m; } else { eq &= ~m; }
      }
      const uint32_t rank = __popc(lt) + __popc(eq & ((1u << lane) - 1u));
      perm[((rank & 3u) << 3) | (rank >> 2)] = (uint8_t)lane;
      __syncwarp();
      src = perm[lane];
    }
    const uint32_t my_i1 = __shfl_sync(0xffffffffu, i1, src);
    const uint64_t my_cell = dstride ? tile0 + (uint64_t)src * dstride
                                     : tile0 + (uint64_t)(src >> 4) * (16u * ncol) + ((src ^ cur.col) & 15u);
    const float* my_xr = reinterpret_cast<const float*>(tiles + b * kTileB + (kOdd ? cur.al + src * kRowB : row_off(src)));
    mbar_wait(bar + b, (phases >> b) & 1u);
    phases ^= 1u << b;
    // the winner's exact score, evaluated by EVERY lane (lanes without a certified winner take row 0 and drop the
    // result): the warp stays converged, so the uniform datapath can feed the constant-bank operands
    // A pair verdict (kPair | a | b << 15, a < b) names the only two columns that can be the argmin: both are evaluated —
    // the second one in a divergent pass that ~8 % of the warps take — and the first minimum wins, as in the reference's scan.
    const bool act = src < cur.nvalid;
    const bool pair = act && (my_i1 & kPair) && !(my_i1 & kUncertain) && (my_i1 & 0x7fffu) < nd && ((my_i1 >> 15) & 0x7fffu) < nd;
    const bool cert = act && (my_i1 < nd || pair);
    uint32_t ki = cert ? (pair ? (my_i1 & 0x7fffu) : my_i1) : 0u;
    double sc = sm_exact<DIM, !kOdd, CIV>(tbl + (size_t)ki * RW, lw[ki], my_xr, iv, ivoff);
    if (pair) {
      const uint32_t kb = (my_i1 >> 15) & 0x7fffu;
      const double sb = sm_exact_again<DIM, !kOdd>(tbl + (size_t)kb * RW, lw[kb], my_xr, iv);
      if (sb < sc) { sc = sb; ki = kb; }               // kb > ki: on equal scores the lower column stays
    }
    if (act) {
      double best = 1e10;                              // min_score's own start, Mixtures.cpp:699
      uint32_t bk = 0;
      bool queued = false;
      if (cert) {
        if (sc < best) { best = sc; bk = ki; }
      } else if (nd) {
        // near-ties are queued for k_tc_rescan (a rescan inside this kernel was measured: one warp walking 256
        // densities holds its CTA back for ~10k cycles, +0.5 ms per 13 M cells)
        const uint32_t ls = p.sm_lq ? atomicAdd(lq_n, 1u) : 0xffffffffu;
        unsigned long long slot = 0ull;
        if (ls < p.sm_lq) {
          lq[2 * ls] = my_cell; lq[2 * ls + 1] = cur.f0 + src;
          queued = true;
        } else if ((slot = atomicAdd(p.rescan, 1ull)) < p.cap) {   // local list full: straight to the global queue
          rq_put(p.rescan, slot, my_cell, cur.f0 + src, s);
          queued = true;
        } else {                                     // both lists full: rescan here
          double rb;
          uint32_t rk;
          sm_rescan<DIM, !kOdd>(tbl, lw, my_xr, iv, nd, RW, 0u, 1u, rb, rk);
          best = rb; bk = rk == 0xffffffffu ? 0u : rk; n_rescans++;
        }
      }
      if (!queued) {
        p.out[my_cell] = best;
        if (p.amin16) p.amin16[my_cell] = (uint16_t)bk;
        if (p.argmin) p.argmin[my_cell] = bk;
      }
    }
    cur = nxt;
  }
  // ---- the item's near-ties: one run of the global queue
  if (p.sm_lq) {
    __syncthreads();
    const uint32_t nq = min(lq_n[0], p.sm_lq);
    if (threadIdx.x == 0 && nq) {
      const unsigned long long base = atomicAdd(p.rescan, (unsigned long long)nq);
      lq_n[2] = (uint32_t)base; lq_n[3] = (uint32_t)(base >> 32);
    }
    __syncthreads();
    const unsigned long long base = ((unsigned long long)lq_n[3] << 32) | lq_n[2];
    for (uint32_t i = threadIdx.x; i < nq; i += nthr) {
      const unsigned long long cell = lq[2 * i], frame = lq[2 * i + 1];
      if (base + i < p.cap) { rq_put(p.rescan, base + i, cell, frame, s); continue; }
      // the global queue is full: rescan here, from the features in global memory
      double rb;
      uint32_t rk;
      sm_rescan<DIM, !kOdd>(tbl, lw, p.x + frame * dim, iv, nd, RW, 0u, 1u, rb, rk);
      p.out[cell] = rb;
      if (p.amin16) p.amin16[cell] = rk == 0xffffffffu ? (uint16_t)0 : (uint16_t)rk;
      if (p.argmin) p.argmin[cell] = rk == 0xffffffffu ? 0u : rk;
      n_rescans++;
    }
  }
  if (n_rescans && p.fallback_count) atomicAdd(p.fallback_count, (unsigned long long)n_rescans);
}

// one warp per queued near-tie cell: exact f64 rescan of the whole state, first minimum wins
__global__ void __launch_bounds__(256) k_tc_rescan(RefineParams p) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t n_done = 0;   // cells this warp finished: ONE atomic per warp at the end (same-address atomics serialise)
  const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const unsigned long long count = min(p.rescan[0], p.cap);
  for (uint64_t i = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < count; i += nwarps) {
    uint64_t e, frame;
    uint32_t s;
    rq_get(p.rescan, i, e, frame, s);
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
      n_done++;
    }
  }
  if (lane == 0 && n_done && p.fallback_count) atomicAdd(p.fallback_count, (unsigned long long)n_done);
}

// Two-stage rescan of the queued near-tie cells, one warp per cell.  Stage 1 scores every density of the state in
// f32 (direct form c + sum h (x - mu)^2 on the FFMA scorer's tables, lane l takes densities l, l + 32, ...) and finds
// the f32 minimum m.  A density whose f32 score exceeds m by more than k_emission_ffma's rigorous bound on the
// evaluation error of both (k_score_ffma.cu, header) has a strictly larger f64 score than the f32 winner and cannot
// be the reference's argmin; stage 2 evaluates only the others (typically two or three) in f64, in the reference's
// operation order, first minimum wins.  Same results as k_tc_rescan with ~1 % of its f64 work and half its bytes.
// Pooled models only (one h row per state).
template <int DV4>
__global__ void __launch_bounds__(256) k_tc_rescan2(RefineParams p) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t n_done = 0;   // cells this warp finished: ONE atomic per warp at the end (same-address atomics serialise)
  const unsigned long long count = min(p.rescan[0], p.cap);
  const float eps = 2.384185791015625e-07f;   // 2^-22, as k_emission_ffma
  const float dterm = (float)(p.m.dim + 8);
  // A CTA takes kRun CONSECUTIVE queue entries at a time (its eight warps walk them together): the state-major
  // refinement appends an item's near-ties as one run, so the entries of a run share a state and its f32 tables (36 KB
  // at 256 x 33) stay in this SM's L1 instead of crossing the L2 once per cell (measured: the kernel was L2-bound).
  constexpr uint64_t kRun = 64;
  const uint64_t wpb = blockDim.x >> 5;
  for (uint64_t i0 = (uint64_t)blockIdx.x * kRun; i0 < count; i0 += (uint64_t)gridDim.x * kRun)
  for (uint64_t i = i0 + (threadIdx.x >> 5); i < min((uint64_t)count, i0 + kRun); i += wpb) {
    uint64_t e, frame;
    uint32_t s;
    rq_get(p.rescan, i, e, frame, s);
    const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
    const float* xf = p.x + frame * p.m.dim;
    float xr[4 * DV4];
#pragma unroll
    for (int q = 0; q < 4 * DV4; q++) xr[q] = q < (int)p.m.dim ? __ldg(xf + q) : 0.f;
    const float4 g = __ldg(p.sguard + s);     // {cmin, cabs, Mmax, -}
    // pooled models only: every density of the state has the same h row -> the first one's, broadcast from L1
    const float4* hrow = p.h32 + (size_t)k0 * DV4;
    constexpr int kKeep = 8;                    // f32 scores of the first 8 rounds stay in registers for stage 2
    float keep[kKeep];
    auto score32 = [&](uint32_t k) {
      float acc = 0.f;
#pragma unroll
      for (int v = 0; v < DV4; v++) {
        const float4 m = __ldg(p.mu32T + (size_t)v * p.m.Dtot + k), h = __ldg(hrow + v);
        float t;
        t = xr[4 * v + 0] - m.x; acc = fmaf(t * h.x, t, acc);
        t = xr[4 * v + 1] - m.y; acc = fmaf(t * h.y, t, acc);
        t = xr[4 * v + 2] - m.z; acc = fmaf(t * h.z, t, acc);
        t = xr[4 * v + 3] - m.w; acc = fmaf(t * h.w, t, acc);
      }
      return acc + __ldg(p.c32 + k);
    };
    float m32 = INFINITY;
#pragma unroll
    for (int j = 0; j < kKeep; j++) {
      const uint32_t k = k0 + lane + 32 * j;
      keep[j] = k < k1 ? score32(k) : INFINITY;
      m32 = fminf(m32, keep[j]);                // fminf drops NaNs; they are kept below
    }
    for (uint32_t k = k0 + lane + 32 * kKeep; k < k1; k += 32) m32 = fminf(m32, score32(k));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m32 = fminf(m32, __shfl_xor_sync(0xffffffffu, m32, o));
    double best = 1e10;
    uint32_t bk = 0xffffffffu;
    // stage 2, per group of 8 rounds (256 densities): ballot the densities the f32 bound cannot exclude, hand the i-th of
    // them to lane i and evaluate all of a group's survivors in ONE divergent pass from the row-major f64 tables
    // (sgmm_exact_score_rm: the reference's operation order).  Evaluating them where they fell — lane by lane, round by
    // round — was eight divergent single-lane passes per cell: ~30 % of the kernel's issued instructions.
    auto keep_test = [&](uint32_t k, float sc) {
      if (k >= k1) return false;
      const float dist = fmaxf(sc - g.x, 0.f);
      const float guard = 2.f * eps * (dterm * dist + 2.f * sqrtf(dist * g.z) + g.y);
      return !((sc - m32) > guard);             // NaN scores are kept
    };
    for (uint32_t gbase = k0; gbase < k1; gbase += 32 * kKeep) {
      uint32_t masks[kKeep];
      uint32_t total = 0;
#pragma unroll
      for (int j = 0; j < kKeep; j++) {
        const uint32_t k = gbase + 32 * j + lane;
        bool take = false;
        if (gbase + 32 * j < k1) take = keep_test(k, gbase == k0 ? keep[j] : (k < k1 ? score32(k) : INFINITY));
        masks[j] = __ballot_sync(0xffffffffu, take);
        total += __popc(masks[j]);
      }
      for (uint32_t done = 0; done < total; done += 32) {
        uint32_t want = done + lane, mine = 0xffffffffu;   // the (done + lane)-th surviving density of the group
        bool found = false;
#pragma unroll
        for (int j = 0; j < kKeep; j++) {
          const uint32_t cnt = __popc(masks[j]);
          if (!found) {
            if (want < cnt) { mine = gbase + 32 * j + __fns(masks[j], 0, want + 1); found = true; }
            else want -= cnt;
          }
        }
        if (mine != 0xffffffffu) {
          const double sd = sgmm_exact_score_rm(p.m, xf, mine);
          if (sd < best || (sd == best && mine - k0 < bk)) { best = sd; bk = mine - k0; }   // first minimum wins
        }
      }
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
      n_done++;
    }
  }
  if (lane == 0 && n_done && p.fallback_count) atomicAdd(p.fallback_count, (unsigned long long)n_done);
}

// ---- tolerance mode (experiment, option "tolerance"; NOT bit-identical to the reference) -------------------------------
// The filter's epilogue left  r = min_k (c_k - sum_d x_d mu_kd / var_d)  per cell (f32, cell-indexed like the emission
// matrix); the state's score is  r + Q,  Q = sum_d x_d^2 / (2 var_sd)  — the x^2 block that pooled variances make common to
// all densities of a state and that the GEMM therefore leaves out.  One CTA per granule of 128 frames of one utterance:
// the frames' rows in shared memory, thread = frame walks the utterance's columns; a warp's 32 frames of a column are
// two 64-byte runs of the tiled layout.  h = 1 / (2 var) comes from the FFMA scorer's f32 table.
template <int DV4>
__global__ void __launch_bounds__(128) k_tol_finish(RefineParams p, float* __restrict__ e32) {
  extern __shared__ __align__(16) float tf_sm[];            // [nu][4 DV4] h rows of the utterance's states, then [128][dim | 1] feature rows
  const uint4 t = p.tasks[blockIdx.x];
  const uint64_t uf0 = p.frame_off[t.x];
  const uint32_t nrows = min(128u, (uint32_t)(p.frame_off[t.x + 1] - uf0) - t.y);
  const uint32_t ncol = p.nuniq_pad[t.x], nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
  const uint32_t* ulist = p.uniq + p.uniq_off[t.x];
  const uint32_t dim = p.m.dim, XS = dim | 1u;
  float* hs = tf_sm;
  float* xs = tf_sm + (size_t)nu * 4 * DV4;
  for (uint32_t i = threadIdx.x; i < nu * 4 * DV4; i += 128) {
    const uint32_t j = i / (4 * DV4), q = i - j * (4 * DV4);
    const uint32_t s = ulist[j], k0 = p.state_off[s];
    // pooled: the state's one h = 1 / (2 var) row (the FFMA scorer's table, zero-padded to 4 DV4); NaN marks an empty mixture
    hs[i] = p.state_off[s + 1] > k0 ? __ldg(reinterpret_cast<const float*>(p.h32 + (size_t)k0 * DV4) + q) : __int_as_float(0x7fc00000);
  }
  for (uint32_t i = threadIdx.x; i < nrows * dim; i += 128) {
    const uint32_t r = i / dim, q = i - r * dim;
    xs[r * XS + q] = __ldg(p.x + (uf0 + t.y + r) * dim + q);
  }
  __syncthreads();
  const uint32_t r = threadIdx.x;
  if (r >= nrows) return;
  float x2[4 * DV4];
#pragma unroll
  for (int d = 0; d < 4 * DV4; d++) { const float v = d < (int)dim ? xs[r * XS + d] : 0.f; x2[d] = v * v; }
  float* out = e32 + p.E_off[t.x] + (uint64_t)t.y * ncol;   // t.y is a multiple of 16: first cell of its tile
  for (uint32_t j = 0; j < nu; j++) {
    const float4* h = reinterpret_cast<const float4*>(hs + (size_t)j * 4 * DV4);
    float q = 0.f;
#pragma unroll
    for (int v = 0; v < DV4; v++) {
      const float4 hv = h[v];
      q = fmaf(x2[4 * v], hv.x, q); q = fmaf(x2[4 * v + 1], hv.y, q); q = fmaf(x2[4 * v + 2], hv.z, q); q = fmaf(x2[4 * v + 3], hv.w, q);
    }
    const uint64_t cell = sgmm_cell_rel(ncol, r, j);
    out[cell] = q == q ? out[cell] + q : 1e10f;             // empty mixture: min_score's start value
  }
}

// ---- sum scores with an f32 pruner (option max_approx = 0) ---------------------------------------------
// score = -log sum_k exp(-s_k) (Mixtures.cpp:719-728).  A density whose score exceeds the state's minimum by more than
// kSumDelta = 45 adds less than e^-45 = 2.9e-20 of the sum; 65535 of them stay below a quarter of an f64 ulp, so leaving them
// out changes the result by less than one rounding (and by nothing at all when the minimum is >= 700: their exp
// underflows to exactly 0).  One warp per (frame, state) cell: stage 1 scores every density in f32 (lanes =
// densities, the FFMA scorer's tables) and takes the minimum m; a density with  f32 score - m > 45 + guard  (guard =
// k_emission_ffma's rigorous bound on the evaluation error of both scores) is certainly more than 45 above the true
// minimum; stage 2 evaluates the others in f64 (the reference's operation order) and adds their exp(-s) in mixture
// order.  NaN scores fail the test and are evaluated.
constexpr float kSumDelta = 45.0f;

template <int DV4>
__global__ void __launch_bounds__(256, 3) k_sum_pruned(RefineParams p) {
  // One CTA per granule of 32 consecutive frames (of one utterance, or of the dense range); the eight warps walk
  // the granule's states TOGETHER (warp w takes frames w, w + 8, ...), so a state's f32 table (36 KB at 256 x 33) is
  // pulled into L1 once per 32 cells instead of once per cell.
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float eps = 2.384185791015625e-07f;   // 2^-22, as k_emission_ffma
  const float dterm = (float)(p.m.dim + 8);
  uint64_t fbase, obase;
  uint32_t nvalid, nu, ostride;
  const uint32_t* ulist = nullptr;
  if (p.dense == 2) {              // every frame against its aligned state (calc_am_score): out[frame]
    const uint64_t r0 = (uint64_t)blockIdx.x * 32;
    fbase = r0;
    nvalid = (uint32_t)min((uint64_t)32, p.n_cells - r0);
    nu = 1; ostride = 1; obase = r0;
  } else if (p.dense) {
    const uint64_t r0 = (uint64_t)blockIdx.x * 32;
    fbase = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)32, p.n_cells / p.n_states - r0);
    nu = p.n_states; ostride = p.n_states; obase = r0 * p.n_states;
  } else {
    const uint4 t = p.tasks[blockIdx.x];
    const uint64_t uf0 = p.frame_off[t.x];
    fbase = uf0 + t.y;
    nvalid = min(32u, (uint32_t)(p.frame_off[t.x + 1] - uf0) - t.y);
    ulist = p.uniq + p.uniq_off[t.x];
    nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
    ostride = p.nuniq_pad[t.x];
    obase = p.E_off[t.x] + (uint64_t)t.y * ostride;
  }
  for (uint32_t cell = warp; cell < nu * 4 * 8; cell += 8) {
    const uint32_t j = cell / 32, fr = cell % 32;       // all warps share j: fr = warp, warp + 8, warp + 16, warp + 24
    if (fr >= nvalid) continue;
    const uint64_t frame = fbase + fr, e = obase + (p.dense ? (uint64_t)fr * ostride + j : sgmm_cell_rel(ostride, fr, j));
    const uint32_t s = p.dense == 2 ? min((uint32_t)p.ali[frame].state, p.n_states - 1) : (ulist ? ulist[j] : j);
    const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
    const float* xf = p.x + frame * p.m.dim;
    float xr[4 * DV4];
#pragma unroll
    for (int q = 0; q < 4 * DV4; q++) xr[q] = q < (int)p.m.dim ? __ldg(xf + q) : 0.f;
    const float4 g = __ldg(p.sguard + s);     // {cmin, cabs, Mmax, -}
    auto score32 = [&](uint32_t k) {
      const float4* hrow = p.h32 + (size_t)(p.pooled ? k0 : k) * DV4;   // pooled: the state's one row, broadcast
      float acc = 0.f;
#pragma unroll
      for (int v = 0; v < DV4; v++) {
        const float4 m = __ldg(p.mu32T + (size_t)v * p.m.Dtot + k), h = __ldg(hrow + v);
        float t;
        t = xr[4 * v + 0] - m.x; acc = fmaf(t * h.x, t, acc);
        t = xr[4 * v + 1] - m.y; acc = fmaf(t * h.y, t, acc);
        t = xr[4 * v + 2] - m.z; acc = fmaf(t * h.z, t, acc);
        t = xr[4 * v + 3] - m.w; acc = fmaf(t * h.w, t, acc);
      }
      return acc + __ldg(p.c32 + k);
    };
    constexpr int kKeep = 8;                    // f32 scores of the first 8 rounds stay in registers for stage 2
    float keep[kKeep];
    float m32 = INFINITY;
#pragma unroll
    for (int j = 0; j < kKeep; j++) {
      const uint32_t k = k0 + lane + 32 * j;
      keep[j] = k < k1 ? score32(k) : INFINITY;
      m32 = fminf(m32, keep[j]);
    }
    for (uint32_t k = k0 + lane + 32 * kKeep; k < k1; k += 32) m32 = fminf(m32, score32(k));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m32 = fminf(m32, __shfl_xor_sync(0xffffffffu, m32, o));
    double acc = 0.0;                           // Mixtures.cpp:721
    // stage 2, per group of 8 rounds (256 densities, in density order): ballot the kept densities, hand the i-th kept
    // one to lane i (all of a group's survivors are evaluated in ONE divergent pass: row-major f64 mean rows, the
    // reference's operation order), then add their exp(-s) in density order through shuffles
    auto keep_test = [&](uint32_t k, float sc) {
      if (k >= k1) return false;
      const float dist = fmaxf(sc - g.x, 0.f);
      const float guard = 2.f * eps * (dterm * dist + 2.f * sqrtf(dist * g.z) + g.y);
      return !((sc - m32) > kSumDelta + guard);   // NaN scores are kept
    };
    for (uint32_t gbase = k0; gbase < k1; gbase += 32 * kKeep) {
      uint32_t masks[kKeep];
      uint32_t total = 0;
#pragma unroll
      for (int j = 0; j < kKeep; j++) {
        const uint32_t k = gbase + 32 * j + lane;
        bool take = false;
        if (gbase + 32 * j < k1) take = keep_test(k, gbase == k0 ? keep[j] : (k < k1 ? score32(k) : INFINITY));
        masks[j] = __ballot_sync(0xffffffffu, take);
        total += __popc(masks[j]);
      }
      for (uint32_t done = 0; done < total; done += 32) {
        // the (done + lane)-th kept density of the group
        uint32_t want = done + lane, mine = 0xffffffffu;
        bool found = false;
#pragma unroll
        for (int j = 0; j < kKeep; j++) {
          const uint32_t c = __popc(masks[j]);
          if (!found) {
            if (want < c) { mine = gbase + 32 * j + __fns(masks[j], 0, want + 1); found = true; }
            else want -= c;
          }
        }
        double ev = 0.0;
        if (mine != 0xffffffffu) ev = exp(-sgmm_exact_score_rm(p.m, xf, mine));
        const uint32_t nb = min(32u, total - done);
        for (uint32_t l = 0; l < nb; l++) acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, ev, l));
      }
    }
    if (lane == 0) {
      p.out[e] = -log(acc);                     // :727
      if (p.argmin) p.argmin[e] = 0;
      if (p.amin16) p.amin16[e] = 0;
    }
  }
}

}  // namespace sgmm_tc

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------

// c_iv is one table per device and process, shared by every context.  A context that finds another owner's rows there
// (or an older model's) replaces them — after the last kernel that read the old rows (an event recorded behind every
// launch that uses the table) — on its own stream, in front of its launch.  The mutex is held from bind to release so
// that bind + launch + event are one step.
namespace {
using sgmm_tc::c_iv;
using sgmm_tc::kCivDoubles;
struct CivSlot {
  std::mutex mu;
  uint64_t owner = 0, version = 0;
  cudaEvent_t last_use = nullptr;
  bool used = false;
};
constexpr int kCivDevices = 64;
CivSlot g_civ[kCivDevices];

int refine_civ_bind(sgmm_ctx* c, bool* civ) {
  *civ = false;
  const size_t n = (size_t)c->n_var_rows * c->dimP;
  if (!c->tc_civ || n == 0 || n > kCivDoubles || c->device < 0 || c->device >= kCivDevices) return 0;
  CivSlot& g = g_civ[c->device];
  g.mu.lock();
  if (!g.last_use && cudaEventCreateWithFlags(&g.last_use, cudaEventDisableTiming) != cudaSuccess) {
    g.mu.unlock();
    SGMM_FAIL(c, "refine: cannot create the constant-table event");
  }
  if (g.owner != c->serial || g.version != c->model_version) {
    cudaError_t e = g.used ? cudaStreamWaitEvent(c->stream, g.last_use, 0) : cudaSuccess;
    if (e == cudaSuccess)
      e = cudaMemcpyToSymbolAsync(c_iv, c->d_invU.p, n * sizeof(double), 0, cudaMemcpyDeviceToDevice, c->stream);
    if (e != cudaSuccess) {
      g.owner = 0;
      g.mu.unlock();
      SGMM_FAIL(c, std::string("refine: constant-table upload: ") + cudaGetErrorString(e));
    }
    g.owner = c->serial; g.version = c->model_version;
  }
  *civ = true;
  return 0;
}

int refine_civ_release(sgmm_ctx* c) {
  CivSlot& g = g_civ[c->device];
  const cudaError_t e = cudaEventRecord(g.last_use, c->stream);
  g.used = true;
  g.mu.unlock();
  if (e != cudaSuccess) SGMM_FAIL(c, std::string("refine: constant-table event: ") + cudaGetErrorString(e));
  return 0;
}
}  // namespace
namespace {

// shared memory: Alo + the B ring (Ahi lives in tensor memory)
constexpr size_t kSmemLimit = 227 * 1024;
// A (hi, lo) of every half + half-stage slots + barriers + exchange (three times as large with sixteen epilogue warps per
// unit) + the per-state metadata cache
size_t tc_smem256(uint32_t ks, int nhalf, int nslot = 3, bool epi16 = true) {
  return 1024 + (size_t)nhalf * 2 * ks * 4096 + (size_t)nslot * ks * 8192 + 256 +
         (epi16 ? 3 : 1) * (2 * 2 * 128 * 5 * sizeof(float)) + sgmm_tc::kMetaCache * sizeof(sgmm_tc::TcMeta);
}
size_t tc_smem(uint32_t stage_bytes, int nhalf, int nstage, int a_smem) {   // A of every half + the B ring + barriers
  return 1024 + (size_t)nhalf * (stage_bytes / 2) * (a_smem == 1 ? 2 : 1) + (size_t)nstage * stage_bytes + 256 +
         2 * 2 * 128 * 5 * sizeof(float);
}

}  // namespace

int sgmm_launch_emission_tc(sgmm_ctx* c, bool dense, uint64_t dense_f0, uint64_t dense_nframes, double* d_scores,
                            uint32_t* d_argmin) {
  using namespace sgmm_tc;
  if (!c->tc_ok) SGMM_FAIL(c, "tensor-core scorer not available for this model shape");
  const bool wide = c->tc_tile_rows == 256;   // N = 256 kernel (B image built for it by sgmm_set_model)
  const bool f16 = wide && c->tc_f16_ok && c->tc_f16 != 0;   // fp16 hi / lo operands: 16 terms per k-step
  const uint32_t ks = f16 ? c->tc_ks16 : c->tc_ksteps, KP = f16 ? 8 * ks : c->tc_KP;   // KP: 4-byte words per A row
  // two frame halves per CTA when their Ahi operands (KP columns each) fit the 128 TMEM columns set aside
  // (wide kernel: when both halves' A parts fit shared memory next to the ring)
  const int nhalf = wide ? ((c->tc_halves == 1 || tc_smem256(ks, 2) > kSmemLimit) ? 1 : 2)
                         : ((c->tc_halves == 1 || 2 * KP > 128) ? 1 : 2);
  const uint32_t stage_bytes = wide ? ks * 8192 : 2 * ks * 128 * 32;
  // both A parts in shared memory when that still leaves three B stages (option tc_debug & 64 forces the TMEM form)
  int a_smem = (c->tc_debug >> 6) & 3;   // experiment: tc_debug bits 6-7 select where the A parts live
  if (a_smem == 1 && tc_smem(stage_bytes, nhalf, 2, 1) > kSmemLimit) a_smem = 0;
  const bool epi16 = f16 && ks == 3 && c->tc_epi16;   // the one instantiation with sixteen epilogue warps per unit
  int nstage = 1;
  while (nstage < sgmm_tc::kMaxStage && tc_smem(stage_bytes, nhalf, nstage + 1, a_smem) <= kSmemLimit) nstage++;
  if (wide) {
    nstage = 3;
    while (nstage < sgmm_tc::kMaxStage && tc_smem256(ks, nhalf, nstage + 1, epi16) <= kSmemLimit) nstage++;
    if (!f16) nstage = 3;   // (the tf32 form was tuned with three slots)
  }
  const uint32_t rows_per_task = 128u * nhalf;
  // Work is cut into frame ranges.  Normally there is one; while a chunked feature upload is in flight
  // (option "async_upload") there is one per chunk, each launched behind that chunk's copy event, so that the
  // GEMM of the first chunks runs under the copies of the later ones.
  struct Range { uint64_t f_end; int ev; };
  std::vector<Range> ranges;
  if (c->feat_pending && !c->tc_A_valid) {   // (dense pass too: recognition right behind an upload)
    for (int i = 0; i < c->feat_pending; i++) ranges.push_back({c->feat_chunk_end[i], i});
  } else {
    if (sgmm_wait_features(c)) return 1;
    ranges.push_back({c->n_frames, -1});
  }
  c->feat_pending = 0;
  // the A rows (split / scaled features) depend on the features and on the operand form, never on the parameter
  // values: a new model of the same shape (every EM iteration) reuses them
  float eps_now = f16 ? c->tc_eps16 : c->tc_eps;
  uint32_t eps_bits;
  memcpy(&eps_bits, &eps_now, 4);
  const uint64_t a_key = ((uint64_t)KP << 40) ^ ((uint64_t)(f16 ? 1 : 0) << 36) ^ ((uint64_t)(c->tc_full ? 1 : 0) << 37) ^
                         ((uint64_t)c->dim << 44) ^ ((uint64_t)(f16 ? c->tc_rem16 : 0u) << 56) ^ eps_bits;
  const bool need_prep = !c->tc_A_valid || c->tc_A_key != a_key;
  c->tc_A_key = a_key;
  c->tc_A_form = f16 ? 1 : 0;
  if (need_prep) {
    SGMM_CHECK(c, c->d_Ahi.reserve((size_t)c->n_frames * KP));
    SGMM_CHECK(c, c->d_Alo.reserve((size_t)c->n_frames * KP));
    SGMM_CHECK(c, c->d_xnorm.reserve(c->n_frames));
    c->tc_A_valid = true;
  }
  const uint64_t n_cells = dense ? dense_nframes * c->n_states : c->E_elems;
  const uint64_t n_tasks = dense ? (dense_nframes + rows_per_task - 1) / rows_per_task
                                 : (nhalf == 2 ? c->n_tasks256 : c->n_tasks128);
  if (n_tasks == 0) return 0;
  SGMM_CHECK(c, c->d_ci1.reserve(dense ? (dense_nframes + 15) / 16 * 16 * c->n_states : n_cells));   // (verdicts: whole 16-frame tiles)
  const size_t cap = (size_t)std::max<uint64_t>(1024, n_cells / 8);
  SGMM_CHECK(c, c->d_rescan.reserve(2 * cap + 2));   // [0] entries, [1] -, then {cell, frame | state << 40} per entry
  c->rescan_cap = cap;
  SGMM_CHECK(c, cudaMemsetAsync(c->d_rescan.p, 0, sizeof(unsigned long long), c->stream));

  TcParams p{};
  p.ksteps = ks; p.nstage = (uint32_t)nstage; p.stage_bytes = stage_bytes; p.nhalf = (uint32_t)nhalf;
  p.debug = (uint32_t)c->tc_debug; p.a_smem = (uint32_t)a_smem;
  p.krem = (f16 && c->tc_rem16) ? 1u : 0u;
  p.prof = (c->tc_debug & 32) ? c->d_fallbacks.p : nullptr;
  p.Ahi = c->d_Ahi.p; p.Alo = c->d_Alo.p; p.n_frames = c->n_frames;
  p.Bimg = c->d_Bimg.p; p.tile_first = c->d_tile_first.p; p.tile_info = c->d_tile_info.p;
  p.xnorm = c->d_xnorm.p; p.sguard = c->d_sguard_tc.p; p.eps = c->tc_eps;
  if (f16) {
    p.Bimg = reinterpret_cast<const float*>(c->d_Bimg16.p); p.tile_info = c->d_tile_info16.p;
    p.sguard = c->d_sguard16.p; p.eps = c->tc_eps16;
  }
  p.n_tasks = n_tasks;
  p.tasks = nhalf == 2 ? c->d_tasks256.p : c->d_tasks128.p;
  p.frame_off = c->d_frame_off.p; p.uniq = c->d_uniq.p; p.uniq_off = c->d_uniq_off.p;
  p.nuniq_pad = c->d_nuniq_pad.p; p.E_off = c->d_E_off.p;
  p.pack_desc = c->d_pack_desc.p; p.pack_uniq = c->d_pack_uniq.p; p.pack_rows = c->d_pack_rows.p;
  p.pack_colmap = c->d_pack_colmap.p;
  p.dense = dense ? 1 : 0; p.f0 = dense_f0; p.nframes = dense_nframes; p.n_states = c->n_states;
  p.cand = c->d_ci1.p;
  p.pairs = c->tc_pairs ? 1 : 0;
  if (c->tol && !dense && !c->tc_sum_pass) {
    if (!(f16 && !epi16 && !c->tc_full)) SGMM_FAIL(c, "tolerance mode: needs the fp16 tensor form (pooled model, 256-column tiles, tc_epi16 = 0)");
    SGMM_CHECK(c, c->d_e32.reserve(n_cells + 2));
    p.tol = 1; p.e32 = c->d_e32.p;
  }
  const size_t smem = wide ? tc_smem256(ks, nhalf, nstage, epi16) : tc_smem(stage_bytes, nhalf, nstage, a_smem);
  RefineParams r{};
  r.m.meanT = c->d_meanT.p; r.m.invT = c->d_invT.p; r.m.mean_tail = c->d_mean_tail.p; r.m.inv_tail = c->d_inv_tail.p;
  r.m.norm = c->d_norm.p; r.m.logw = c->d_logw.p; r.m.dim = c->dim; r.m.npair = c->npair; r.m.Dtot = c->Dtot;
  r.m.meanR = c->d_meanR.p; r.m.invU = c->d_invU.p; r.m.var_of = c->d_var_of.p; r.m.dimP = c->dimP;
  r.state_off = c->d_state_off.p; r.x = c->d_x.p; r.cand = c->d_ci1.p; r.n_cells = n_cells;
  r.dense = dense ? 1 : 0; r.f0 = dense_f0; r.n_states = c->n_states; r.n_utts = c->n_utts;
  r.frame_off = c->d_frame_off.p; r.E_off = c->d_E_off.p; r.uniq = c->d_uniq.p; r.uniq_off = c->d_uniq_off.p;
  r.nuniq_pad = c->d_nuniq_pad.p;
  r.out = dense ? d_scores : c->d_E.p; r.argmin = dense ? d_argmin : nullptr;
  r.amin16 = dense ? nullptr : c->d_Amin.p;
  r.rescan = c->d_rescan.p; r.cap = cap; r.fallback_count = c->d_fallbacks.p;
  const uint4* gemm_tasks = nhalf == 2 ? c->d_tasks256.p : c->d_tasks128.p;
  const std::vector<uint64_t>& gemm_end = nhalf == 2 ? c->h_t256_end : c->h_t128_end;
  r.pooled = c->tc_full ? 0 : 1;
  // refinement shared memory: 128 feature rows + (pooled) one inverse-variance row per state of the granule
  const size_t rsm = 128 * (2 * (size_t)c->npair + 2) * sizeof(float) +
                     (r.pooled ? (size_t)(dense ? c->n_states : c->max_uniq) * c->dimP * sizeof(double) : 0);
  if (rsm > 200 * 1024) r.pooled = 0;
  const size_t rsm_used = r.pooled ? rsm : 128 * (2 * (size_t)c->npair + 2) * sizeof(float);
  // state-major refinement: gathered, pooled, unrolled dimensions, and the widest state's table fits shared memory
  uint32_t sm_rw = c->dim + 1;
  while ((sm_rw & 1) || !((sm_rw >> 1) & 1)) sm_rw++;        // 2 * odd doubles per row
  uint32_t sm_maxnd = 0;
  for (uint32_t s = 0; s + 1 < c->h_state_off.size(); s++) sm_maxnd = std::max(sm_maxnd, c->h_state_off[s + 1] - c->h_state_off[s]);
  // one CTA per SM: as many warps (each with its own PAIR of 32-row feature tiles, mbarriers and lane permutation) as
  // fit next to the table — the carve-up of k_tc_refine_sm
  const size_t sm_fixed = ((size_t)sm_maxnd * sm_rw + ((sm_maxnd + 1u) & ~1u) + c->dimP) * sizeof(double) +
                          kSmMaxEntries * (sizeof(uint4) + sizeof(uint2)) + kSmMaxChunks * sizeof(uint16_t);
  const uint32_t sm_grp = (!(c->dim & 3u) && (c->dim / 4) % 8 == 4) ? 2u : 1u;   // rows per bulk copy (kGrp of k_tc_refine_sm)
  const size_t sm_tile = (c->dim & 1u) ? 32 * (size_t)c->dim * sizeof(float) + 32
                                       : (32 / sm_grp) * (sm_grp * (size_t)c->dim * sizeof(float) + 16);
  const size_t sm_warp = 304 + 2 * sm_tile;
  int sm_warps = sm_fixed + 8 * sm_warp <= kSmemLimit ? (int)std::min<size_t>(16, (kSmemLimit - sm_fixed) / sm_warp) : 0;
  const size_t sm_tiles = sm_fixed + (size_t)sm_warps * sm_warp;
  const size_t lq_want = c->tc_lq <= 0 ? 0 : (c->tc_lq == 1 ? 1024 : (size_t)c->tc_lq);
  const uint32_t sm_lq = (lq_want && sm_tiles + 16 + 64 * 16 <= kSmemLimit) ? (uint32_t)std::min<size_t>(lq_want, (kSmemLimit - sm_tiles - 16) / 16) : 0u;
  const size_t sm_smem = sm_tiles + (sm_lq ? (size_t)sm_lq * 16 + 16 : 0);
  const bool sm_dim = c->dim == 24 || c->dim == 25 || c->dim == 32 || c->dim == 33 || c->dim == 48 || c->dim == 49;
  const bool sm_can = !dense && !c->tc_full && sm_dim && sm_warps >= 8 && c->n_sm_items > 0 && c->n_frames < 0xffffffffull &&
                      c->max_uniq < 0xffffu;
  // dense pass (sgmm_score_all, recognition): the same kernel over items built here — (block of <= 64 k frames, state),
  // one entry each, row-major cells — instead of the frame-major gather (272 B of L2 traffic per cell)
  const bool sm_dense = dense && !c->tc_full && sm_dim && sm_warps >= 8 && c->tc_refine != 1 && c->tc_dense_sm &&
                        dense_f0 + dense_nframes < 0xffffffffull && c->n_states < 0xffffu;
  const bool use_sm = (sm_can && c->tc_refine != 1) || sm_dense;
  if (c->tc_sum_pass && !(f16 && !epi16 && nhalf == 2 && sm_can && sm_maxnd <= 256)) {
    c->tc_sum_pass = false;   // not this shape: the caller falls back to k_sum_pruned (nothing has been launched yet)
    return 0;
  }
  if (c->tc_sum_pass) {       // 64 B per cell: four {mask, minimum, width} entries
    SGMM_CHECK(c, c->d_summask.reserve(4 * (n_cells + 2)));
    p.summask = c->d_summask.p;
  }
  r.sm_items = c->d_sm_items.p; r.sm_ent_a = c->d_sm_ent_a.p; r.sm_ent_b = c->d_sm_ent_b.p;
  r.sm_rw = sm_rw; r.sm_maxnd = sm_maxnd; r.sm_lq = sm_lq;
  const bool one = ranges.size() == 1;
  sgmm_stage_begin(c, SGMM_T_SCORE_FILTER);
  uint64_t f_prev = 0, g_prev = 0;
  // several ranges: each on its own side stream, so that the next range's CTAs fill the SMs that the tail of the
  // previous one leaves idle; all of them rejoin the main stream before the rescan
  if (!one) SGMM_CHECK(c, cudaEventRecord(c->ev_fork, c->stream));
  static_assert(sgmm_ctx::kFeatChunks <= sgmm_ctx::kAux, "one side stream per upload chunk");
  for (size_t ri = 0; ri < ranges.size(); ri++) {
    const Range& rg = ranges[ri];
    cudaStream_t st = one ? c->stream : c->aux[ri];
    if (!one) SGMM_CHECK(c, cudaStreamWaitEvent(st, c->ev_fork, 0));
    if (rg.ev >= 0) SGMM_CHECK(c, cudaStreamWaitEvent(st, c->feat_ev[rg.ev], 0));
    if (need_prep && rg.f_end > f_prev) {
      const uint64_t nf = rg.f_end - f_prev;
      if (f16)
        k_tc_prep16<<<(unsigned)((nf + 7) / 8), 256, 0, st>>>(c->d_x.p + f_prev * c->dim, nf, c->dim, 2 * KP, c->tc_rem16, 0.5f / c->tc_eps16,
                                                                    reinterpret_cast<__half*>(c->d_Ahi.p + f_prev * KP),
                                                                    reinterpret_cast<__half*>(c->d_Alo.p + f_prev * KP),
                                                                    c->d_xnorm.p + f_prev);
      else
        k_tc_prep<<<(unsigned)((nf + 7) / 8), 256, 0, st>>>(c->d_x.p + f_prev * c->dim, nf, c->dim, c->tc_full, KP,
                                                                  c->d_Ahi.p + f_prev * KP, c->d_Alo.p + f_prev * KP,
                                                                  c->d_xnorm.p + f_prev);
      SGMM_LAUNCH_CHECK(c);
    }
    f_prev = rg.f_end;
    // tasks whose frames have all arrived
    // (dense: task i is frames dense_f0 + [i * rows, + rows); a range takes the whole tasks below its last frame)
    const uint64_t g_end = !dense ? std::upper_bound(gemm_end.begin(), gemm_end.end(), rg.f_end) - gemm_end.begin()
                           : (one || rg.f_end >= dense_f0 + dense_nframes) ? n_tasks
                           : (rg.f_end <= dense_f0 ? 0 : (rg.f_end - dense_f0) / rows_per_task);
    if (g_end > g_prev) {
      const unsigned grid = (unsigned)(g_end - g_prev);
      p.tasks = dense ? nullptr : gemm_tasks + g_prev;
      if (dense) {   // the kernel counts dense tasks from blockIdx.x = 0: shift its origin instead
        p.f0 = dense_f0 + g_prev * rows_per_task;
        p.nframes = dense_nframes - g_prev * rows_per_task;
        p.cand = c->d_ci1.p + g_prev * rows_per_task * c->n_states;
      }
      if (epi16) {   // d = 32..46 pooled, fp16 operands, sixteen epilogue warps per unit
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<3, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm256<3, true, true><<<grid, 576, smem, st>>>(p);
      } else if (f16 && ks == 3) {   // d = 32..46 pooled, fp16 operands
        if (p.summask) {
          SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<3, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          k_tc_gemm256<3, true, false, true><<<grid, 576, smem, st>>>(p);
        } else {
          SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          k_tc_gemm256<3, true><<<grid, 576, smem, st>>>(p);
        }
      } else if (f16) {
        if (p.summask) {
          SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<0, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          k_tc_gemm256<0, true, false, true><<<grid, 576, smem, st>>>(p);
        } else {
          SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          k_tc_gemm256<0, true><<<grid, 576, smem, st>>>(p);
        }
      } else if (wide && ks == 5) {
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<5, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm256<5, false><<<grid, 576, smem, st>>>(p);
      } else if (wide) {
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm256<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm256<0, false><<<grid, 576, smem, st>>>(p);
      } else if (ks == 5) {          // d = 33 pooled (BASELINE configs 2-4)
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm<5><<<grid, 576, smem, st>>>(p);
      } else if (ks == 7) {   // d = 48 pooled (config 5), d = 25 unpooled
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm<7><<<grid, 576, smem, st>>>(p);
      } else {
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_gemm<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tc_gemm<0><<<grid, 576, smem, st>>>(p);
      }
      SGMM_LAUNCH_CHECK(c);
      g_prev = g_end;
    }
    if (!one) SGMM_CHECK(c, cudaEventRecord(c->ev_join[ri], st));
  }
  if (!one)
    for (size_t ri = 0; ri < ranges.size(); ri++) SGMM_CHECK(c, cudaStreamWaitEvent(c->stream, c->ev_join[ri], 0));
  sgmm_stage_end(c, SGMM_T_SCORE_FILTER);
  if (c->tc_sum_pass) {
    // every candidate of every cell evaluated exactly, summed in density order: the state-major kernel in SUM form
    sgmm_stage_begin(c, SGMM_T_SCORE_REFINE);
    const unsigned sgrid = (unsigned)c->n_sm_items;
    SGMM_CHECK(c, c->d_sm_ivoff.reserve(sgrid));
    k_sm_ivoff<<<(sgrid + 255) / 256, 256, 0, c->stream>>>(c->d_sm_items.p, sgrid, c->d_state_off.p, c->d_var_of.p, c->dimP, c->d_sm_ivoff.p);
    SGMM_LAUNCH_CHECK(c);
    r.sm_ivoff = c->d_sm_ivoff.p; r.summask = c->d_summask.p; r.sm_lq = 0;
    const size_t sum_smem = sm_tiles;
    bool civ = false;
    if (refine_civ_bind(c, &civ)) return 1;
#define SGMM_SUM_CASE(D)                                                                                                  \
  case D:                                                                                                                 \
    if (civ) {                                                                                                            \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sum_smem));  \
      k_tc_refine_sm<D, true, true><<<sgrid, 32 * sm_warps, sum_smem, c->stream>>>(r);                                    \
    } else {                                                                                                              \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sum_smem)); \
      k_tc_refine_sm<D, false, true><<<sgrid, 32 * sm_warps, sum_smem, c->stream>>>(r);                                   \
    }                                                                                                                     \
    break;
    switch (c->dim) {
      SGMM_SUM_CASE(24) SGMM_SUM_CASE(25) SGMM_SUM_CASE(32) SGMM_SUM_CASE(33) SGMM_SUM_CASE(48) SGMM_SUM_CASE(49)
      default: if (civ) refine_civ_release(c); SGMM_FAIL(c, "sum scores on the tensor path: unsupported dimension");
    }
#undef SGMM_SUM_CASE
    SGMM_LAUNCH_CHECK(c);
    if (civ && refine_civ_release(c)) return 1;
    sgmm_stage_end(c, SGMM_T_SCORE_REFINE);
    return 0;
  }
  if (p.tol) {
    // tolerance mode: e32 holds f32 scores minus the x^2 term; add it, and that is the emission matrix Viterbi reads
    // (the verdicts stay in d_ci1: the E-step looks the aligned cells' winners up there, k_rescore_verdict)
    sgmm_stage_begin(c, SGMM_T_SCORE_REFINE);
    r.tasks = c->d_tasks128.p;
    r.h32 = c->d_h32.p;
    const size_t fsm = ((size_t)c->max_uniq * 4 * c->dv4 + 128 * (size_t)(c->dim | 1u)) * sizeof(float);
    if (fsm > kSmemLimit) SGMM_FAIL(c, "tolerance mode: too many distinct states in one utterance");
#define SGMM_TOL_CASE(V)                                                                                        \
  case V:                                                                                                       \
    SGMM_CHECK(c, cudaFuncSetAttribute(k_tol_finish<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm)); \
    k_tol_finish<V><<<(unsigned)c->n_tasks128, 128, fsm, c->stream>>>(r, c->d_e32.p);                            \
    break;
    switch (c->dv4) {
      SGMM_TOL_CASE(7) SGMM_TOL_CASE(9) SGMM_TOL_CASE(12) SGMM_TOL_CASE(16)
      default: SGMM_FAIL(c, "tolerance mode: unsupported feature dimension");
    }
#undef SGMM_TOL_CASE
    SGMM_LAUNCH_CHECK(c);
    sgmm_stage_end(c, SGMM_T_SCORE_REFINE);
    c->tol_active = true;
    return 0;
  }
  sgmm_stage_begin(c, SGMM_T_SCORE_REFINE);
  // refinement of every cell in one launch (alternating GEMM and refinement launches per range wastes the SMs at
  // every switch: a GEMM CTA needs the whole shared memory of an SM, so it waits for all refinement CTAs there)
  if (use_sm) {
    unsigned sgrid = (unsigned)c->n_sm_items;
    const uint4* d_items = c->d_sm_items.p;
    if (sm_dense) {
      const uint64_t nblk = (dense_nframes + kSmBlockFrames - 1) / kSmBlockFrames;
      std::vector<uint4> items, ea;
      std::vector<uint2> eb;
      for (uint64_t bk = 0; bk < nblk; bk++) {
        const uint64_t t0 = bk * kSmBlockFrames, T = std::min<uint64_t>(kSmBlockFrames, dense_nframes - t0);
        const uint64_t base = t0 * c->n_states;                       // first cell of the block's first row
        for (uint32_t st = 0; st < c->n_states; st++) {
          items.push_back(make_uint4(st, (uint32_t)ea.size(), 1u, (uint32_t)((T + 31) / 32)));
          ea.push_back(make_uint4((uint32_t)base, (uint32_t)(base >> 32), st << 16, 0u));
          eb.push_back(make_uint2((uint32_t)(dense_f0 + t0), (uint32_t)T));
        }
      }
      sgrid = (unsigned)items.size();
      SGMM_CHECK(c, c->d_dsm_items.reserve(items.size()));
      SGMM_CHECK(c, c->d_dsm_ent_a.reserve(ea.size()));
      SGMM_CHECK(c, c->d_dsm_ent_b.reserve(eb.size()));
      // (pageable sources: the copies are staged before the calls return, the vectors may go)
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_dsm_items.p, items.data(), items.size() * sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_dsm_ent_a.p, ea.data(), ea.size() * sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_dsm_ent_b.p, eb.data(), eb.size() * sizeof(uint2), cudaMemcpyHostToDevice, c->stream));
      d_items = c->d_dsm_items.p;
      r.sm_items = d_items; r.sm_ent_a = c->d_dsm_ent_a.p; r.sm_ent_b = c->d_dsm_ent_b.p;
      r.sm_dense_stride = c->n_states; r.sm_dense_f0 = (uint32_t)dense_f0;
    }
    SGMM_CHECK(c, c->d_sm_ivoff.reserve(sgrid));
    k_sm_ivoff<<<(sgrid + 255) / 256, 256, 0, c->stream>>>(d_items, sgrid, c->d_state_off.p, c->d_var_of.p, c->dimP, c->d_sm_ivoff.p);
    SGMM_LAUNCH_CHECK(c);
    r.sm_ivoff = c->d_sm_ivoff.p;
    bool civ = false;
    if (refine_civ_bind(c, &civ)) return 1;
#define SGMM_SM_CASE(D)                                                                                              \
  case D:                                                                                                            \
    if (sm_dense && civ) {                                                                                           \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_smem)); \
      k_tc_refine_sm<D, true, false, true><<<sgrid, 32 * sm_warps, sm_smem, c->stream>>>(r);                          \
    } else if (sm_dense) {                                                                                           \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_smem)); \
      k_tc_refine_sm<D, false, false, true><<<sgrid, 32 * sm_warps, sm_smem, c->stream>>>(r);                         \
    } else if (civ) {                                                                                                \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_smem)); \
      k_tc_refine_sm<D, true><<<sgrid, 32 * sm_warps, sm_smem, c->stream>>>(r);                                      \
    } else {                                                                                                         \
      SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine_sm<D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_smem)); \
      k_tc_refine_sm<D, false><<<sgrid, 32 * sm_warps, sm_smem, c->stream>>>(r);                                     \
    }                                                                                                                \
    break;
    switch (c->dim) {
      SGMM_SM_CASE(24) SGMM_SM_CASE(25) SGMM_SM_CASE(32) SGMM_SM_CASE(33) SGMM_SM_CASE(48) SGMM_SM_CASE(49)
      default: SGMM_FAIL(c, "state-major refinement: unsupported dimension");
    }
#undef SGMM_SM_CASE
    SGMM_LAUNCH_CHECK(c);
    if (civ && refine_civ_release(c)) return 1;
  } else {
    const unsigned rgrid = dense ? (unsigned)((dense_nframes + 127) / 128) : (unsigned)c->n_tasks128;
    r.tasks = dense ? nullptr : c->d_tasks128.p;
    switch (c->npair) {   // common feature dimensions get a fully unrolled row walk
      case 16:   // d = 32, 33
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsm_used));
        k_tc_refine<16><<<rgrid, 256, rsm_used, c->stream>>>(r); break;
      case 24:   // d = 48, 49
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsm_used));
        k_tc_refine<24><<<rgrid, 256, rsm_used, c->stream>>>(r); break;
      case 12:   // d = 24, 25
        SGMM_CHECK(c, cudaFuncSetAttribute(k_tc_refine<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsm_used));
        k_tc_refine<12><<<rgrid, 256, rsm_used, c->stream>>>(r); break;
      default: k_tc_refine<0><<<rgrid, 256, 0, c->stream>>>(r); break;
    }
    SGMM_LAUNCH_CHECK(c);
  }
  r.mu32T = c->d_mu32T.p; r.h32 = c->d_h32.p; r.c32 = c->d_c32.p; r.sguard = c->d_sguard.p;
  // option tc_rescan 1: the plain full-f64 rescan (same results)
  switch ((c->tc_rescan == 1 || c->tc_full) ? 0u : c->dv4) {
    case 7: k_tc_rescan2<7><<<148 * 16, 256, 0, c->stream>>>(r); break;
    case 9: k_tc_rescan2<9><<<148 * 16, 256, 0, c->stream>>>(r); break;
    case 12: k_tc_rescan2<12><<<148 * 16, 256, 0, c->stream>>>(r); break;
    case 16: k_tc_rescan2<16><<<148 * 16, 256, 0, c->stream>>>(r); break;
    default: k_tc_rescan<<<148 * 16, 256, 0, c->stream>>>(r); break;
  }
  SGMM_LAUNCH_CHECK(c);
  sgmm_stage_end(c, SGMM_T_SCORE_REFINE);
  return 0;
}

// Sum scores of every cell (dense: nframes x n_states, gathered: the ragged emission matrix) through k_sum_pruned.
int sgmm_launch_emission_sum_pruned(sgmm_ctx* c, bool dense, uint64_t dense_f0, uint64_t dense_nframes, double* d_scores,
                                    uint32_t* d_argmin, bool* handled) {
  using namespace sgmm_tc;
  *handled = false;
  if (!c->sum_prune || !(c->dv4 == 7 || c->dv4 == 9 || c->dv4 == 12 || c->dv4 == 16) || !c->d_mu32T.p) return 0;
  if (!dense && c->sum_tc && c->tc_ok && sgmm_effective_scorer(c) == 3) {
    // the tensor filter names the densities within 45 of each cell's best score, the state-major kernel sums them
    c->tc_sum_pass = true;
    const int rc = sgmm_launch_emission_tc(c, false, 0, 0, nullptr, nullptr);
    const bool done = c->tc_sum_pass;
    c->tc_sum_pass = false;
    if (rc) return rc;
    if (done) { *handled = true; return 0; }
  }
  RefineParams r{};
  r.m.meanT = c->d_meanT.p; r.m.invT = c->d_invT.p; r.m.mean_tail = c->d_mean_tail.p; r.m.inv_tail = c->d_inv_tail.p;
  r.m.norm = c->d_norm.p; r.m.logw = c->d_logw.p; r.m.dim = c->dim; r.m.npair = c->npair; r.m.Dtot = c->Dtot;
  r.m.meanR = c->d_meanR.p; r.m.invU = c->d_invU.p; r.m.var_of = c->d_var_of.p; r.m.dimP = c->dimP;
  r.state_off = c->d_state_off.p; r.x = c->d_x.p;
  r.n_cells = dense ? dense_nframes * c->n_states : c->E_elems;
  r.dense = dense ? 1 : 0; r.f0 = dense_f0; r.n_states = c->n_states; r.n_utts = c->n_utts;
  r.frame_off = c->d_frame_off.p; r.E_off = c->d_E_off.p; r.uniq = c->d_uniq.p; r.uniq_off = c->d_uniq_off.p;
  r.nuniq_pad = c->d_nuniq_pad.p;
  r.out = dense ? d_scores : c->d_E.p; r.argmin = dense ? d_argmin : nullptr;
  r.amin16 = dense ? nullptr : c->d_Amin.p;
  r.pooled = c->pooled_states ? 1 : 0;
  r.mu32T = c->d_mu32T.p; r.h32 = c->d_h32.p; r.c32 = c->d_c32.p; r.sguard = c->d_sguard.p;
  if (r.n_cells == 0) { *handled = true; return 0; }
  if (!dense)   // padding columns of the ragged matrix are never written by the kernel: keep them defined
    SGMM_CHECK(c, cudaMemsetAsync(c->d_E.p, 0, c->E_elems * sizeof(double), c->stream));
  r.tasks = dense ? nullptr : c->d_tasks.p;       // granules of 32 frames
  const unsigned grid = dense ? (unsigned)((dense_nframes + 31) / 32) : (unsigned)c->n_tasks;
  switch (c->dv4) {
    case 7: k_sum_pruned<7><<<grid, 256, 0, c->stream>>>(r); break;
    case 9: k_sum_pruned<9><<<grid, 256, 0, c->stream>>>(r); break;
    case 12: k_sum_pruned<12><<<grid, 256, 0, c->stream>>>(r); break;
    default: k_sum_pruned<16><<<grid, 256, 0, c->stream>>>(r); break;
  }
  SGMM_LAUNCH_CHECK(c);
  *handled = true;
  return 0;
}

// Sum score of every frame against its aligned state through k_sum_pruned (Trainer::calc_am_score with sum scores).
int sgmm_launch_sum_aligned_pruned(sgmm_ctx* c, bool* handled) {
  using namespace sgmm_tc;
  *handled = false;
  if (!c->sum_prune || !(c->dv4 == 7 || c->dv4 == 9 || c->dv4 == 12 || c->dv4 == 16) || !c->d_mu32T.p) return 0;
  *handled = true;
  if (c->n_frames == 0) return 0;
  RefineParams r{};
  r.m.meanT = c->d_meanT.p; r.m.invT = c->d_invT.p; r.m.mean_tail = c->d_mean_tail.p; r.m.inv_tail = c->d_inv_tail.p;
  r.m.norm = c->d_norm.p; r.m.logw = c->d_logw.p; r.m.dim = c->dim; r.m.npair = c->npair; r.m.Dtot = c->Dtot;
  r.m.meanR = c->d_meanR.p; r.m.invU = c->d_invU.p; r.m.var_of = c->d_var_of.p; r.m.dimP = c->dimP;
  r.state_off = c->d_state_off.p; r.x = c->d_x.p; r.ali = c->d_ali.p;
  r.n_cells = c->n_frames; r.dense = 2; r.n_states = c->n_states;
  r.out = c->d_fscore.p;
  r.pooled = c->pooled_states ? 1 : 0;
  r.mu32T = c->d_mu32T.p; r.h32 = c->d_h32.p; r.c32 = c->d_c32.p; r.sguard = c->d_sguard.p;
  SGMM_CHECK(c, cudaMemsetAsync(c->d_fdens.p, 0xff, c->n_frames * sizeof(uint32_t), c->stream));   // no per-frame density
  const unsigned grid = (unsigned)((c->n_frames + 31) / 32);
  switch (c->dv4) {
    case 7: k_sum_pruned<7><<<grid, 256, 0, c->stream>>>(r); break;
    case 9: k_sum_pruned<9><<<grid, 256, 0, c->stream>>>(r); break;
    case 12: k_sum_pruned<12><<<grid, 256, 0, c->stream>>>(r); break;
    default: k_sum_pruned<16><<<grid, 256, 0, c->stream>>>(r); break;
  }
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
