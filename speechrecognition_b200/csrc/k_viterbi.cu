// K2: batched Viterbi alignment over linear 0-1-2 HMMs with time-distortion penalties and a device
// backtrace — Aligner::align_sequence_full (reference src/sietill/Alignment.cpp:50-144), one warp per
// utterance.
//
//  * The trellis column lives in registers: lane l owns the C contiguous positions [l*C, (l+1)*C); only
//    three f64 values cross lanes per frame (warp shuffles).  Path costs are f64 and every add/compare is
//    the reference's, in its order (current = local + best; loop, then forward, then skip with strict
//    '<'), so costs and backpointers are bit-identical.  The reference's 0-1-2 band (:73-74,82,87) is an
//    optimisation only: cells outside it are +inf or cannot reach (T-1,S-1), so the un-banded DP below
//    visits a superset of cells with identical values on the band.
//  * Emission scores (f64, one row of the utterance's distinct states per frame) are streamed from HBM
//    by TMA bulk copies (cp.async.bulk -> shared, mbarrier completion) SGMM_STAGES tiles ahead of the DP.
//  * Backpointers are 2-bit jumps packed 16 frames per 32-bit word per position; they stay in shared
//    memory, so the backtrace never touches HBM.  The backtrace writes AlignmentItem{1, state, 1.0f}
//    (Alignment.cpp:131-133).
#include <math_constants.h>

#include "sgmm_internal.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
// 1-D TMA bulk copy global -> shared (SASS: UBLKCP); completion is signalled on the mbarrier.
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct VitParams {
  const uint32_t* utts;        // utterances of this launch class
  uint32_t n_utts;
  const uint64_t* frame_off;
  const uint32_t* aut;
  const uint32_t* aut_off;
  const uint16_t* pos2uniq;    // per automaton position: column in the utterance's emission row
  const uint32_t* nuniq_pad;
  const uint64_t* E_off;
  const double* E;             // emission matrix; f32 == 1: the same cells as floats (tolerance mode: the filter's scores)
  int f32;
  double tdp_loop, tdp_forward, tdp_skip;
  uint32_t silence_state;
  double beam;                 // DEST only: positions costlier than (frame best + beam) die; +inf = no pruning
  sgmm_alignment_item* ali;
  uint16_t* fcol;              // per frame: emission column of the aligned position
  double* costs;
};

template <int C>
__device__ __forceinline__ double pick(const double (&v)[C], int c) {
  double r = v[0];
#pragma unroll
  for (int i = 1; i < C; i++) r = (c == i) ? v[i] : r;
  return r;
}

// DEST = false: align_sequence_full semantics (source-state TDP, tie order loop > forward > skip, end at S-1).
// DEST = true : align_sequence_pruned (destination-state TDP, Alignment.cpp:191; candidates arrive skip,
//               forward, loop and replace on strictly smaller cost, :199-204; after every frame t >= 1 the
//               positions with cost > best + threshold die, :227-239 — a dead position is +inf here, which
//               never wins a strict comparison, exactly like a predecessor the reference skips; the
//               backtrace starts from the highest position alive in the last frame, :245-256).
template <int C, bool DEST, typename ET = double>   // ET: element type of the emission matrix (float: tolerance mode)
__global__ void __launch_bounds__(32) k_viterbi(VitParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x;
  if (blockIdx.x >= p.n_utts) return;
  const uint32_t u = p.utts[blockIdx.x];
  const uint64_t f0 = p.frame_off[u];
  const uint32_t T = (uint32_t)(p.frame_off[u + 1] - f0);
  const uint32_t a0 = p.aut_off[u];
  const uint32_t S = p.aut_off[u + 1] - a0;
  const uint32_t ncol = p.nuniq_pad[u];
  const ET* Eu = reinterpret_cast<const ET*>(p.E) + p.E_off[u];
  const double INF = CUDART_INF;

  // shared memory carve-up
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                          // SGMM_STAGES mbarriers
  ET* tiles = reinterpret_cast<ET*>(smem + 64);                        // STAGES x TILE_T x ncol
  const size_t tile_elems = (size_t)SGMM_TILE_T * ncol;
  uint32_t* bpw = reinterpret_cast<uint32_t*>(tiles + SGMM_STAGES * tile_elems);  // ceil(T/16) x 32C words
  uint16_t* sref = reinterpret_cast<uint16_t*>(bpw + (size_t)((T + 15) / 16) * 32 * C);  // automaton states
  uint16_t* scol = sref + S;                                                          // emission column per position

  const uint32_t n_tiles = (T + SGMM_TILE_T - 1) / SGMM_TILE_T;
  if (lane == 0) {
    for (int i = 0; i < SGMM_STAGES; i++) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (uint32_t s = lane; s < S; s += 32) { sref[s] = (uint16_t)p.aut[a0 + s]; scol[s] = p.pos2uniq[a0 + s]; }
  __syncwarp();
  if (lane == 0) {
    for (uint32_t i = 0; i < SGMM_STAGES && i < n_tiles; i++) {
      const uint32_t rows = min((uint32_t)SGMM_TILE_T, T - i * SGMM_TILE_T);
      const uint32_t bytes = SGMM_TILE_T * ncol * sizeof(ET);   // whole tiles: [column][16 frames] (sgmm_cell_rel)
      mbar_expect_tx(&bars[i], bytes);
      tma_load_1d(tiles + i * tile_elems, Eu + (size_t)i * tile_elems, bytes, &bars[i]);
    }
  }

  // per-position constants (registers).  TdpModel::score: the silence state pays tdp_forward for every jump.
  //   SOURCE convention: the penalty belongs to the position the transition LEAVES: loop from s, forward from
  //   s-1, skip from s-2.  DESTINATION: to the position it ENTERS (s for all three).
  double q[C], pl[C], pk[C];
  uint32_t col[C], w[C];
  uint32_t validmask = 0;
  const double INFd = INF;
#pragma unroll
  for (int c = 0; c < C; c++) {
    const uint32_t s = lane * C + c;
    q[c] = INFd;
    w[c] = 0;
    col[c] = 0;
    pl[c] = p.tdp_loop;
    pk[c] = p.tdp_skip;
    if (s < S) {
      validmask |= 1u << c;
      const uint32_t cj = p.pos2uniq[a0 + s];
      col[c] = (cj << 4) | (cj & 15u);   // cell of (frame tt, column cj) inside a tile: col[c] ^ tt  (sgmm_cell_rel)
      if (sref[s] == p.silence_state) pl[c] = p.tdp_forward;
      const uint32_t src = DEST ? s : (s >= 2 ? s - 2 : 0);
      if (sref[src] == p.silence_state) pk[c] = p.tdp_forward;
    }
  }
  const double pf = p.tdp_forward;
  const bool prune = DEST && p.beam < INF;
  // Lane 0 has no left neighbour.  Instead of masking the shuffled costs every frame, the penalties of the transitions
  // that would enter from it are +inf: cost + inf = +inf never wins a strict comparison (costs are never -inf), which
  // is what the reference's missing predecessor amounts to.
  const double pf0 = lane == 0 ? INFd : pf;          // forward into this lane's first position
  if (lane == 0) { pk[0] = INFd; if (C > 1) pk[C > 1 ? 1 : 0] = INFd; }

  // One frame of the recursion.  The backpointer is kept as the two raw comparison flags (bit 0 = first comparison,
  // bit 1 = second), set by predicated ORs with the frame's mask; the backtrace decodes them.
  auto frame = [&](const ET* __restrict__ tile, const uint32_t tt, const uint32_t m1) {
    double e[C];
#pragma unroll
    for (int c = 0; c < C; c++) e[c] = (double)tile[col[c] ^ tt];
    // the two costs that cross the lane boundary (positions s-1 and s-2 of this lane's first position)
    const double Q1 = __shfl_up_sync(0xffffffffu, q[C - 1], 1);
    const double Q2 = __shfl_up_sync(0xffffffffu, q[C - 2], 1);
    double nq[C];
#pragma unroll
    for (int c = 0; c < C; c++) {
      const double p1 = c > 0 ? q[c > 0 ? c - 1 : 0] : Q1;
      const double p2 = c > 1 ? q[c > 1 ? c - 2 : 0] : (c == 1 ? Q1 : Q2);
      const double pfc = c == 0 ? pf0 : pf;
      bool f1, f2;
      if (!DEST) {
        // current = local + min{loop, forward, skip}, ties loop > forward > skip (Alignment.cpp:96-113)
        const double lv = __dadd_rn(q[c], pl[c]);
        const double fv = __dadd_rn(p1, pfc);
        const double kv = __dadd_rn(p2, pk[c]);
        f1 = fv < lv;
        const double mn = f1 ? fv : lv;
        f2 = kv < mn;
        nq[c] = __dadd_rn(e[c], f2 ? kv : mn);      // jump = f2 ? 2 : f1
      } else {
        // (prev + tdp) + emission; candidates arrive skip, forward, loop and replace on strictly smaller cost
        // (Alignment.cpp:190-204).  Dead predecessors are +inf and never win a strict comparison.
        const double ck = __dadd_rn(__dadd_rn(p2, pk[c]), e[c]);
        const double cf = __dadd_rn(__dadd_rn(p1, pfc), e[c]);
        const double cl = __dadd_rn(__dadd_rn(q[c], pl[c]), e[c]);
        f1 = cf < ck;
        const double mn = f1 ? cf : ck;
        f2 = cl < mn;
        nq[c] = f2 ? cl : mn;                       // jump = f2 ? 0 : (f1 ? 1 : 2)
      }
      if (f1) w[c] |= m1;
      if (f2) w[c] |= m1 << 1;
    }
    if (prune) {
      // best over the automaton's own positions only (lanes past S hold junk from column 0)
      double best = INFd;
#pragma unroll
      for (int c = 0; c < C; c++)
        if ((validmask >> c) & 1) best = fmin(best, nq[c]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
      const double upper = __dadd_rn(best, p.beam);
#pragma unroll
      for (int c = 0; c < C; c++) nq[c] = nq[c] > upper ? INFd : nq[c];
    }
#pragma unroll
    for (int c = 0; c < C; c++) q[c] = nq[c];
  };

  // SGMM_TILE_T = 16 frames per emission tile = one 32-bit backpointer word per position and tile
  static_assert(SGMM_TILE_T == 16, "one backpointer word per tile");
  for (uint32_t ti = 0; ti < n_tiles; ti++) {
    const uint32_t slot = ti % SGMM_STAGES;
    mbar_wait(&bars[slot], (ti / SGMM_STAGES) & 1);
    const ET* tile = tiles + slot * tile_elems;
    const uint32_t rows = min((uint32_t)SGMM_TILE_T, T - ti * SGMM_TILE_T);
    if (ti > 0 && rows == SGMM_TILE_T) {
      // full tile: unrolled, the frame masks are immediates
#pragma unroll(C <= 4 ? 16 : (C <= 8 ? 4 : 2))
      for (int tt = 0; tt < SGMM_TILE_T; tt++) frame(tile, (uint32_t)tt, 1u << (2 * tt));
    } else {
      uint32_t tt = 0;
      if (ti == 0) {
        // Q(0,0) = emission(0, ref[0]); every other position starts at +inf (Alignment.cpp:62-63,77)
        const double e0 = (double)tile[col[0]];
#pragma unroll
        for (int c = 0; c < C; c++) q[c] = (lane == 0 && c == 0) ? e0 : INFd;
        tt = 1;
      }
#pragma unroll 1
      for (; tt < rows; tt++) frame(tile, tt, 1u << (2 * tt));
    }
#pragma unroll
    for (int c = 0; c < C; c++) { bpw[(size_t)ti * 32 * C + c * 32 + lane] = w[c]; w[c] = 0; }
    __syncwarp();
    const uint32_t nxt = ti + SGMM_STAGES;
    if (lane == 0 && nxt < n_tiles) {
      const uint32_t nrows = min((uint32_t)SGMM_TILE_T, T - nxt * SGMM_TILE_T);
      const uint32_t bytes = SGMM_TILE_T * ncol * sizeof(ET);
      mbar_expect_tx(&bars[slot], bytes);
      tma_load_1d(tiles + slot * tile_elems, Eu + (size_t)nxt * tile_elems, bytes, &bars[slot]);
    }
  }
  __syncwarp();

  // end position and path cost
  uint32_t end_s = S - 1;
  if (DEST) {
    uint32_t hi = 0;
#pragma unroll
    for (int c = 0; c < C; c++)
      if (((validmask >> c) & 1) && q[c] < INF) hi = lane * C + c + 1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    end_s = hi ? hi - 1 : 0;
  }
  const double mine = pick<C>(q, (int)(end_s % C));
  // a one-frame utterance never enters the reference's frame loop: align_sequence_full returns current_costs[S-1],
  // still +inf (Alignment.cpp:62-63,142); the pruned aligner returns the start node's score
  const double cost = (!DEST && T == 1) ? INF : __shfl_sync(0xffffffffu, mine, (int)(end_s / C));

  if (lane == 0) {
    if (p.costs) p.costs[u] = cost;
    uint2* out = reinterpret_cast<uint2*>(p.ali + f0);
    const uint32_t one = __float_as_uint(1.0f);
    uint32_t s = end_s;
    for (uint32_t t = T - 1; t >= 1; t--) {
      out[t] = make_uint2(1u | ((uint32_t)sref[s] << 16), one);
      p.fcol[f0 + t] = scol[s];
      const uint32_t word = bpw[(size_t)(t >> 4) * 32 * C + (s % C) * 32 + s / C];
      const uint32_t v = (word >> (2 * (t & 15))) & 3u;           // raw comparison flags of the forward pass
      const uint32_t j = DEST ? ((v & 2u) ? 0u : ((v & 1u) ? 1u : 2u)) : ((v & 2u) ? 2u : v);
      s = s >= j ? s - j : 0;
    }
    // frame 0: the full aligner writes reference[state reached] (= position 0 on any valid path); the
    // pruned one writes reference[0] explicitly (Alignment.cpp:272-275)
    out[0] = make_uint2(1u | ((uint32_t)sref[DEST ? 0 : s] << 16), one);
    p.fcol[f0] = scol[DEST ? 0 : s];
  }
}

// Multi-warp form for LONG utterances (T >= kVitMwMinFrames, sgmm_set_corpus): the automaton's positions are split over
// W = 4 warps — thread g of the CTA owns positions [g C, (g + 1) C) — so that a frame costs C <= 4 positions per lane on
// FOUR sub-partitions' FP64 pipes instead of up to 16 positions per lane on one (the one-warp kernel walked the 1399-frame
// utterances at ~340 cycles per frame: they were the critical path of the whole stage).  Inside a warp the two boundary
// costs cross lanes by shuffle as before; across warps lane 31 leaves {q[C-1], q[C-2]} in a double-buffered shared-memory
// slot and ONE CTA barrier per frame makes them visible.  Same additions and comparisons in the same order as k_viterbi:
// same bits.  The beam of the destination convention needs the frame's minimum over all warps: every warp leaves its
// minimum next to the boundary costs, and each thread prunes its own costs AND the neighbour's boundary costs it reads in
// the next frame with the same bound — still one barrier per frame.
constexpr int kVitW = SGMM_VIT_MW_WARPS;
template <int C, bool DEST>
__global__ void __launch_bounds__(32 * kVitW) k_viterbi_mw(VitParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int NT = 32 * kVitW;
  const int g = threadIdx.x, lane = g & 31, warp = g >> 5;
  if (blockIdx.x >= p.n_utts) return;
  const uint32_t u = p.utts[blockIdx.x];
  const uint64_t f0 = p.frame_off[u];
  const uint32_t T = (uint32_t)(p.frame_off[u + 1] - f0);
  const uint32_t a0 = p.aut_off[u];
  const uint32_t S = p.aut_off[u + 1] - a0;
  const uint32_t ncol = p.nuniq_pad[u];
  const double* Eu = p.E + p.E_off[u];
  const double INF = CUDART_INF;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                          // SGMM_STAGES mbarriers
  double* xq = reinterpret_cast<double*>(smem + 64);                           // [2 frames][W][q[C-1], q[C-2], warp min, -]
  uint32_t* sh_hi = reinterpret_cast<uint32_t*>(xq + 2 * kVitW * 4);           // highest live position (DEST), + pad
  double* tiles = reinterpret_cast<double*>(sh_hi + 4);                        // STAGES x TILE_T x ncol
  const size_t tile_elems = (size_t)SGMM_TILE_T * ncol;
  uint32_t* bpw = reinterpret_cast<uint32_t*>(tiles + SGMM_STAGES * tile_elems);   // ceil(T/16) x NT x C words
  uint16_t* sref = reinterpret_cast<uint16_t*>(bpw + (size_t)((T + 15) / 16) * NT * C);
  uint16_t* scol = sref + S;

  const uint32_t n_tiles = (T + SGMM_TILE_T - 1) / SGMM_TILE_T;
  if (g == 0) {
    for (int i = 0; i < SGMM_STAGES; i++) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    *sh_hi = 0;
  }
  for (uint32_t s = g; s < S; s += NT) { sref[s] = (uint16_t)p.aut[a0 + s]; scol[s] = p.pos2uniq[a0 + s]; }
  __syncthreads();
  if (g == 0) {
    for (uint32_t i = 0; i < SGMM_STAGES && i < n_tiles; i++) {
      const uint32_t rows = min((uint32_t)SGMM_TILE_T, T - i * SGMM_TILE_T);
      const uint32_t bytes = SGMM_TILE_T * ncol * sizeof(double);   // whole tiles: [column][16 frames] (sgmm_cell_rel)
      mbar_expect_tx(&bars[i], bytes);
      tma_load_1d(tiles + i * tile_elems, Eu + (size_t)i * tile_elems, bytes, &bars[i]);
    }
  }

  double q[C], pl[C], pk[C];
  uint32_t col[C], w[C];
  uint32_t validmask = 0;
#pragma unroll
  for (int c = 0; c < C; c++) {
    const uint32_t s = g * C + c;
    q[c] = INF; w[c] = 0; col[c] = 0;
    pl[c] = p.tdp_loop; pk[c] = p.tdp_skip;
    if (s < S) {
      validmask |= 1u << c;
      col[c] = ((uint32_t)scol[s] << 4) | (scol[s] & 15u);   // cell inside a tile: col[c] ^ tt  (sgmm_cell_rel)
      if (sref[s] == p.silence_state) pl[c] = p.tdp_forward;
      const uint32_t src = DEST ? s : (s >= 2 ? s - 2 : 0);
      if (sref[src] == p.silence_state) pk[c] = p.tdp_forward;
    }
  }
  const double pf = p.tdp_forward;
  const bool prune = DEST && p.beam < INF;
  double upper_prev = INF;        // pruning bound of the previous frame (applied to the neighbour's boundary costs)

  for (uint32_t ti = 0; ti < n_tiles; ti++) {
    const uint32_t slot = ti % SGMM_STAGES;
    mbar_wait(&bars[slot], (ti / SGMM_STAGES) & 1);
    const double* tile = tiles + slot * tile_elems;
    const uint32_t t0 = ti * SGMM_TILE_T;
    const uint32_t rows = min((uint32_t)SGMM_TILE_T, T - t0);
    for (uint32_t tt = 0; tt < rows; tt++) {
      const uint32_t t = t0 + tt;
      double e[C];
#pragma unroll
      for (int c = 0; c < C; c++) e[c] = tile[col[c] ^ tt];
      double* xw = xq + ((t & 1) * kVitW + warp) * 4;
      if (t == 0) {
#pragma unroll
        for (int c = 0; c < C; c++) q[c] = (g == 0 && c == 0) ? e[0] : INF;
      } else {
        double Q1 = __shfl_up_sync(0xffffffffu, q[C - 1], 1);
        double Q2 = __shfl_up_sync(0xffffffffu, q[C - 2], 1);
        if (lane == 0) {
          if (warp == 0) { Q1 = INF; Q2 = INF; }
          else {
            const double* xr = xq + (((t - 1) & 1) * kVitW + warp - 1) * 4;
            Q1 = xr[0]; Q2 = xr[1];
            if (prune) { Q1 = Q1 > upper_prev ? INF : Q1; Q2 = Q2 > upper_prev ? INF : Q2; }
          }
        }
        double nq[C];
        const uint32_t sh = 2 * (t & 15);
#pragma unroll
        for (int c = 0; c < C; c++) {
          const double p1 = c > 0 ? q[c > 0 ? c - 1 : 0] : Q1;
          const double p2 = c > 1 ? q[c > 1 ? c - 2 : 0] : (c == 1 ? Q1 : Q2);
          uint32_t j;
          if (!DEST) {
            const double lv = __dadd_rn(q[c], pl[c]);
            const double fv = __dadd_rn(p1, pf);
            const double kv = __dadd_rn(p2, pk[c]);
            const bool f1 = fv < lv;
            const double m1 = f1 ? fv : lv;
            const bool f2 = kv < m1;
            const double best = f2 ? kv : m1;
            j = f2 ? 2u : (f1 ? 1u : 0u);
            nq[c] = __dadd_rn(e[c], best);
          } else {
            const double ck = __dadd_rn(__dadd_rn(p2, pk[c]), e[c]);
            const double cf = __dadd_rn(__dadd_rn(p1, pf), e[c]);
            const double cl = __dadd_rn(__dadd_rn(q[c], pl[c]), e[c]);
            const bool f1 = cf < ck;
            const double m1 = f1 ? cf : ck;
            const bool f2 = cl < m1;
            nq[c] = f2 ? cl : m1;
            j = f2 ? 0u : (f1 ? 1u : 2u);
          }
          w[c] |= j << sh;
        }
        if (prune) {
          double best = INF;
#pragma unroll
          for (int c = 0; c < C; c++)
            if ((validmask >> c) & 1) best = fmin(best, nq[c]);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
          if (lane == 0) xw[2] = best;
        }
#pragma unroll
        for (int c = 0; c < C; c++) q[c] = nq[c];
      }
      if (lane == 31) { xw[0] = q[C - 1]; xw[1] = q[C - 2]; }      // (unpruned; the reader applies the frame's bound)
      __syncthreads();                                             // the frame's one barrier
      if (prune && t >= 1) {
        const double* xb = xq + (t & 1) * kVitW * 4;
        double best = xb[2];
#pragma unroll
        for (int ww = 1; ww < kVitW; ww++) best = fmin(best, xb[ww * 4 + 2]);
        const double upper = __dadd_rn(best, p.beam);
#pragma unroll
        for (int c = 0; c < C; c++) q[c] = q[c] > upper ? INF : q[c];
        upper_prev = upper;
      }
      if (t >= 1 && ((t & 15) == 15 || t == T - 1)) {
#pragma unroll
        for (int c = 0; c < C; c++) { bpw[(size_t)(t >> 4) * NT * C + c * NT + g] = w[c]; w[c] = 0; }
      }
    }
    // every thread has passed the barrier of the tile's last frame: the slot can be refilled
    const uint32_t nxt = ti + SGMM_STAGES;
    if (g == 0 && nxt < n_tiles) {
      const uint32_t nrows = min((uint32_t)SGMM_TILE_T, T - nxt * SGMM_TILE_T);
      const uint32_t bytes = SGMM_TILE_T * ncol * sizeof(double);
      mbar_expect_tx(&bars[slot], bytes);
      tma_load_1d(tiles + slot * tile_elems, Eu + (size_t)nxt * tile_elems, bytes, &bars[slot]);
    }
  }

  // end position and path cost
  uint32_t end_s = S - 1;
  if (DEST) {
    uint32_t hi = 0;
#pragma unroll
    for (int c = 0; c < C; c++)
      if (((validmask >> c) & 1) && q[c] < INF) hi = g * C + c + 1;
    if (hi) atomicMax(sh_hi, hi);
    __syncthreads();
    hi = *sh_hi;
    end_s = hi ? hi - 1 : 0;
  }
  if ((uint32_t)g == end_s / C && p.costs) p.costs[u] = (!DEST && T == 1) ? INF : pick<C>(q, (int)(end_s % C));
  __syncthreads();   // the last backpointer words are in shared memory

  if (g == 0) {
    uint2* out = reinterpret_cast<uint2*>(p.ali + f0);
    const uint32_t one = __float_as_uint(1.0f);
    uint32_t s = end_s;
    for (uint32_t t = T - 1; t >= 1; t--) {
      out[t] = make_uint2(1u | ((uint32_t)sref[s] << 16), one);
      p.fcol[f0 + t] = scol[s];
      const uint32_t word = bpw[(size_t)(t >> 4) * NT * C + (s % C) * NT + s / C];
      const uint32_t j = (word >> (2 * (t & 15))) & 3u;
      s = s >= j ? s - j : 0;
    }
    out[0] = make_uint2(1u | ((uint32_t)sref[DEST ? 0 : s] << 16), one);
    p.fcol[f0] = scol[DEST ? 0 : s];
  }
}

template <int C>
int launch_mw(sgmm_ctx* c, const VitParams& p, size_t smem, int conv, cudaStream_t stream) {
  if (conv == SGMM_TDP_SOURCE) {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_viterbi_mw<C, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_viterbi_mw<C, false><<<p.n_utts, 32 * kVitW, smem, stream>>>(p);
  } else {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_viterbi_mw<C, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_viterbi_mw<C, true><<<p.n_utts, 32 * kVitW, smem, stream>>>(p);
  }
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

template <int C>
int launch_c(sgmm_ctx* c, const VitParams& p, size_t smem, int conv, cudaStream_t stream) {
  if (p.f32) {   // tolerance mode (experiment): the full aligner on the filter's f32 scores
    if (conv != SGMM_TDP_SOURCE) SGMM_FAIL(c, "viterbi: the tolerance mode supports the full aligner only");
    SGMM_CHECK(c, cudaFuncSetAttribute(k_viterbi<C, false, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_viterbi<C, false, float><<<p.n_utts, 32, smem, stream>>>(p);
  } else if (conv == SGMM_TDP_SOURCE) {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_viterbi<C, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_viterbi<C, false><<<p.n_utts, 32, smem, stream>>>(p);
  } else {
    SGMM_CHECK(c, cudaFuncSetAttribute(k_viterbi<C, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_viterbi<C, true><<<p.n_utts, 32, smem, stream>>>(p);
  }
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

}  // namespace

int sgmm_launch_viterbi(sgmm_ctx* c, int conv) {
  // the shape classes are independent: fork them over the side streams, join on the main stream
  SGMM_CHECK(c, cudaEventRecord(c->ev_fork, c->stream));
  int n_launched = 0;
  for (ViterbiClass& vc : c->vclasses) {
    if (vc.utts.empty()) continue;
    cudaStream_t stream = c->aux[n_launched % sgmm_ctx::kAux];
    if (n_launched < sgmm_ctx::kAux) SGMM_CHECK(c, cudaStreamWaitEvent(stream, c->ev_fork, 0));
    n_launched++;
    VitParams p{};
    p.utts = vc.d_utts.p; p.n_utts = (uint32_t)vc.utts.size();
    p.frame_off = c->d_frame_off.p; p.aut = c->d_aut.p; p.aut_off = c->d_aut_off.p;
    p.pos2uniq = c->d_pos2uniq.p; p.nuniq_pad = c->d_nuniq_pad.p; p.E_off = c->d_E_off.p; p.E = c->d_E.p;
    if (c->tol_active) { p.E = reinterpret_cast<const double*>(c->d_e32.p); p.f32 = 1; }
    p.tdp_loop = c->tdp_loop; p.tdp_forward = c->tdp_forward; p.tdp_skip = c->tdp_skip;
    p.silence_state = c->silence_state;
    p.beam = c->beam;
    p.ali = c->d_ali.p; p.fcol = c->d_fcol.p; p.costs = c->d_costs.p;
    int rc = 1;
    if (vc.W == kVitW) {
      if (p.f32) SGMM_FAIL(c, "viterbi: the tolerance mode needs option vit_mw = 0 before sgmm_set_corpus");
      switch (vc.C) {
        case 2: rc = launch_mw<2>(c, p, vc.smem, conv, stream); break;
        case 3: rc = launch_mw<3>(c, p, vc.smem, conv, stream); break;
        case 4: rc = launch_mw<4>(c, p, vc.smem, conv, stream); break;
        default: SGMM_FAIL(c, "viterbi: unsupported positions-per-lane class of the multi-warp kernel");
      }
      if (rc) return rc;
      continue;
    }
    switch (vc.C) {
      case 2: rc = launch_c<2>(c, p, vc.smem, conv, stream); break;
      case 3: rc = launch_c<3>(c, p, vc.smem, conv, stream); break;
      case 4: rc = launch_c<4>(c, p, vc.smem, conv, stream); break;
      case 5: rc = launch_c<5>(c, p, vc.smem, conv, stream); break;
      case 6: rc = launch_c<6>(c, p, vc.smem, conv, stream); break;
      case 8: rc = launch_c<8>(c, p, vc.smem, conv, stream); break;
      case 10: rc = launch_c<10>(c, p, vc.smem, conv, stream); break;
      case 12: rc = launch_c<12>(c, p, vc.smem, conv, stream); break;
      case 16: rc = launch_c<16>(c, p, vc.smem, conv, stream); break;
      default: SGMM_FAIL(c, "viterbi: unsupported positions-per-lane class");
    }
    if (rc) return rc;
  }
  for (int i = 0; i < n_launched && i < sgmm_ctx::kAux; i++) {
    SGMM_CHECK(c, cudaEventRecord(c->ev_join[i], c->aux[i]));
    SGMM_CHECK(c, cudaStreamWaitEvent(c->stream, c->ev_join[i], 0));
  }
  return 0;
}
