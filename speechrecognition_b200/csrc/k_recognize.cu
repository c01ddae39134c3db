// K6: batched word-loop recognition — Recognizer::recognizeSequence_pruned (Recognizer.cpp:103-232) and the
// per-segment loop of Recognizer::recognize (:46-47), one warp per utterance.
//
// The reference pushes hypotheses from a table indexed max_len * word + position, visiting sources in index
// order and replacing a target on strict improvement.  Here every TARGET slot gathers its candidates in that
// same source order, so the survivor (score, word-start time) is the one the reference keeps, including its
// pre-emission rejection test (:142-145, :172-175) that matters when emission scores are negative:
//   - position 0 / 1 of word w is entered from the LAST position of every word v (a hypothesis there can only
//     leave, :131) with word_penalty(w) + tdp(first_state(w), position + 1) + emission(first_state(w)) — the
//     word's FIRST state also for position 1 (:136-151);
//   - inside a word: jumps 0/1/2 from non-final positions with the DESTINATION state's tdp (:162-187);
//   - per frame: prune against best + am_threshold (:193-198), best word end with strict '>' (:201-207);
//   - backtrace over word ends, silence not emitted (:222-231).
// Emission scores come from the dense K1 pass ([frames][states] f64, exactly min_score().first).
// Two kernels: k_recognize (any lexicon up to 4096 table slots: hypothesis tables in shared memory) and k_recognize_reg
// (<= 256 slots, <= 32 words: the table in registers, see there); both launch the longest utterances first, and
// k_rec_compact packs the word lists so that the host reads words, not a frame-sized array.
#include <algorithm>
#include <type_traits>
#include <vector>

#include "sgmm_internal.cuh"

namespace {

constexpr int kRecWarps = 4;       // utterances per CTA
constexpr int kMaxSlotsPerLane = 16;
#ifndef SGMM_REC_LEAN
#define SGMM_REC_LEAN 1
#endif
constexpr bool kLeanBody = SGMM_REC_LEAN != 0;
#ifndef SGMM_REC_MINB
#define SGMM_REC_MINB 3
#endif

struct RecParams {
  const double* E;             // [chunk frames][n_states] emission scores, row 0 = frame chunk_f0
  uint64_t chunk_f0;
  uint32_t n_states;
  const uint64_t* frame_off;   // [n_utts + 1] over the resident features
  const uint32_t* perm;        // [n_utts(all)] launch order: within a launch the longest utterance first (it ends last)
  uint32_t u0, n_utts;         // utterances [u0, u0 + n_utts) in this launch
  // lexicon tables (device)
  uint32_t n_words, n_live, silence_word;
  const uint4* slot_tab;       // [n_live] in processing order (the word-entry slots, positions 0 and 1, first):
                               //   x = word | pos << 8 | len << 16 | (pos < len) << 31,  y = slot | word's first slot << 16,
                               //   z = automaton state of the slot,  w = first state of the word
  const uint32_t* word_end;    // [n_words] slot of the word's last position
  int stage_rows;              // emission rows staged in shared memory one frame ahead (cp.async)
  double loop, forward, skip, word_penalty, am_threshold;
  uint32_t silence_state;
  // outputs
  uint32_t* tb;                // [frames + n_utts] per utterance T+1 entries: word | bkp << 16
  const uint64_t* tb_off;      // [n_utts(all)] offset of the utterance's traceback
  uint32_t* words_raw;         // [frames] recognised words of utterance u at frame_off[u] ...
  uint32_t* n_words_out;       // [n_utts(all)]
  double* score_out;           // [n_utts(all)]
};

__device__ __forceinline__ double tdp_of(const RecParams& p, uint32_t state, uint32_t jump) {
  if (state == p.silence_state) return p.forward;            // TdpModel.cpp:19-29
  return jump == 0 ? p.loop : (jump == 1 ? p.forward : p.skip);
}

// doubles of shared memory one warp needs: 2 score buffers, 2 u16 back-pointer buffers, the word-end scores, the rows
__host__ __device__ inline size_t rec_warp_doubles(size_t nl, size_t W, size_t row_d) {
  return 2 * nl + ((2 * nl + 3) / 4) + W + row_d;
}

__global__ void __launch_bounds__(kRecWarps * 32) k_recognize(RecParams p) {
  extern __shared__ unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t nl = p.n_live, W = p.n_words;
  uint4* s_tab = reinterpret_cast<uint4*>(smem_raw);
  uint32_t* s_end = reinterpret_cast<uint32_t*>(s_tab + nl);
  double* warp_mem = reinterpret_cast<double*>(smem_raw + (((size_t)nl * 16 + (size_t)W * 4 + 15) & ~(size_t)15));
  const size_t row_d = p.stage_rows ? (size_t)2 * p.n_states : 0;   // two staged emission rows
  double* sc0 = warp_mem + warp * rec_warp_doubles(nl, W, row_d);
  double* sc1 = sc0 + nl;
  uint16_t* bk0 = reinterpret_cast<uint16_t*>(sc1 + nl);
  uint16_t* bk1 = bk0 + nl;
  double* we = sc0 + (size_t)2 * nl + ((2 * nl + 3) / 4);   // scores of the word ends after the previous frame
  double* srow = we + W;
  for (uint32_t i = threadIdx.x; i < nl; i += blockDim.x) s_tab[i] = p.slot_tab[i];
  for (uint32_t i = threadIdx.x; i < W; i += blockDim.x) s_end[i] = p.word_end[i];
  __syncthreads();
  const uint32_t ul = blockIdx.x * kRecWarps + warp;
  if (ul >= p.n_utts) return;
  const uint32_t u = p.perm[p.u0 + ul];
  const uint64_t f0 = p.frame_off[u];
  const uint32_t T = (uint32_t)(p.frame_off[u + 1] - f0);
  const double* E = p.E + (f0 - p.chunk_f0) * p.n_states;
  uint32_t* tb = p.tb + p.tb_off[u];
  const double inf = __longlong_as_double(0x7ff0000000000000ll);

  for (uint32_t i = lane; i < nl; i += 32) { sc0[i] = inf; sc1[i] = inf; bk0[i] = 0; bk1[i] = 0; }
  __syncwarp();
  if (lane == 0) { sc0[0] = 0.0; tb[0] = 0; }     // start hypothesis: word 0, position 0, score 0 (:121)
  __syncwarp();
  for (uint32_t v = lane; v < W; v += 32) we[v] = sc0[s_end[v]];

  double* prev = sc0; double* next = sc1;
  uint16_t* pbk = bk0; uint16_t* nbk = bk1;
  // The reference prunes the table after every frame (:193-198).  Here the scores stay as computed and every reader of
  // the previous frame applies that frame's limit (same values, one pass over the table less per frame).
  double prev_limit = inf;
  // emission row of frame t - 1 -> srow[(t - 1) & 1], copied one frame ahead (8-byte cp.async, any n_states)
  auto stage_row = [&](uint32_t fr) {
    const double* src = E + (size_t)fr * p.n_states;
    double* dst = srow + (size_t)(fr & 1u) * p.n_states;
    for (uint32_t k = lane; k < p.n_states; k += 32)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + k)), "l"(src + k) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (p.stage_rows && T > 0) stage_row(0);
  for (uint32_t t = 1; t <= T; t++) {
    const double* am = E + (size_t)(t - 1) * p.n_states;
    if (p.stage_rows) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      if (t < T) stage_row(t);          // the other buffer: its last readers finished before the previous frame's end
      am = srow + (size_t)((t - 1) & 1u) * p.n_states;
    }
    double my_best = inf;
    for (uint32_t oi = lane; oi < nl; oi += 32) {
      const uint4 sl = s_tab[oi];
      const uint32_t info = sl.x, i = sl.y & 0xffffu, base = sl.y >> 16;
      const uint32_t w = info & 0xffu, pos = (info >> 8) & 0xffu, L = (info >> 16) & 0xffu;
      const bool in_word = (info >> 31) != 0;          // pos < L (a real position of the word)
      double cur = inf;
      uint16_t bk = 0;
      double am_e = 0.0, pre_e = 0.0;
      const bool entry = pos <= 1;
      if (entry) {
        am_e = am[sl.w];
        pre_e = tdp_of(p, sl.w, pos + 1);
      }
      const double wp = (w != p.silence_word) ? p.word_penalty : 0.0;
      // one word-end candidate (:136-151): rejected before the emission is added when it is already worse
      auto enter = [&](double S) {
        if (S == inf) return;
        double ns = S + wp + pre_e;
        if (ns > cur) return;
        ns += am_e;
        if (cur > ns) { cur = ns; bk = (uint16_t)(t - 1); }
      };
      // With a non-negative emission the early rejection never changes the outcome and every word-end candidate
      // carries the same back-pointer, so a run of them reduces to its lowest score (fp addition is monotone).
      const bool collapse = am_e >= 0.0;
      double run_a = inf, run_b = inf;
      if (entry && collapse)
        for (uint32_t v = 0; v < W; v++) {
          const double S = we[v];
          if (v < w) run_a = fmin(run_a, S); else run_b = fmin(run_b, S);
        }
      // sources in table order: word ends of the words before w ...
      if (entry) {
        if (collapse) enter(run_a);
        else for (uint32_t v = 0; v < w; v++) enter(we[v]);
      }
      // ... then the word's own positions, ascending: pos-2, pos-1, pos (never its last position, which only exits) ...
      if (in_word) {
        const uint32_t st = sl.z;
        const double am_s = am[st];
#pragma unroll
        for (int jump = 2; jump >= 0; jump--) {
          if ((int)pos - jump < 0) continue;
          const uint32_t q = pos - (uint32_t)jump;
          if (q == L - 1) continue;
          const double S = prev[base + q];
          if (S == inf || S > prev_limit) continue;
          double ns = S + tdp_of(p, st, (uint32_t)jump);
          if (ns > cur) continue;
          ns += am_s;
          if (cur > ns) { cur = ns; bk = pbk[base + q]; }
        }
      }
      // ... then the word ends of w itself and of the words after it
      if (entry) {
        if (collapse) enter(run_b);
        else for (uint32_t v = w; v < W; v++) enter(we[v]);
      }
      next[i] = cur;
      nbk[i] = bk;
      my_best = fmin(my_best, cur);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_best = fmin(my_best, __shfl_xor_sync(0xffffffffu, my_best, o));
    const double limit = my_best + p.am_threshold;
    __syncwarp();
    // best word end of this frame: lowest score, lowest word on ties (strict '>' in table order)
    double e_best = inf;
    uint32_t e_word = 0xffffffffu;
    for (uint32_t v = lane; v < W; v += 32) {
      double S = next[s_end[v]];
      if (S > limit) S = inf;
      we[v] = S;
      if (S < e_best) { e_best = S; e_word = v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double os = __shfl_xor_sync(0xffffffffu, e_best, o);
      const uint32_t ow = __shfl_xor_sync(0xffffffffu, e_word, o);
      if (os < e_best || (os == e_best && ow < e_word)) { e_best = os; e_word = ow; }
    }
    if (lane == 0) {
      uint32_t rec = 0;                                    // no finite word end: word 0, bkp 0 (:117-119)
      if (e_best < inf) rec = e_word | ((uint32_t)nbk[s_end[e_word]] << 16);
      tb[t] = rec;
      if (t == T && p.score_out) p.score_out[u] = e_best;
    }
    prev_limit = limit;
    double* ts = prev; prev = next; next = ts;
    uint16_t* tk = pbk; pbk = nbk; nbk = tk;
    __syncwarp();
  }
  if (lane == 0) {
    if (T == 0 && p.score_out) p.score_out[u] = 0.0;
    uint32_t n = 0;
    for (uint32_t t = T; t > 0; t = tb[t] >> 16)
      if ((tb[t] & 0xffffu) != p.silence_word) n++;
    uint32_t* out = p.words_raw + f0;
    uint32_t k = n;
    for (uint32_t t = T; t > 0; t = tb[t] >> 16)
      if ((tb[t] & 0xffffu) != p.silence_word) out[--k] = tb[t] & 0xffffu;
    p.n_words_out[u] = n;
  }
}

// ---- small lexicons (<= 256 slots, <= 32 words: the sietill digits have 212 and 12): hypotheses in registers ----
// Lane l owns the K consecutive slots l * K .. l * K + K - 1 of the table, so a slot's in-word sources (positions
// pos, pos - 1, pos - 2) are the lane's own registers or the last two of the lane to its left (shuffles); nothing of the
// table goes through shared memory and the frame body is predicated instead of branched.  A long utterance is one warp
// running alone for most of the launch (utterance lengths 279 .. 1399), so the frame LATENCY is what counts: the
// general kernel above takes 7.3 k cycles per frame there.
// Word entries use the exclusive prefix / inclusive suffix minima of the word-end scores (a warp scan per frame) when
// the entered state's emission is non-negative (see `collapse` above); otherwise that slot walks the word ends in
// table order exactly as the reference does.
struct RegParams {
  RecParams r;
  const uint4* slot_nat;       // [n_live] table order: x = flags (below), y = state, z = first state of the word, w = 0
};
enum : uint32_t {
  kFEntry = 1u, kFJ0 = 2u, kFJ1 = 4u, kFJ2 = 8u, kFSil = 16u, kFFirstSil = 32u, kFPos1 = 64u, kFEnd = 128u,
  kFPenalty = 1u << 16   // bits 8..15: word
};

template <int K>
__global__ void __launch_bounds__(kRecWarps * 32, SGMM_REC_MINB) k_recognize_reg(RegParams q) {
  const RecParams& p = q.r;
  extern __shared__ unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t nl = p.n_live, W = p.n_words, N = p.n_states;
  // per warp: we[32], pa[32], sb[32] doubles, webk[32] u32 (16 doubles), two emission rows
  double* wm = reinterpret_cast<double*>(smem_raw) + (size_t)warp * (112 + 2 * (size_t)N);
  double* we = wm; double* pa_s = wm + 32; double* sb_s = wm + 64;
  uint32_t* webk = reinterpret_cast<uint32_t*>(wm + 96);
  double* srow = wm + 112;
  const uint32_t ul = blockIdx.x * kRecWarps + warp;
  if (ul >= p.n_utts) return;
  const uint32_t u = p.perm[p.u0 + ul];
  const uint64_t f0 = p.frame_off[u];
  const uint32_t T = (uint32_t)(p.frame_off[u + 1] - f0);
  const double* E = p.E + (f0 - p.chunk_f0) * N;
  uint32_t* tb = p.tb + p.tb_off[u];
  const double inf = __longlong_as_double(0x7ff0000000000000ll);

  uint32_t fl[K], st[K], fst[K];
  double sc[K];
  uint32_t bk[K];
#pragma unroll
  for (int k = 0; k < K; k++) {
    const uint32_t s = (uint32_t)lane * K + k;
    uint4 m = make_uint4(0, 0, 0, 0);
    if (s < nl) m = q.slot_nat[s];
    fl[k] = m.x | (m.y == p.silence_state ? kFSil : 0u) | (m.z == p.silence_state ? kFFirstSil : 0u);   // TdpModel.cpp:19-29
    st[k] = m.y; fst[k] = m.z;
    sc[k] = inf; bk[k] = 0;
  }
  if (lane == 0) { sc[0] = 0.0; tb[0] = 0; }     // start hypothesis: word 0, position 0, score 0 (:121)

  // End of a frame.  `sc` holds the frame's scores before the beam; the beam's best score (a butterfly over the lanes)
  // and the prefix / suffix minima of the word-end scores (two scans) are independent chains of the same five shuffle
  // steps, so they run interleaved, and the beam is applied to the minima afterwards: pruning is monotone, the minimum
  // of the pruned scores is the pruned minimum.  `we` keeps the unpruned word-end scores (readers apply `limit`).
  double limit = inf, e_best;
  uint32_t e_word;
  auto finish_frame = [&](double lane_best, bool beam) {
    __syncwarp();
#pragma unroll
    for (int k = 0; k < K; k++)
      if (fl[k] & kFEnd) { we[(fl[k] >> 8) & 0xffu] = sc[k]; webk[(fl[k] >> 8) & 0xffu] = bk[k]; }
    __syncwarp();
    const double x = (uint32_t)lane < W ? we[lane] : inf;
    double ip = x, is = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double a = __shfl_up_sync(0xffffffffu, ip, o), b = __shfl_down_sync(0xffffffffu, is, o);
      lane_best = fmin(lane_best, __shfl_xor_sync(0xffffffffu, lane_best, o));
      if (lane >= o) ip = fmin(ip, a);
      if (lane + o < 32) is = fmin(is, b);
    }
    limit = beam ? lane_best + p.am_threshold : inf;       // beam (:193-198)
    double ex = __shfl_up_sync(0xffffffffu, ip, 1);
    if (lane == 0) ex = inf;
    pa_s[lane] = ex > limit ? inf : ex;
    sb_s[lane] = is > limit ? inf : is;
    const double all = __shfl_sync(0xffffffffu, is, 0);
    const uint32_t hit = __ballot_sync(0xffffffffu, (uint32_t)lane < W && x == all);
    e_best = all > limit ? inf : all;
    e_word = hit ? (uint32_t)__ffs(hit) - 1u : 0xffffffffu;   // lowest word on ties (strict '>' in table order)
#pragma unroll
    for (int k = 0; k < K; k++) sc[k] = sc[k] > limit ? inf : sc[k];
    __syncwarp();
  };
  finish_frame(inf, false);

  auto stage_row = [&](uint32_t fr) {
    const double* src = E + (size_t)fr * N;
    double* dst = srow + (size_t)(fr & 1u) * N;
    for (uint32_t k = lane; k < N; k += 32)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + k)), "l"(src + k) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (T > 0) stage_row(0);
  for (uint32_t t = 1; t <= T; t++) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    if (t < T) stage_row(t);
    const double* am = srow + (size_t)((t - 1) & 1u) * N;
    // the two slots to the left of this lane's first one
    const double l1 = __shfl_up_sync(0xffffffffu, sc[K - 1], 1), l2 = __shfl_up_sync(0xffffffffu, sc[K - 2], 1);
    const uint32_t b1 = __shfl_up_sync(0xffffffffu, bk[K - 1], 1), b2 = __shfl_up_sync(0xffffffffu, bk[K - 2], 1);
    double am_s[K], am_e[K];
    bool slow = false, neg = false;
#pragma unroll
    for (int k = 0; k < K; k++) {
      am_s[k] = am[st[k]];
      am_e[k] = am[fst[k]];
      slow |= (fl[k] & kFEntry) && !(am_e[k] >= 0.0);
      neg |= !(am_s[k] >= 0.0) || !(am_e[k] >= 0.0);
    }
    double ns_[K];
    uint32_t nb_[K];
    double my_best = inf;
    // LEAN: every emission this warp adds in this frame is non-negative, so no candidate can be rejected before its
    // emission is added without also losing after it, and every entry run collapses.  One straight-line block: the K
    // slots' candidate chains are independent and interleave.  Otherwise the exact order of the reference, slot by slot.
    auto body = [&](auto lean_tag) {
      constexpr bool LEAN = decltype(lean_tag)::value;
      const bool any_slow = LEAN ? false : __any_sync(0xffffffffu, slow);
#pragma unroll
      for (int k = 0; k < K; k++) {
        const uint32_t f = fl[k], w = (f >> 8) & 0xffu;
        const bool entry = (f & kFEntry) != 0, collapse = LEAN || am_e[k] >= 0.0;
        double cur = inf;
        uint32_t cb = 0;
        auto cand = [&](bool on, double ns, double a, uint32_t b) {
          const bool rej = LEAN ? false : ns > cur;        // rejected before the emission is added (:142-145, :172-175)
          const double ns2 = ns + a;
          const bool take = on && !rej && (cur > ns2);
          cur = take ? ns2 : cur;
          cb = take ? b : cb;
        };
        const double wp = (f & kFPenalty) ? p.word_penalty : 0.0;
        const double pre = (f & kFFirstSil) ? p.forward : ((f & kFPos1) ? p.skip : p.forward);
        // word ends of the words before w ...
        if (!LEAN && any_slow)
          for (uint32_t v = 0; v < W; v++) cand(entry && !collapse && v < w, ((we[v] > limit ? inf : we[v]) + wp) + pre, am_e[k], t - 1);
        cand(entry && collapse, (pa_s[w] + wp) + pre, am_e[k], t - 1);
        // ... the word's own positions pos - 2, pos - 1, pos ...
        const bool sil = (f & kFSil) != 0;
        const double s2 = k >= 2 ? sc[k >= 2 ? k - 2 : 0] : (k == 1 ? l1 : l2);
        const double s1 = k >= 1 ? sc[k >= 1 ? k - 1 : 0] : l1;
        const uint32_t c2 = k >= 2 ? bk[k >= 2 ? k - 2 : 0] : (k == 1 ? b1 : b2);
        const uint32_t c1 = k >= 1 ? bk[k >= 1 ? k - 1 : 0] : b1;
        cand((f & kFJ2) != 0, s2 + (sil ? p.forward : p.skip), am_s[k], c2);
        cand((f & kFJ1) != 0, s1 + p.forward, am_s[k], c1);
        cand((f & kFJ0) != 0, sc[k] + (sil ? p.forward : p.loop), am_s[k], bk[k]);
        // ... the word ends of w itself and of the words after it
        if (!LEAN && any_slow)
          for (uint32_t v = 0; v < W; v++) cand(entry && !collapse && v >= w, ((we[v] > limit ? inf : we[v]) + wp) + pre, am_e[k], t - 1);
        cand(entry && collapse, (sb_s[w] + wp) + pre, am_e[k], t - 1);
        ns_[k] = cur; nb_[k] = cb;
        my_best = fmin(my_best, cur);
      }
    };
    if (!kLeanBody || __any_sync(0xffffffffu, neg)) body(std::false_type{});
    else body(std::true_type{});
#pragma unroll
    for (int k = 0; k < K; k++) { sc[k] = ns_[k]; bk[k] = nb_[k]; }
    finish_frame(my_best, true);
    if (lane == 0) {
      uint32_t rec = 0;                                    // no finite word end: word 0, bkp 0 (:117-119)
      if (e_best < inf) rec = e_word | (webk[e_word] << 16);
      tb[t] = rec;
      if (t == T && p.score_out) p.score_out[u] = e_best;
    }
  }
  if (lane == 0) {
    if (T == 0 && p.score_out) p.score_out[u] = 0.0;
    uint32_t n = 0;
    for (uint32_t t = T; t > 0; t = tb[t] >> 16)
      if ((tb[t] & 0xffffu) != p.silence_word) n++;
    uint32_t* out = p.words_raw + f0;
    uint32_t k = n;
    for (uint32_t t = T; t > 0; t = tb[t] >> 16)
      if ((tb[t] & 0xffffu) != p.silence_word) out[--k] = tb[t] & 0xffffu;
    p.n_words_out[u] = n;
  }
}

template <int K>
int launch_reg(sgmm_ctx* c, const RegParams& q, uint32_t n) {
  const size_t smem = (size_t)kRecWarps * (112 + 2 * (size_t)c->n_states) * sizeof(double);
  SGMM_CHECK(c, cudaFuncSetAttribute(k_recognize_reg<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_recognize_reg<K><<<(n + kRecWarps - 1) / kRecWarps, kRecWarps * 32, smem, c->stream>>>(q);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

// word lists of all utterances back to back: word_off[u] = words before utterance u (one block: scan of the counts,
// then every thread moves its utterances' words), so the host reads ~5 words per utterance instead of a frame-sized array
__global__ void __launch_bounds__(1024) k_rec_compact(const uint32_t* __restrict__ counts, const uint64_t* __restrict__ frame_off,
                                                      const uint32_t* __restrict__ raw, uint32_t n_utts,
                                                      uint32_t* __restrict__ word_off, uint32_t* __restrict__ out) {
  __shared__ uint32_t part[1024];
  const uint32_t per = (n_utts + 1023u) / 1024u, u0 = threadIdx.x * per, u1 = min(n_utts, u0 + per);
  uint32_t sum = 0;
  for (uint32_t u = u0; u < u1; u++) sum += counts[u];
  part[threadIdx.x] = sum;
  __syncthreads();
  for (uint32_t o = 1; o < 1024; o <<= 1) {
    const uint32_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  uint32_t at = part[threadIdx.x] - sum;
  for (uint32_t u = u0; u < u1; u++) {
    word_off[u] = at;
    const uint32_t n = counts[u];
    const uint32_t* src = raw + frame_off[u];
    for (uint32_t i = 0; i < n; i++) out[at + i] = src[i];
    at += n;
  }
  if (threadIdx.x == 1023) word_off[n_utts] = part[1023];
}

}  // namespace

// Tables of sgmm_set_lexicon: live slots are positions 0 .. max(len, 2) - 1 of every word (position 1 of a one-state
// word receives entries in the reference as well; it never continues, but it takes part in the beam's best score).
int sgmm_build_lexicon(sgmm_ctx* c, uint32_t n_words, const uint32_t* word_off, const uint32_t* word_states,
                       uint32_t silence_word) {
  if (n_words == 0 || n_words > 255) SGMM_FAIL(c, "set_lexicon: need 1..255 words");
  if (silence_word >= n_words && silence_word != 0xffffffffu) SGMM_FAIL(c, "set_lexicon: silence word out of range");
  std::vector<uint32_t> info, state, base(n_words), len(n_words), first(n_words);
  uint32_t max_state = 0;
  for (uint32_t w = 0; w < n_words; w++) {
    if (word_off[w + 1] <= word_off[w]) SGMM_FAIL(c, "set_lexicon: empty word");
    const uint32_t L = word_off[w + 1] - word_off[w];
    if (L > 255) SGMM_FAIL(c, "set_lexicon: word longer than 255 positions");
    base[w] = (uint32_t)info.size(); len[w] = L; first[w] = word_states[word_off[w]];
    for (uint32_t q = 0; q < std::max(L, 2u); q++) {
      const bool real = q < L;
      info.push_back(w | (q << 8) | (L << 16) | (real ? 0x80000000u : 0u));
      state.push_back(real ? word_states[word_off[w] + q] : first[w]);
      if (real) max_state = std::max(max_state, word_states[word_off[w] + q]);
    }
  }
  if (info.size() > 32u * kMaxSlotsPerLane * 8) SGMM_FAIL(c, "set_lexicon: lexicon too large for the per-warp search");
  c->lex_words = n_words; c->lex_live = (uint32_t)info.size(); c->lex_silence = silence_word; c->lex_max_state = max_state;
  auto up = [&](DevBuf<uint32_t>& b, const std::vector<uint32_t>& v) -> int {
    SGMM_CHECK(c, b.reserve(v.size()));
    SGMM_CHECK(c, cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    return 0;
  };
  // processing order: the entry slots (positions 0 / 1: they look at every word end) first, so that they share warp
  // iterations instead of making every iteration wait for one of them
  std::vector<uint32_t> order;
  for (uint32_t i = 0; i < info.size(); i++) if (((info[i] >> 8) & 0xffu) <= 1) order.push_back(i);
  for (uint32_t i = 0; i < info.size(); i++) if (((info[i] >> 8) & 0xffu) > 1) order.push_back(i);
  std::vector<uint32_t> tab(4 * info.size()), end(n_words);
  for (size_t oi = 0; oi < order.size(); oi++) {
    const uint32_t i = order[oi], w = info[i] & 0xffu;
    tab[4 * oi + 0] = info[i]; tab[4 * oi + 1] = i | (base[w] << 16); tab[4 * oi + 2] = state[i]; tab[4 * oi + 3] = first[w];
  }
  for (uint32_t w = 0; w < n_words; w++) end[w] = base[w] + len[w] - 1;
  // table order, for k_recognize_reg
  std::vector<uint32_t> nat(4 * info.size(), 0);
  for (size_t i = 0; i < info.size(); i++) {
    const uint32_t w = info[i] & 0xffu, pos = (info[i] >> 8) & 0xffu, L = (info[i] >> 16) & 0xffu;
    const bool real = (info[i] >> 31) != 0;
    uint32_t f = w << 8;
    if (pos <= 1) f |= kFEntry;
    if (real && pos != L - 1) f |= kFJ0;
    if (real && pos >= 1) f |= kFJ1;
    if (real && pos >= 2) f |= kFJ2;
    if (real && pos == L - 1) f |= kFEnd;
    if (pos == 1) f |= kFPos1;
    if (w != silence_word) f |= kFPenalty;
    nat[4 * i + 0] = f; nat[4 * i + 1] = state[i]; nat[4 * i + 2] = first[w];
  }
  if (up(c->d_lex_info, tab) || up(c->d_lex_base, end) || up(c->d_lex_state, nat)) return 1;
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// one launch over utterances [u0, u0 + n) whose emission scores are in d_dense_scores (row 0 = frame chunk_f0)
int sgmm_launch_recognize(sgmm_ctx* c, uint32_t u0, uint32_t n, uint64_t chunk_f0, double word_penalty,
                          double am_threshold) {
  RecParams p{};
  p.E = c->d_dense_scores.p; p.chunk_f0 = chunk_f0; p.n_states = c->n_states;
  p.frame_off = c->d_rec_frame_off.p; p.perm = c->d_rec_perm.p; p.u0 = u0; p.n_utts = n;
  p.n_words = c->lex_words; p.n_live = c->lex_live; p.silence_word = c->lex_silence;
  p.slot_tab = reinterpret_cast<const uint4*>(c->d_lex_info.p); p.word_end = c->d_lex_base.p;
  p.loop = c->tdp_loop; p.forward = c->tdp_forward; p.skip = c->tdp_skip; p.silence_state = c->silence_state;
  p.word_penalty = word_penalty; p.am_threshold = am_threshold;
  p.tb = c->d_rec_tb.p; p.tb_off = c->d_rec_tb_off.p; p.words_raw = c->d_rec_words.p;
  p.n_words_out = c->d_rec_nwords.p; p.score_out = c->d_rec_score.p;
  const size_t nl = c->lex_live, W = c->lex_words;
  if (c->rec_reg && nl <= 256 && W <= 32 && (size_t)kRecWarps * (112 + 2 * (size_t)c->n_states) * 8 <= 200 * 1024) {
    RegParams q{p, reinterpret_cast<const uint4*>(c->d_lex_state.p)};
    switch (std::max<size_t>(2, (nl + 31) / 32)) {
      case 2: return launch_reg<2>(c, q, n);
      case 3: return launch_reg<3>(c, q, n);
      case 4: return launch_reg<4>(c, q, n);
      case 5: return launch_reg<5>(c, q, n);
      case 6: return launch_reg<6>(c, q, n);
      case 7: return launch_reg<7>(c, q, n);
      default: return launch_reg<8>(c, q, n);
    }
  }
  const size_t lex_bytes = (nl * 16 + W * 4 + 15) & ~(size_t)15;
  size_t per_warp = rec_warp_doubles(nl, W, 2 * (size_t)c->n_states) * sizeof(double);
  p.stage_rows = 1;
  if (lex_bytes + kRecWarps * per_warp > 96 * 1024) {   // many states: read the emission rows from L2 instead
    p.stage_rows = 0;
    per_warp = rec_warp_doubles(nl, W, 0) * sizeof(double);
  }
  const size_t smem = lex_bytes + kRecWarps * per_warp;
  if (smem > 200 * 1024) SGMM_FAIL(c, "recognize: lexicon too large for shared memory");
  SGMM_CHECK(c, cudaFuncSetAttribute(k_recognize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_recognize<<<(n + kRecWarps - 1) / kRecWarps, kRecWarps * 32, smem, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

// after the last search launch: d_rec_woff[n_utts + 1], d_rec_compact[total words]
int sgmm_launch_rec_compact(sgmm_ctx* c, uint32_t n_utts) {
  SGMM_CHECK(c, c->d_rec_woff.reserve((size_t)n_utts + 1));
  SGMM_CHECK(c, c->d_rec_compact.reserve(c->n_frames));     // a word ends at a frame of its own: at most one per frame
  k_rec_compact<<<1, 1024, 0, c->stream>>>(c->d_rec_nwords.p, c->d_rec_frame_off.p, c->d_rec_words.p, n_utts,
                                           c->d_rec_woff.p, c->d_rec_compact.p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
