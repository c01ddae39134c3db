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
#include <algorithm>
#include <vector>

#include "sgmm_internal.cuh"

namespace {

constexpr int kRecWarps = 4;       // utterances per CTA
constexpr int kMaxSlotsPerLane = 16;

struct RecParams {
  const double* E;             // [chunk frames][n_states] emission scores, row 0 = frame chunk_f0
  uint64_t chunk_f0;
  uint32_t n_states;
  const uint64_t* frame_off;   // [n_utts + 1] over the resident features
  uint32_t u0, n_utts;         // utterances [u0, u0 + n_utts) in this launch
  // lexicon tables (device)
  uint32_t n_words, n_live, silence_word;
  const uint32_t* slot_info;   // [n_live] word | pos << 8 | len << 16 | (p < len ? 1 : 0) << 31
  const uint32_t* slot_state;  // [n_live] automaton state of the slot (first state for slots past the word's end)
  const uint32_t* word_base;   // [n_words] first live slot of the word
  const uint32_t* word_len;    // [n_words]
  const uint32_t* word_first;  // [n_words] first state
  const uint32_t* slot_order;  // [n_live] slots in processing order: the word-entry slots (positions 0 and 1) first
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

__global__ void __launch_bounds__(kRecWarps * 32) k_recognize(RecParams p) {
  extern __shared__ unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t nl = p.n_live, W = p.n_words;
  // shared lexicon tables, then per-warp hypothesis buffers
  uint32_t* s_info = reinterpret_cast<uint32_t*>(smem_raw);
  uint32_t* s_state = s_info + nl;
  uint32_t* s_base = s_state + nl;
  uint32_t* s_len = s_base + W;
  uint32_t* s_first = s_len + W;
  uint32_t* s_ord = s_first + W;
  const size_t lex_words = (size_t)3 * nl + 3 * W;
  double* warp_mem = reinterpret_cast<double*>(smem_raw + ((lex_words * 4 + 15) & ~(size_t)15));
  const size_t row_d = p.stage_rows ? (size_t)2 * p.n_states : 0;   // two staged emission rows
  const size_t per_warp = (size_t)2 * nl + ((2 * nl + 3) / 4) + row_d;   // doubles: 2 score buffers + 2 u16 bkp buffers + rows
  double* sc0 = warp_mem + warp * per_warp;
  double* sc1 = sc0 + nl;
  uint16_t* bk0 = reinterpret_cast<uint16_t*>(sc1 + nl);
  uint16_t* bk1 = bk0 + nl;
  double* srow = sc0 + (size_t)2 * nl + ((2 * nl + 3) / 4);
  for (uint32_t i = threadIdx.x; i < nl; i += blockDim.x) {
    s_info[i] = p.slot_info[i]; s_state[i] = p.slot_state[i]; s_ord[i] = p.slot_order[i];
  }
  for (uint32_t i = threadIdx.x; i < W; i += blockDim.x) {
    s_base[i] = p.word_base[i]; s_len[i] = p.word_len[i]; s_first[i] = p.word_first[i];
  }
  __syncthreads();
  const uint32_t ul = blockIdx.x * kRecWarps + warp;
  if (ul >= p.n_utts) return;
  const uint32_t u = p.u0 + ul;
  const uint64_t f0 = p.frame_off[u];
  const uint32_t T = (uint32_t)(p.frame_off[u + 1] - f0);
  const double* E = p.E + (f0 - p.chunk_f0) * p.n_states;
  uint32_t* tb = p.tb + p.tb_off[u];
  const double inf = __longlong_as_double(0x7ff0000000000000ll);

  for (uint32_t i = lane; i < nl; i += 32) { sc0[i] = inf; sc1[i] = inf; bk0[i] = 0; bk1[i] = 0; }
  __syncwarp();
  if (lane == 0) { sc0[0] = 0.0; tb[0] = 0; }     // start hypothesis: word 0, position 0, score 0 (:121)
  __syncwarp();

  double* prev = sc0; double* next = sc1;
  uint16_t* pbk = bk0; uint16_t* nbk = bk1;
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
      const uint32_t i = s_ord[oi];
      const uint32_t info = s_info[i];
      const uint32_t w = info & 0xffu, pos = (info >> 8) & 0xffu, L = (info >> 16) & 0xffu;
      const bool in_word = (info >> 31) != 0;          // pos < L (a real position of the word)
      const uint32_t base = s_base[w];
      double cur = inf;
      uint16_t bk = 0;
      double am_e = 0.0, pre_e = 0.0;
      const bool entry = pos <= 1;
      if (entry) {
        const uint32_t first = s_first[w];
        am_e = am[first];
        pre_e = tdp_of(p, first, pos + 1);
      }
      const double wp = (w != p.silence_word) ? p.word_penalty : 0.0;
      // sources in table order: word ends of the words before w ...
      if (entry)
        for (uint32_t v = 0; v < w; v++) {
          const double S = prev[s_base[v] + s_len[v] - 1];
          if (S == inf) continue;
          double ns = S + wp + pre_e;
          if (ns > cur) continue;
          ns += am_e;
          if (cur > ns) { cur = ns; bk = (uint16_t)(t - 1); }
        }
      // ... then the word's own positions, ascending: pos-2, pos-1, pos (never its last position, which only exits) ...
      if (in_word) {
        const uint32_t st = s_state[i];
        const double am_s = am[st];
#pragma unroll
        for (int jump = 2; jump >= 0; jump--) {
          if ((int)pos - jump < 0) continue;
          const uint32_t q = pos - (uint32_t)jump;
          if (q == L - 1) continue;
          const double S = prev[base + q];
          if (S == inf) continue;
          double ns = S + tdp_of(p, st, (uint32_t)jump);
          if (ns > cur) continue;
          ns += am_s;
          if (cur > ns) { cur = ns; bk = pbk[base + q]; }
        }
      }
      // ... then the word ends of w itself and of the words after it
      if (entry)
        for (uint32_t v = w; v < W; v++) {
          const double S = prev[s_base[v] + s_len[v] - 1];
          if (S == inf) continue;
          double ns = S + wp + pre_e;
          if (ns > cur) continue;
          ns += am_e;
          if (cur > ns) { cur = ns; bk = (uint16_t)(t - 1); }
        }
      next[i] = cur;
      nbk[i] = bk;
      my_best = fmin(my_best, cur);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_best = fmin(my_best, __shfl_xor_sync(0xffffffffu, my_best, o));
    const double limit = my_best + p.am_threshold;
    __syncwarp();
    for (uint32_t i = lane; i < nl; i += 32)
      if (next[i] > limit) next[i] = inf;
    __syncwarp();
    // best word end of this frame: lowest score, lowest word on ties (strict '>' in table order)
    double e_best = inf;
    uint32_t e_word = 0xffffffffu;
    for (uint32_t v = lane; v < W; v += 32) {
      const double S = next[s_base[v] + s_len[v] - 1];
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
      if (e_best < inf) rec = e_word | ((uint32_t)nbk[s_base[e_word] + s_len[e_word] - 1] << 16);
      tb[t] = rec;
      if (t == T && p.score_out) p.score_out[u] = e_best;
    }
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
  // processing order: the entry slots (positions 0 / 1: they walk every word end) first, so that they share warp
  // iterations instead of making every iteration wait for one of them
  std::vector<uint32_t> order;
  for (uint32_t i = 0; i < info.size(); i++) if (((info[i] >> 8) & 0xffu) <= 1) order.push_back(i);
  for (uint32_t i = 0; i < info.size(); i++) if (((info[i] >> 8) & 0xffu) > 1) order.push_back(i);
  if (up(c->d_lex_info, info) || up(c->d_lex_state, state) || up(c->d_lex_base, base) || up(c->d_lex_len, len) ||
      up(c->d_lex_first, first) || up(c->d_lex_order, order))
    return 1;
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// one launch over utterances [u0, u0 + n) whose emission scores are in d_dense_scores (row 0 = frame chunk_f0)
int sgmm_launch_recognize(sgmm_ctx* c, uint32_t u0, uint32_t n, uint64_t chunk_f0, double word_penalty,
                          double am_threshold) {
  RecParams p{};
  p.E = c->d_dense_scores.p; p.chunk_f0 = chunk_f0; p.n_states = c->n_states;
  p.frame_off = c->d_rec_frame_off.p; p.u0 = u0; p.n_utts = n;
  p.n_words = c->lex_words; p.n_live = c->lex_live; p.silence_word = c->lex_silence;
  p.slot_info = c->d_lex_info.p; p.slot_state = c->d_lex_state.p; p.word_base = c->d_lex_base.p;
  p.word_len = c->d_lex_len.p; p.word_first = c->d_lex_first.p; p.slot_order = c->d_lex_order.p;
  p.loop = c->tdp_loop; p.forward = c->tdp_forward; p.skip = c->tdp_skip; p.silence_state = c->silence_state;
  p.word_penalty = word_penalty; p.am_threshold = am_threshold;
  p.tb = c->d_rec_tb.p; p.tb_off = c->d_rec_tb_off.p; p.words_raw = c->d_rec_words.p;
  p.n_words_out = c->d_rec_nwords.p; p.score_out = c->d_rec_score.p;
  const size_t nl = c->lex_live, W = c->lex_words;
  const size_t lex_bytes = ((3 * nl + 3 * W) * 4 + 15) & ~(size_t)15;
  size_t per_warp = (2 * nl + ((2 * nl + 3) / 4) + 2 * (size_t)c->n_states) * sizeof(double);
  p.stage_rows = 1;
  if (lex_bytes + kRecWarps * per_warp > 96 * 1024) {   // many states: read the emission rows from L2 instead
    p.stage_rows = 0;
    per_warp = (2 * nl + ((2 * nl + 3) / 4)) * sizeof(double);
  }
  const size_t smem = lex_bytes + kRecWarps * per_warp;
  if (smem > 200 * 1024) SGMM_FAIL(c, "recognize: lexicon too large for shared memory");
  SGMM_CHECK(c, cudaFuncSetAttribute(k_recognize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_recognize<<<(n + kRecWarps - 1) / kRecWarps, kRecWarps * 32, smem, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
