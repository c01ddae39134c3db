// The compute entry points: score_all (K1), align (K1 gathered + K2), accumulate (K3 + K4 [+ NCCL]).
#include <algorithm>
#include <cstring>
#include <vector>

#include "sgmm_internal.cuh"

namespace {

int require(sgmm_ctx* c, bool model, bool feats, bool corpus) {
  if (model && c->Dtot == 0) SGMM_FAIL(c, "no model: call sgmm_set_model first");
  if (feats && c->n_frames == 0) SGMM_FAIL(c, "no features: call sgmm_set_features first");
  if (model && feats && c->feat_dim != c->dim) SGMM_FAIL(c, "feature dimension differs from the model's");
  if (corpus) {
    if (c->n_utts == 0) SGMM_FAIL(c, "no corpus: call sgmm_set_corpus first");
    if (c->h_frame_off.back() != c->n_frames) SGMM_FAIL(c, "corpus frame offsets do not cover the resident features");
    if (model && c->max_aut_state >= c->n_states) SGMM_FAIL(c, "automaton names a state the model does not have");
  }
  return 0;
}

// [count D | sum_x D*d | sum_xx D*d | score_sum | frames]: one contiguous block = the all-reduce payload
inline size_t acc_len(const sgmm_ctx* c) { return (size_t)c->Dtot * (1 + 2 * (size_t)c->dim) + 2; }

void reset_stage(sgmm_ctx* c) { sgmm_stage_reset(c); }

// emission scoring through whichever scorer is selected
int launch_emission_any(sgmm_ctx* c, bool dense, uint64_t f0, uint64_t nf, double* scores, uint32_t* argmin) {
  if (!c->max_approx) {   // score() = sum_score (Mixtures.cpp:737-744): every density contributes, exact f64 only
    if (sgmm_wait_features(c)) return 1;
    bool handled = false;   // f32-pruned kernel when the model has f32 tables (dim <= 64), else the plain f64 sum
    if (sgmm_launch_emission_sum_pruned(c, dense, f0, nf, scores, argmin, &handled)) return 1;
    if (handled) return 0;
    if (dense) return sgmm_launch_emission_sum(c, nullptr, (nf + 31) / 32, true, f0, nf, scores, argmin);
    return sgmm_launch_emission_sum(c, c->d_tasks.p, c->n_tasks, false, 0, 0, nullptr, nullptr);
  }
  const int sc = sgmm_effective_scorer(c);
  if (sc == 3) return sgmm_launch_emission_tc(c, dense, f0, nf, scores, argmin);   // consumes upload chunks itself
  if (sgmm_wait_features(c)) return 1;
  if (sc == 2) {
    if (dense) return sgmm_launch_emission_ffma(c, nullptr, (nf + 63) / 64, true, f0, nf, scores, argmin);
    return sgmm_launch_emission_ffma(c, c->d_tasks64.p, c->n_tasks64, false, 0, 0, nullptr, nullptr);
  }
  if (dense) return sgmm_launch_emission(c, nullptr, (nf + 31) / 32, true, f0, nf, scores, argmin);
  return sgmm_launch_emission(c, c->d_tasks.p, c->n_tasks, false, 0, 0, nullptr, nullptr);
}

}  // namespace

// 1 = exact f64, 2 = FP32 FFMA + in-kernel refinement, 3 = tcgen05 3xTF32 GEMM + refinement kernels.
// Automatic: an f32 filter pays off once a state has enough densities to amortise the mandatory f64
// evaluation of the winner; the tensor-core form needs enough columns per state to fill MMA tiles.
int sgmm_effective_scorer(const sgmm_ctx* c) {
  if (c->scorer == 1) return 1;
  if (c->scorer == 3) return c->tc_ok ? 3 : (c->dv4 ? 2 : 1);
  if (c->scorer == 2) return c->dv4 ? 2 : 1;
  const double avg = c->n_states ? (double)c->Dtot / c->n_states : 0.0;
  if (avg >= 48.0 && c->tc_ok) return 3;
  return (avg >= 6.0 && c->dv4) ? 2 : 1;
}

extern "C" {

int sgmm_score_all(sgmm_ctx* c, uint64_t f0, uint64_t f1, double* scores, uint32_t* argmin) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (require(c, true, true, false)) return 1;
  if (f1 > c->n_frames || f0 > f1) SGMM_FAIL(c, "score_all: frame range out of bounds");
  const uint32_t N = c->n_states;
  // chunked so that the device staging buffers stay bounded (<= ~256 MB of scores)
  const uint64_t chunk = std::max<uint64_t>(64, ((256ull << 20) / (N * sizeof(double))) & ~63ull);
  SGMM_CHECK(c, c->d_dense_scores.reserve((size_t)std::min<uint64_t>(chunk, f1 - f0 + 1) * N));
  SGMM_CHECK(c, c->d_dense_argmin.reserve((size_t)std::min<uint64_t>(chunk, f1 - f0 + 1) * N));
  for (uint64_t b = f0; b < f1; b += chunk) {
    const uint64_t nf = std::min(chunk, f1 - b);
    {
      StageTimer st(c, SGMM_T_SCORE);
      if (launch_emission_any(c, true, b, nf, c->d_dense_scores.p, argmin ? c->d_dense_argmin.p : nullptr))
        return 1;
    }
    StageTimer st(c, SGMM_T_D2H);
    if (scores)
      SGMM_CHECK(c, cudaMemcpyAsync(scores + (b - f0) * N, c->d_dense_scores.p, nf * N * sizeof(double),
                                    cudaMemcpyDeviceToHost, c->stream));
    if (argmin)
      SGMM_CHECK(c, cudaMemcpyAsync(argmin + (b - f0) * N, c->d_dense_argmin.p, nf * N * sizeof(uint32_t),
                                    cudaMemcpyDeviceToHost, c->stream));
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  }
  return 0;
}

int sgmm_align(sgmm_ctx* c, int conv, sgmm_alignment_item* alignment, double* costs) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (require(c, true, true, true)) return 1;
  if (conv != SGMM_TDP_SOURCE && conv != SGMM_TDP_DESTINATION) SGMM_FAIL(c, "align: bad tdp_convention");
  SGMM_CHECK(c, c->d_E.reserve(c->E_elems + 2));
  SGMM_CHECK(c, c->d_Amin.reserve(c->E_elems + 2));
  SGMM_CHECK(c, c->d_fcol.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_ali.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_costs.reserve(c->n_utts));
  c->tol_active = false;
  {
    StageTimer st(c, SGMM_T_SCORE);
    if (launch_emission_any(c, false, 0, 0, nullptr, nullptr)) return 1;
  }
  {
    StageTimer st(c, SGMM_T_VITERBI);
    if (sgmm_launch_viterbi(c, conv)) return 1;
  }
  {
    StageTimer st(c, SGMM_T_D2H);
    if (alignment)
      SGMM_CHECK(c, cudaMemcpyAsync(alignment, c->d_ali.p, c->n_frames * sizeof(sgmm_alignment_item),
                                    cudaMemcpyDeviceToHost, c->stream));
    if (costs)
      SGMM_CHECK(c, cudaMemcpyAsync(costs, c->d_costs.p, c->n_utts * sizeof(double), cudaMemcpyDeviceToHost,
                                    c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  c->have_alignment = true;
  c->ali_has_cols = true;
  c->E_model_version = c->tol_active ? 0 : c->model_version;   // tolerance mode: no exact emission matrix to look up
  c->tol_model_version = c->tol_active ? c->model_version : 0;
  c->E_max_approx = c->max_approx;
  return 0;
}

int sgmm_set_lexicon(sgmm_ctx* c, uint32_t n_words, const uint32_t* word_off, const uint32_t* word_states,
                     uint32_t silence_word) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (!word_off || !word_states) SGMM_FAIL(c, "set_lexicon: null argument");
  if (word_off[0] != 0) SGMM_FAIL(c, "set_lexicon: offsets must start at 0");
  return sgmm_build_lexicon(c, n_words, word_off, word_states, silence_word);
}

int sgmm_recognize(sgmm_ctx* c, uint32_t n_utts, const uint64_t* frame_off, double word_penalty, double am_threshold,
                   uint32_t* words, uint64_t max_words, uint32_t* word_off_out, double* scores) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (require(c, true, true, false)) return 1;
  if (c->lex_words == 0) SGMM_FAIL(c, "no lexicon: call sgmm_set_lexicon first");
  if (c->lex_max_state >= c->n_states) SGMM_FAIL(c, "lexicon names a state the model does not have");
  if (!frame_off || !word_off_out || (!words && max_words)) SGMM_FAIL(c, "recognize: null argument");
  if (frame_off[0] != 0 || frame_off[n_utts] != c->n_frames)
    SGMM_FAIL(c, "recognize: frame offsets do not cover the resident features");
  std::vector<uint64_t> tb_off(n_utts + 1, 0);
  for (uint32_t u = 0; u < n_utts; u++) {
    if (frame_off[u + 1] <= frame_off[u]) SGMM_FAIL(c, "recognize: empty utterance");
    if (frame_off[u + 1] - frame_off[u] > 65535) SGMM_FAIL(c, "recognize: utterance longer than 65535 frames");
    tb_off[u + 1] = tb_off[u] + (frame_off[u + 1] - frame_off[u]) + 1;
  }
  SGMM_CHECK(c, c->d_rec_frame_off.reserve(n_utts + 1));
  SGMM_CHECK(c, c->d_rec_tb_off.reserve(n_utts + 1));
  SGMM_CHECK(c, c->d_rec_tb.reserve(tb_off[n_utts]));
  SGMM_CHECK(c, c->d_rec_words.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_rec_nwords.reserve(n_utts));
  SGMM_CHECK(c, c->d_rec_score.reserve(n_utts));
  SGMM_CHECK(c, cudaMemcpyAsync(c->d_rec_frame_off.p, frame_off, (n_utts + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  SGMM_CHECK(c, cudaMemcpyAsync(c->d_rec_tb_off.p, tb_off.data(), (n_utts + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  // dense emission scores for runs of whole utterances (<= ~4 GB of scores per run: the search is one warp per
  // utterance and latency-bound, so a run should hold thousands of utterances), then one warp per utterance
  const uint32_t N = c->n_states;
  const uint64_t budget = c->rec_run_frames > 0 ? (uint64_t)c->rec_run_frames
                                                : std::max<uint64_t>(65536, (4096ull << 20) / (N * sizeof(double)));
  // runs of utterances, and within each run the launch order: longest first (the search is one warp per utterance, so
  // the longest one should start with the launch, not in its last wave)
  std::vector<uint32_t> run_end, perm(n_utts);
  for (uint32_t u0 = 0; u0 < n_utts;) {
    uint32_t u1 = u0 + 1;
    while (u1 < n_utts && frame_off[u1 + 1] - frame_off[u0] <= budget) u1++;
    for (uint32_t u = u0; u < u1; u++) perm[u] = u;
    std::stable_sort(perm.begin() + u0, perm.begin() + u1, [&](uint32_t a, uint32_t b) {
      return frame_off[a + 1] - frame_off[a] > frame_off[b + 1] - frame_off[b];
    });
    run_end.push_back(u1);
    u0 = u1;
  }
  SGMM_CHECK(c, c->d_rec_perm.reserve(n_utts));
  SGMM_CHECK(c, cudaMemcpyAsync(c->d_rec_perm.p, perm.data(), n_utts * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));   // perm and tb_off are stack vectors
  uint32_t run = 0;
  for (uint32_t u0 = 0; u0 < n_utts;) {
    const uint32_t u1 = run_end[run++];
    const uint64_t f0 = frame_off[u0], nf = frame_off[u1] - f0;
    SGMM_CHECK(c, c->d_dense_scores.reserve((size_t)nf * N));
    {
      StageTimer st(c, SGMM_T_SCORE);
      if (launch_emission_any(c, true, f0, nf, c->d_dense_scores.p, nullptr)) return 1;
    }
    {
      StageTimer st(c, SGMM_T_VITERBI);
      if (sgmm_launch_recognize(c, u0, u1 - u0, f0, word_penalty, am_threshold)) return 1;
    }
    u0 = u1;
  }
  {
    StageTimer st(c, SGMM_T_VITERBI);
    if (sgmm_launch_rec_compact(c, n_utts)) return 1;
  }
  {
    StageTimer st(c, SGMM_T_D2H);
    SGMM_CHECK(c, cudaMemcpyAsync(word_off_out, c->d_rec_woff.p, ((size_t)n_utts + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    if (scores) SGMM_CHECK(c, cudaMemcpyAsync(scores, c->d_rec_score.p, n_utts * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
    const uint64_t total = word_off_out[n_utts];
    if (total > max_words) SGMM_FAIL(c, "recognize: word buffer too small");
    if (total) SGMM_CHECK(c, cudaMemcpyAsync(words, c->d_rec_compact.p, total * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int sgmm_set_alignment(sgmm_ctx* c, const sgmm_alignment_item* alignment, uint64_t n_frames) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (n_frames != c->n_frames) SGMM_FAIL(c, "set_alignment: length differs from the resident features");
  if (c->n_states)
    for (uint64_t f = 0; f < n_frames; f++)
      if (alignment[f].state >= c->n_states) SGMM_FAIL(c, "set_alignment: state index out of range");
  SGMM_CHECK(c, c->d_ali.reserve(n_frames));
  {
    StageTimer st(c, SGMM_T_H2D);
    SGMM_CHECK(c, cudaMemcpyAsync(c->d_ali.p, alignment, n_frames * sizeof(sgmm_alignment_item),
                                  cudaMemcpyHostToDevice, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  c->have_alignment = true;
  c->ali_has_cols = false;
  return 0;
}

static int rescore_common(sgmm_ctx* c, bool first_pass) {
  if (require(c, true, true, false)) return 1;
  if (!c->have_alignment) SGMM_FAIL(c, "no alignment resident: call sgmm_align or sgmm_set_alignment first");
  if (sgmm_wait_features(c)) return 1;
  SGMM_CHECK(c, c->d_fdens.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_fscore.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_cnt_u32.reserve(c->Dtot + 2));
  SGMM_CHECK(c, c->d_acc.reserve(acc_len(c)));
  SGMM_CHECK(c, cudaMemsetAsync(c->d_cnt_u32.p, 0, (c->Dtot + 1) * sizeof(uint32_t), c->stream));
  StageTimer st(c, SGMM_T_RESCORE);
  // the alignment pass already scored these cells with this very model: look the results up
  // (the resident emission matrix must hold min scores + argmins: not after an alignment with sum scores)
  const bool cached = !first_pass && c->ali_has_cols && c->E_model_version == c->model_version && c->use_cache &&
                      c->E_max_approx;
  // tolerance mode: the filter's verdicts of the alignment pass name the winner of (almost) every aligned cell
  const bool by_verdict = !first_pass && c->tol_active && c->ali_has_cols && c->tol_model_version == c->model_version;
  if (by_verdict ? sgmm_launch_rescore_verdict(c)
                 : (cached ? sgmm_launch_rescore_cached(c) : sgmm_launch_rescore(c, first_pass))) return 1;
  if (sgmm_launch_score_sum(c)) return 1;
  return 0;
}

int sgmm_accumulate(sgmm_ctx* c, int first_pass, double* count, double* sum_x, double* sum_xx,
                    double* score_sum) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (rescore_common(c, first_pass != 0)) return 1;
  if (sgmm_launch_accumulate(c)) return 1;
  if (sgmm_nccl_allreduce(c, c->d_acc.p, acc_len(c))) return 1;
  c->stats_valid = true; c->stats_model_D = c->Dtot;
  const size_t D = c->Dtot, d = c->dim;
  {
    StageTimer st(c, SGMM_T_D2H);
    if (count) SGMM_CHECK(c, cudaMemcpyAsync(count, c->d_acc.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_x) SGMM_CHECK(c, cudaMemcpyAsync(sum_x, c->d_acc.p + D, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_xx) SGMM_CHECK(c, cudaMemcpyAsync(sum_xx, c->d_acc.p + D + D * d, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (score_sum) SGMM_CHECK(c, cudaMemcpyAsync(score_sum, c->d_acc.p + D + 2 * D * d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int sgmm_accumulate_soft(sgmm_ctx* c, double* count, double* sum_x, double* sum_xx, double* score_sum) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (require(c, true, true, false)) return 1;
  if (!c->have_alignment) SGMM_FAIL(c, "no alignment resident: call sgmm_align or sgmm_set_alignment first");
  if (sgmm_wait_features(c)) return 1;
  SGMM_CHECK(c, c->d_fdens.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_fscore.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_acc.reserve(acc_len(c)));
  if (sgmm_launch_accumulate_soft(c)) return 1;
  {
    StageTimer st(c, SGMM_T_RESCORE);
    if (sgmm_launch_score_sum(c)) return 1;
  }
  if (sgmm_nccl_allreduce(c, c->d_acc.p, acc_len(c))) return 1;
  c->stats_valid = true; c->stats_model_D = c->Dtot;
  const size_t D = c->Dtot, d = c->dim;
  {
    StageTimer st(c, SGMM_T_D2H);
    if (count) SGMM_CHECK(c, cudaMemcpyAsync(count, c->d_acc.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_x) SGMM_CHECK(c, cudaMemcpyAsync(sum_x, c->d_acc.p + D, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_xx) SGMM_CHECK(c, cudaMemcpyAsync(sum_xx, c->d_acc.p + D + D * d, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (score_sum) SGMM_CHECK(c, cudaMemcpyAsync(score_sum, c->d_acc.p + D + 2 * D * d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// The statistics of the last sgmm_accumulate* call, still resident (and all-reduced): host copies on demand, so that a
// caller running the M-step on the device (sgmm_mstep) pays the 2 D (d + 1) doubles only when it needs them (.mix
// output, split / eliminate).
int sgmm_get_statistics(sgmm_ctx* c, double* count, double* sum_x, double* sum_xx, double* score_sum) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (!c->stats_valid || c->stats_model_D != c->Dtot) SGMM_FAIL(c, "get_statistics: no statistics resident");
  const size_t D = c->Dtot, d = c->dim;
  if (count) SGMM_CHECK(c, cudaMemcpyAsync(count, c->d_acc.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (sum_x) SGMM_CHECK(c, cudaMemcpyAsync(sum_x, c->d_acc.p + D, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (sum_xx) SGMM_CHECK(c, cudaMemcpyAsync(sum_xx, c->d_acc.p + D + D * d, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (score_sum) SGMM_CHECK(c, cudaMemcpyAsync(score_sum, c->d_acc.p + D + 2 * D * d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// sgmm_align followed by sgmm_accumulate(first_pass = 0) in ONE call — the realign + estimate pair of
// Trainer::train (Training.cpp:168-201).  Same results as the two calls; the host copies of the alignment and the path
// costs travel on the copy stream while the E-step kernels run, and there is one synchronisation instead of two.
int sgmm_align_accumulate(sgmm_ctx* c, int conv, sgmm_alignment_item* alignment, double* costs, double* count,
                          double* sum_x, double* sum_xx, double* score_sum) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (require(c, true, true, true)) return 1;
  if (conv != SGMM_TDP_SOURCE && conv != SGMM_TDP_DESTINATION) SGMM_FAIL(c, "align: bad tdp_convention");
  SGMM_CHECK(c, c->d_E.reserve(c->E_elems + 2));
  SGMM_CHECK(c, c->d_Amin.reserve(c->E_elems + 2));
  SGMM_CHECK(c, c->d_fcol.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_ali.reserve(c->n_frames));
  SGMM_CHECK(c, c->d_costs.reserve(c->n_utts));
  c->tol_active = false;
  {
    StageTimer st(c, SGMM_T_SCORE);
    if (launch_emission_any(c, false, 0, 0, nullptr, nullptr)) return 1;
  }
  {
    StageTimer st(c, SGMM_T_VITERBI);
    if (sgmm_launch_viterbi(c, conv)) return 1;
  }
  c->have_alignment = true;
  c->ali_has_cols = true;
  c->E_model_version = c->tol_active ? 0 : c->model_version;   // tolerance mode: no exact emission matrix to look up
  c->tol_model_version = c->tol_active ? c->model_version : 0;
  c->E_max_approx = c->max_approx;
  // alignment + costs go home on the copy stream, behind the Viterbi launches, beside the E-step
  SGMM_CHECK(c, cudaEventRecord(c->ev_fork, c->stream));
  SGMM_CHECK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_fork, 0));
  if (alignment)
    SGMM_CHECK(c, cudaMemcpyAsync(alignment, c->d_ali.p, c->n_frames * sizeof(sgmm_alignment_item),
                                  cudaMemcpyDeviceToHost, c->copy_stream));
  if (costs)
    SGMM_CHECK(c, cudaMemcpyAsync(costs, c->d_costs.p, c->n_utts * sizeof(double), cudaMemcpyDeviceToHost,
                                  c->copy_stream));
  if (rescore_common(c, false)) return 1;
  if (sgmm_launch_accumulate(c)) return 1;
  if (sgmm_nccl_allreduce(c, c->d_acc.p, acc_len(c))) return 1;
  c->stats_valid = true; c->stats_model_D = c->Dtot;
  const size_t D = c->Dtot, d = c->dim;
  {
    StageTimer st(c, SGMM_T_D2H);
    if (count) SGMM_CHECK(c, cudaMemcpyAsync(count, c->d_acc.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_x) SGMM_CHECK(c, cudaMemcpyAsync(sum_x, c->d_acc.p + D, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sum_xx) SGMM_CHECK(c, cudaMemcpyAsync(sum_xx, c->d_acc.p + D + D * d, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (score_sum) SGMM_CHECK(c, cudaMemcpyAsync(score_sum, c->d_acc.p + D + 2 * D * d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->copy_stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int sgmm_am_score_sum(sgmm_ctx* c, double* score_sum) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  reset_stage(c);
  if (!c->max_approx) {   // calc_am_score through score() = sum_score
    if (require(c, true, true, false)) return 1;
    if (!c->have_alignment) SGMM_FAIL(c, "no alignment resident: call sgmm_align or sgmm_set_alignment first");
    if (sgmm_wait_features(c)) return 1;
    SGMM_CHECK(c, c->d_fdens.reserve(c->n_frames));
    SGMM_CHECK(c, c->d_fscore.reserve(c->n_frames));
    SGMM_CHECK(c, c->d_acc.reserve(acc_len(c)));
    StageTimer st(c, SGMM_T_RESCORE);
    bool handled = false;
    if (sgmm_launch_sum_aligned_pruned(c, &handled)) return 1;
    if (!handled && sgmm_launch_sum_aligned(c)) return 1;
    if (sgmm_launch_score_sum(c)) return 1;
  } else if (rescore_common(c, false)) return 1;
  double* slot = c->d_acc.p + (size_t)c->Dtot * (1 + 2 * (size_t)c->dim);
  if (sgmm_nccl_allreduce(c, slot, 2)) return 1;
  if (score_sum) SGMM_CHECK(c, cudaMemcpyAsync(score_sum, slot, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int sgmm_last_frame_results(sgmm_ctx* c, uint32_t* density, double* score) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (c->d_fdens.cap < c->n_frames || c->n_frames == 0) SGMM_FAIL(c, "no per-frame results yet");
  if (score) SGMM_CHECK(c, cudaMemcpyAsync(score, c->d_fscore.p, c->n_frames * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  std::vector<uint32_t> glob;
  if (density) {
    glob.resize(c->n_frames);
    SGMM_CHECK(c, cudaMemcpyAsync(glob.data(), c->d_fdens.p, c->n_frames * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  if (density) {
    // global density id -> position k within its state (state_off is sorted)
    for (uint64_t f = 0; f < c->n_frames; f++) {
      const uint32_t g = glob[f];
      if (g >= c->Dtot) { density[f] = 0; continue; }   // empty mixture: min_score returns index 0
      const uint32_t s = (uint32_t)(std::upper_bound(c->h_state_off.begin(), c->h_state_off.end(), g) - c->h_state_off.begin()) - 1;
      density[f] = g - c->h_state_off[s];
    }
  }
  return 0;
}

}  // extern "C"
