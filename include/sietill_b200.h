/* sietill_b200.h — C-ABI of the B200-native GMM acoustic-model hot path.
 *
 * Drop-in boundary for the hot path of kkromberg/SpeechRecognition's src/sietill:
 *   GMM max-approx scoring  ->  Viterbi alignment over linear HMMs with TDPs  ->  EM accumulation.
 * Plain pointers and sizes only; every buffer passed in is HOST memory owned by the caller, every
 * device buffer is owned by the context.  All functions return 0 on success, non-zero on failure
 * (the text is available from sgmm_last_error); no C++ exception crosses this boundary.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference
 * repository, src/sietill/).  The C++ classes of speechrecognition_b200/csrc/host/ (MixtureModel,
 * Aligner, TdpModel, Trainer — same public signatures as the reference's) are the callers;
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 */
#ifndef SIETILL_B200_H
#define SIETILL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sgmm_ctx sgmm_ctx;

/* AlignmentItem, Types.hpp:29-38 (8 bytes; only slot [0] of each frame is used, num-max-aligns = 1). */
typedef struct {
  uint16_t count;
  uint16_t state;
  float    weight;
} sgmm_alignment_item;

/* TDP lookup convention.  SOURCE: Aligner::align_sequence_full looks the penalty up with the state
 * of the position the transition LEAVES (Alignment.cpp:95,100,108).  DESTINATION: the pruned aligner
 * and the recogniser use the state it ENTERS (Alignment.cpp:191, Recognizer.cpp:141,170). */
enum { SGMM_TDP_SOURCE = 0, SGMM_TDP_DESTINATION = 1 };

/* ---- context -------------------------------------------------------------------------------- */
/* One context per GPU (not thread-safe; replaces the single shared MixtureModel/Aligner pair that
 * main() builds, SieTill.cpp:107-113). */
int         sgmm_create(int device, sgmm_ctx** out);
void        sgmm_destroy(sgmm_ctx* ctx);
const char* sgmm_last_error(const sgmm_ctx* ctx); /* ctx may be NULL: error of a failed sgmm_create */
const char* sgmm_version(void);

/* ---- model tables (K5) ---------------------------------------------------------------------- */
/* Uploads the parameter tables that MixtureModel::finalize/calculate_variance produce
 * (Mixtures.cpp:251-275,374-461), already resolved through mixtures_[s][k] into state-major order:
 * density j = state_off[s] + k is the k-th entry of mixtures_[s].  mean / inv_var are [D*dim] f64
 * (means_ / vars_inv_ rows), norm[D] = norm_[var_idx], log_weight[D] = mean_weights_log_[mean_idx].
 * The device keeps them in f64 and evaluates density_score_sse's exact operation order
 * (Mixtures.cpp:645-690).  One upload of the four arrays; every derived table (transposes, f32 / tf32 / fp16 operand
 * images, guard constants) is built by device kernels.  Transactional: the new model is built beside the current one
 * and swapped in on success — after a failed call the context still holds the previous model. */
int sgmm_set_model(sgmm_ctx* ctx, uint32_t dim, uint32_t n_states, const uint32_t* state_off,
                   const double* mean, const double* inv_var, const double* norm,
                   const double* log_weight);

/* MixtureModel::finalize + calculate_variance (Mixtures.cpp:374-461, :251-275) ON THE DEVICE, from the
 * statistics the last sgmm_accumulate / sgmm_accumulate_soft / sgmm_align_accumulate call left resident (already
 * all-reduced over the ranks): mean = sum_x / count, weight = count / (sum of the state's counts), log weight,
 * variance = (var_seed + sum_xx [pooled over the state / over everything]) / count - mean_ref^2 around the density's own
 * (pooling 0 = NO_POOLING), the state's (1 = MIXTURE_POOLING) or the global (2 = GLOBAL_POOLING) mean, its inverse and
 * norm = (d log 2 pi + sum_i log var_i) / 2 — the reference's operation order (sums over densities and dimensions
 * in index order; divisions, products and square roots are IEEE-exact on both sides; `log` is the device's, <= 1 ulp from
 * glibc's).  var_seed is minimal_variance_value_ (1e-4, Mixtures.cpp:243), added once whatever the number of GPUs.
 * The new parameters replace the context's model exactly as sgmm_set_model would — every derived table is rebuilt on
 * the device — without leaving the GPU; the topology (state_off) is unchanged.  Split / eliminate stay host-side
 * topology edits (fetch what they need with sgmm_get_statistics / sgmm_get_model, then sgmm_set_model). */
int sgmm_mstep(sgmm_ctx* ctx, int pooling, double var_seed);

/* Host copies of the current model's parameters, state-major as in sgmm_set_model (any pointer may be NULL); var[D*dim]
 * and weight[D] (vars_ / mean_weights_) exist only after sgmm_mstep. */
int sgmm_get_model(sgmm_ctx* ctx, double* mean, double* inv_var, double* norm, double* log_weight, double* var,
                   double* weight);

/* Host copies of the statistics of the last accumulate call (same layout and NULL rules as sgmm_accumulate's outputs);
 * they stay resident on the device until the next accumulate call. */
int sgmm_get_statistics(sgmm_ctx* ctx, double* count, double* sum_x, double* sum_xx, double* score_sum);

/* Test hook: the raw bytes of one derived table of the current model ("meanT", "invT", "meanR", "invU", "mu32", "h32",
 * "mu32T", "c32", "sguard", "Bimg", "sguard_tc", "Bimg16", "sguard16", ...).  *bytes receives the size; out may be NULL. */
int sgmm_debug_table(sgmm_ctx* ctx, const char* name, void* out, size_t cap, size_t* bytes);

/* ---- corpus --------------------------------------------------------------------------------- */
/* Corpus::features_ (Corpus.hpp:75-81): one contiguous row-major float[n_frames * dim]; one H2D copy,
 * resident until replaced. */
int sgmm_set_features(sgmm_ctx* ctx, const float* x, uint64_t n_frames, uint32_t dim);

/* Per-utterance frame ranges (segment_offsets, Training.cpp:67-76) and utterance automata
 * (Trainer::build_segment_automaton, Training.cpp:239-253) concatenated: utterance u owns frames
 * [frame_off[u], frame_off[u+1]) and positions automata[aut_off[u] .. aut_off[u+1]). */
int sgmm_set_corpus(sgmm_ctx* ctx, uint32_t n_utts, const uint64_t* frame_off,
                    const uint32_t* automata, const uint32_t* aut_off);

/* TdpModel (TdpModel.cpp:19-29): silence_state always pays `forward`. */
int sgmm_set_tdp(sgmm_ctx* ctx, double loop, double forward, double skip, uint32_t silence_state);

/* Beam of Aligner::align_sequence_pruned (Alignment.cpp:227-239, "pruning-threshold"): after every frame t >= 1 the
 * positions whose cost exceeds (best cost of the frame + pruning_threshold) die.  Applies to sgmm_align with
 * SGMM_TDP_DESTINATION only (the reference's full aligner has no beam); +inf (the default) = exact search. */
int sgmm_set_beam(sgmm_ctx* ctx, double pruning_threshold);

/* Load-time feature processing on the device: SignalAnalysis::process_features (SignalAnalysis.cpp:379-399) for every
 * utterance frame_off[u] .. frame_off[u+1] of `raw` ([n_frames][n_file] f32, the frames as they lie in the .mm2 files,
 * Corpus.cpp:95-101): first derivatives of the first n_first columns and second derivatives of the first n_second
 * ("n-features-first", "n-features-second", "deriv-step"; add_deltas, :320-337), then — if mean / stddev
 * ([n_file + n_first + n_second] f64, the two blobs of the normalization file, :371-375) are given — (x - mean) / stddev
 * with the reference's two f32 roundings (:389-394), then, if energy_max_norm, column 0 minus its maximum over the
 * utterance (:341-350).  Bit-identical to the reference.  The processed [n_frames][total] matrix becomes the
 * context's resident feature matrix exactly as after sgmm_set_features (dim = total); `out` (may be NULL) receives
 * a host copy.  Requires n_second <= n_first <= n_file and more than deriv_step frames per utterance (the reference
 * reads out of bounds otherwise). */
int sgmm_process_features(sgmm_ctx* ctx, const float* raw, uint64_t n_frames, uint32_t n_file, uint32_t n_first,
                          uint32_t n_second, uint32_t deriv_step, uint32_t n_utts, const uint64_t* frame_off,
                          const double* mean, const double* stddev, int energy_max_norm, float* out);

/* ---- scoring (K1) --------------------------------------------------------------------------- */
/* MixtureModel::min_score (Mixtures.cpp:696-713; sum_score, :719-728, when option "max_approx" is 0) for every frame
 * in [f0, f1) and every state:
 * scores[(f-f0)*n_states + s] (f64, bit-identical to the reference), argmin likewise = position k
 * within mixtures_[s].  Either output may be NULL.  Replaces the per-cell score() calls of
 * Alignment.cpp:90 / Recognizer.cpp:149,179 by one batched pass (the prepare_sequence idiom of
 * FeatureScorer.hpp:12-16). */
int sgmm_score_all(sgmm_ctx* ctx, uint64_t f0, uint64_t f1, double* scores, uint32_t* argmin);

/* ---- alignment (K1 gathered + K2) ----------------------------------------------------------- */
/* Aligner::align_sequence_full (Alignment.cpp:50-144) for every utterance of the corpus in one call
 * (the loop of Training.cpp:168-183).  The alignment stays resident on the device for
 * sgmm_accumulate; `alignment` (n_frames items, may be NULL) and `costs` (n_utts, may be NULL)
 * receive host copies.  An utterance whose end state is unreachable (S > 2T-1) gets cost +inf. */
int sgmm_align(sgmm_ctx* ctx, int tdp_convention, sgmm_alignment_item* alignment, double* costs);

/* Recognition lexicon: word w is the linear HMM word_states[word_off[w] .. word_off[w+1]) (automaton positions ->
 * state ids), as Lexicon::get_automaton_for_word returns it (Lexicon.hpp:16-33, Lexicon.cpp:11-23,70-85).
 * silence_word: Lexicon::silence_idx().  At most 255 words of at most 255 positions. */
int sgmm_set_lexicon(sgmm_ctx* ctx, uint32_t n_words, const uint32_t* word_off, const uint32_t* word_states,
                     uint32_t silence_word);

/* Replaces Recognizer::recognizeSequence_pruned (Recognizer.cpp:103-232) and the per-segment OpenMP loop of
 * Recognizer::recognize (:46-47): time-synchronous word-loop Viterbi over the lexicon with the TDPs of sgmm_set_tdp
 * (destination-state convention, :167), word_penalty ("word-penalty", 0 for silence) and beam am_threshold
 * ("am-threshold"; >= 1e30 disables pruning), for every utterance frame_off[u] .. frame_off[u+1] of the resident
 * features.  words[word_off_out[u] .. word_off_out[u+1]) = the recognised word ids (silence dropped), scores[u]
 * (optional) = score of the best word end at the last frame.  Fails if more than max_words words are produced. */
int sgmm_recognize(sgmm_ctx* ctx, uint32_t n_utts, const uint64_t* frame_off, double word_penalty,
                   double am_threshold, uint32_t* words, uint64_t max_words, uint32_t* word_off_out /* [n_utts+1] */,
                   double* scores /* [n_utts] or NULL */);

/* Upload an alignment produced elsewhere (e.g. the linear segmentation of Training.cpp:85-99). */
int sgmm_set_alignment(sgmm_ctx* ctx, const sgmm_alignment_item* alignment, uint64_t n_frames);

/* ---- EM accumulation (K3 + K4) -------------------------------------------------------------- */
/* MixtureModel::accumulate with max_approx (Mixtures.cpp:278-372) over the resident alignment:
 * for frame f with aligned state m the density is k = 0 (first_pass) or min_score(x_f, m).second with
 * the CURRENT model; then count[j] += 1, sum_x[j*dim..] += x, sum_xx[j*dim..] += x*x for
 * j = state_off[m] + k.  Outputs are per density in state-major order (count[D], sum_x[D*dim],
 * sum_xx[D*dim], all f64, starting from 0 — the caller adds the reference's 1e-4 variance seed,
 * Mixtures.cpp:243, and folds pooled variances).  score_sum receives sum_f min_score(x_f, m).first,
 * i.e. Trainer::calc_am_score's numerator (Training.cpp:585-612) for the model the E-step used
 * (0 when first_pass).  If a communicator is attached the four outputs are summed over all ranks
 * with ONE ncclAllReduce before they are copied out.  Any output may be NULL. */
int sgmm_accumulate(sgmm_ctx* ctx, int first_pass, double* count, double* sum_x, double* sum_xx,
                    double* score_sum);

/* sgmm_align followed by sgmm_accumulate(first_pass = 0) in one call — the realign + estimate pair of Trainer::train
 * (Training.cpp:168-201).  Results are those of the two calls; the alignment / cost copies overlap the E-step. */
int sgmm_align_accumulate(sgmm_ctx* ctx, int tdp_convention, sgmm_alignment_item* alignment, double* costs,
                          double* count, double* sum_x, double* sum_xx, double* score_sum);

/* The posterior-weighted branch of MixtureModel::accumulate (max_approx == false, not the first pass;
 * Mixtures.cpp:300-357) over the resident alignment: for frame f with aligned state m,
 * p_k = exp(-density_score(x_f, m, k)) normalised by their in-order sum (:310-324), entries below 1e-8 skipped
 * (:336-338); count[j] += p_k, sum_x[j*dim..] += p_k * x, sum_xx[j*dim..] += (p_k * x) * x for j = state_off[m] + k.
 * Same output layout, NULL rules and all-reduce as sgmm_accumulate.  score_sum receives
 * sum_f sum_score(x_f, m) (Mixtures.cpp:719-728) — the normaliser is the same sum — i.e. calc_am_score's numerator
 * for a model built with max_approx == false.  Density scores are bit-identical to the reference; exp / log are the
 * device's (<= 1 ulp), the statistics agree to ~1e-13 relative. */
int sgmm_accumulate_soft(sgmm_ctx* ctx, double* count, double* sum_x, double* sum_xx, double* score_sum);

/* Trainer::calc_am_score's numerator alone (Training.cpp:585-612): sum over the resident alignment of
 * score(x_f, state_f) with the current model — min_score(...).first, or sum_score when option "max_approx" is 0
 * (all-reduced if a communicator is attached). */
int sgmm_am_score_sum(sgmm_ctx* ctx, double* score_sum);

/* Per-frame results of the last sgmm_accumulate / sgmm_am_score_sum (tests, diagnostics):
 * density[f] = k chosen for frame f, score[f] = its min score.  Either may be NULL. */
int sgmm_last_frame_results(sgmm_ctx* ctx, uint32_t* density, double* score);

/* Allocation-free variant of the deterministic score sum (the numerator of Trainer::calc_am_score as sgmm_am_score_sum /
 * sgmm_accumulate compute it).  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Unlike every
 * other entry point, d_values[n] (f64) and d_sum[1] are DEVICE pointers on the context's GPU, and the call only ENQUEUES
 * its two kernels on `stream` (a cudaStream_t; NULL = the context's own stream) without synchronising: *d_sum is valid
 * once the stream has passed that point.  It calls no cudaMalloc / cudaFree or any allocating routine: the partial sums
 * live in a file-scope 128 KB __device__ array (one per GPU that loads the library, held until the process exits), which
 * limits a call to n <= 67 108 864 values (16 384 partials of 4 096 values) — larger n is rejected with an error.  Every
 * call shares that array: calls that overlap on the GPU (several streams / threads / concurrently replayed graphs) race
 * and must be serialised by the caller; nothing in it has to be cleared between calls (a call overwrites exactly the
 * partials it reads).  The first call on a device looks the array's address up (cudaGetSymbolAddress): make it outside
 * a graph capture (n = 0 is a valid warm-up).  The summation tree is fixed by n alone (bit-reproducible run to run); it
 * differs from sgmm_am_score_sum's tree, so the two agree to ~1e-15 relative, not bit for bit. */
int sgmm_score_sum_static(sgmm_ctx* ctx, const double* d_values, uint64_t n, double* d_sum, void* stream);

/* ---- options / counters -------------------------------------------------------------------- */
/* "max_approx" (default 1): what FeatureScorer::score means (MixtureModel's constructor argument, Mixtures.cpp:737-744):
 *   1 = min_score (maximum approximation), 0 = sum_score = -log sum_k exp(-density_score_k) (:719-728).  With 0,
 *   sgmm_score_all / sgmm_align / sgmm_recognize / sgmm_am_score_sum use sum scores (exact-f64 kernel; argmin outputs
 *   are 0).  Independent of which accumulate entry point is called, as in the reference (Training.cpp:145-149).
 * "sum_prune" (default 1): sum scores through the f32-pruned kernel (densities certainly more than 45 above the state's
 *   minimum are left out: at most one rounding from the full sum); 0 = plain f64 sum over every density.
 * "scorer": 0 = automatic choice by model shape, 1 = exact-f64 kernel, 2 = FP32-FFMA kernel with certified
 * f64 refinement, 3 = tcgen05 3xTF32 tensor-core GEMM with certified f64 refinement.  Every scorer returns
 * results bit-identical to the reference; the choice only moves time.
 * "rescore_cache" (default 1): sgmm_accumulate after sgmm_align with an unchanged model looks the aligned cells' scores up.
 * "async_upload" (default 0): sgmm_set_features queues chunked copies ("upload_chunks", 1..16, default 16) and returns;
 *   the host buffer must stay untouched until the next compute call returns (use sgmm_host_alloc memory).
 *   The tensor-core filter of the next sgmm_align* / sgmm_score_all / sgmm_recognize call starts behind the first chunk.
 * "tc_halves" (0 auto, 1, 2): 128-frame halves per tensor-core CTA.  "tc_pack" (default 1; read by sgmm_set_corpus):
 *   pack short utterance tails into shared 128-row halves.
 * "tc_refine" (0 auto, 1, 2): refinement kernel of the tensor-core scorer — 1 frame-major (winner's mean row gathered
 *   from L2), 2 state-major (the state's mean table in shared memory; gathered scoring of pooled models only).
 *   "tc_rescan" (0, 1): 0 two-stage rescan of near-tie cells (f32 prefilter, f64 candidates), 1 full f64 scan.
 * "tc_wide" (-1 auto, 0, 1; read by sgmm_set_model): 256-density GEMM tiles (the N = 256 kernel) instead of 128;
 *   auto picks them when some state has more than 128 densities and the tile ring fits shared memory.
 * "tc_f16" (-1 auto = on when the model allows it, 0, 1; read by sgmm_set_model and at every launch): fp16 hi / lo operands for the 256-wide tiles of pooled
 *   models (kind::f16 MMAs, 16 terms per k-step); 0 = the 3 x TF32 form.  "tc_rem" (default 1; read by sgmm_set_model): pack
 *   the K remainder of the fp16 form into one MMA (7 instead of 9 MMAs per tile and 128-frame half).
 * "tc_dense_sm" (default 1): the dense pass (sgmm_score_all, sgmm_recognize) is refined by the state-major kernel; 0 = the
 *   frame-major one.
 * "tc_epi16" (default 0): sixteen instead of eight epilogue warps per GEMM unit (measured slower; kept as an A/B switch).
 * "host_tables" (default 0; read by sgmm_set_model): 1 = derive the model tables on the host (the previous builder, kept
 *   as the checker of the device builder: same tables, same bits).
 * "allreduce" (default 1): 0 = sgmm_accumulate* skip the ncclAllReduce and return this rank's partial sums (checks).
 * "tc_pairs" (default 1): the tensor filter names the TWO candidate densities of a near-tie (both refined exactly in place)
 *   instead of declaring the cell uncertain.  "tc_civ" (default 1): the state-major refinement reads the inverse variances
 *   from the constant bank.  "tc_lq" (0 none, 1 default, n entries): per-item list of near-ties.  All three: same results.
 * "rec_run_frames" (default 0 = about 4 GB of scores): sgmm_recognize scores and searches runs of whole utterances of at
 *   most this many frames at a time (a single longer utterance is a run of its own).
 * "rec_reg" (default 1): sgmm_recognize searches lexicons of at most 256 table slots and 32 words with the
 *   register-resident kernel (same results; 0 = always the general shared-memory kernel).
 * "vit_mw" (read by sgmm_set_corpus; 0 never, 1 default = 1000, n frames): utterances at least this long are aligned by the
 *   four-warp Viterbi kernel.  Same results.
 * "sum_tc" (default 1): with max_approx = 0, the gathered alignment pass takes its sum scores from the tensor filter's
 *   candidate masks where the model shape allows (pooled, 129..256 densities per state); 0 = k_sum_pruned everywhere.
 * "tolerance" (default 0; an EXPERIMENT, NOT bit-identical to the reference): sgmm_align runs
 *   Viterbi on the filter's f32 scores without refinement (needs "vit_mw" = 0 before sgmm_set_corpus, a pooled model on the
 *   fp16 tensor form); path costs then agree to ~2e-6 relative, the E-step stays exact for the alignment it is given.
 * "tc_debug": timing experiments for the tensor-core kernel — any bit other than 32 (cycle profile, read with the
 *   "tc_prof1".."tc_prof12" counters) makes results INVALID; never set in production.  The experiments and the counters
 *   exist only in libraries built with EXTRA_NVFLAGS="-DSGMM_TC_PROF=1 -DSGMM_TC_TWOPASS_REF=1" (csrc/Makefile). */
int sgmm_set_option(sgmm_ctx* ctx, const char* key, int64_t value);
/* "fallbacks": (frame, state) cells whose f32 top-2 gap was inside the error bound and were re-scanned in
 * f64 since the counter was last read; "scorer": the scorer the automatic choice resolves to; "tc_form": 1 when the wide
 * GEMM runs fp16 hi/lo operands; "packs": packed GEMM tasks of the resident corpus; "frames_global": frames behind the
 * last all-reduced statistics / AM-score sum (every rank's shard). */
int sgmm_get_counter(sgmm_ctx* ctx, const char* key, int64_t* value);

/* ---- multi-GPU ------------------------------------------------------------------------------ */
/* One process per GPU.  Rank 0 calls sgmm_comm_unique_id (ncclGetUniqueId), ships the 128 bytes to
 * the other ranks by any side channel (torch.distributed store, MPI, a file), then every rank calls
 * sgmm_comm_init.  The reference has no counterpart (it is single-process). */
int sgmm_comm_unique_id(void* id_out, size_t id_bytes);
int sgmm_comm_init(sgmm_ctx* ctx, int rank, int world_size, const void* id, size_t id_bytes);

/* ---- instrumentation ------------------------------------------------------------------------ */
/* CUDA-event times (ms) of the stages of the most recent call, mirroring the Timer buckets of
 * Trainer::train (Training.cpp:45-49,230-234).  Index with the SGMM_T_* constants. */
enum {
  SGMM_T_H2D = 0, SGMM_T_SCORE = 1, SGMM_T_VITERBI = 2, SGMM_T_RESCORE = 3, SGMM_T_SORT = 4,
  SGMM_T_REDUCE = 5, SGMM_T_ALLREDUCE = 6, SGMM_T_D2H = 7,
  SGMM_T_SCORE_FILTER = 8,   /* part of SCORE: the f32 / tensor-core candidate kernel (k_tc_gemm) */
  SGMM_T_SCORE_REFINE = 9,   /* part of SCORE: f64 certification + rescan kernels */
  SGMM_T_MSTEP = 10,         /* sgmm_mstep / sgmm_set_model: parameter kernels + device table build (K5) */
  SGMM_T_COUNT = 11
};
int sgmm_stage_ms(sgmm_ctx* ctx, float* ms /* [SGMM_T_COUNT] */);
/* Device-side stopwatch on the context's stream (CUDA events): start, then stop returns the elapsed ms
 * of everything enqueued in between (copies included). */
int sgmm_timer_start(sgmm_ctx* ctx);
int sgmm_timer_stop(sgmm_ctx* ctx, float* ms);
/* Pinned (page-locked) host memory for callers that want asynchronous-speed H2D/D2H copies. */
void* sgmm_host_alloc(size_t bytes);
void  sgmm_host_free(void* p);
/* Number of kernel launches issued by this context since creation. */
uint64_t sgmm_launch_count(const sgmm_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* SIETILL_B200_H */
