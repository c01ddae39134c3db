// Internal declarations shared by the .cu files of libsietill_b200.so (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "sietill_b200.h"

#define SGMM_WARP 32
#define SGMM_TILE_T 16      // Viterbi: time steps per emission tile
#define SGMM_STAGES 3       // Viterbi: emission tiles in flight (cp.async.bulk ring)
#define SGMM_VIT_MW_WARPS 4        // k_viterbi_mw: warps per utterance
#define SGMM_VIT_MW_MIN_FRAMES 1000  // utterances at least this long take the multi-warp Viterbi kernel
// ---- layout of the gathered (ragged) emission matrix, shared by the GEMM verdicts (d_ci1) and the argmin array (d_Amin):
// utterance u holds ceil(T_u / 16) tiles of [column][16 frames]; inside a column's 16-frame run (128 B of doubles, one
// full line) frame t sits at (t ^ column) & 15:
//     cell(u, t, j) = E_off[u] + (t >> 4) * 16 * ncol_u + j * 16 + ((t ^ j) & 15)
//   * a warp whose lanes are 32 consecutive frames of one column (GEMM epilogue: lane = TMEM row; refinement: lane = frame)
//     touches two full lines, instead of 32 sectors in the frame-major [T][ncol] form;
//   * a Viterbi tile (16 frames, all columns) is still ONE contiguous run of 16 * ncol doubles = one bulk copy, and the
//     xor spreads a frame's columns over the 16 eight-byte bank pairs exactly as the row-major tile did.
// Frames T_u .. of the last tile are padding: never written, never read by the recursion.
#define SGMM_E_TILE 16
__host__ __device__ __forceinline__ uint64_t sgmm_cell_rel(uint32_t ncol, uint32_t t, uint32_t j) {
  return (uint64_t)(t >> 4) * (16u * ncol) + (j << 4) + ((t ^ j) & 15u);
}
__host__ __device__ __forceinline__ uint64_t sgmm_cells_of(uint32_t T, uint32_t ncol) {
  return (uint64_t)((T + 15u) >> 4) * 16u * ncol;
}
#define SGMM_SORT_CHUNK 2048  // radix sort: keys per warp-chunk
#define SGMM_RED_BLOCK 256    // segmented reduce: frames per partial

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;  // elements
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = n + n / 8 + 64;
    cudaError_t e = cudaMalloc(&p, want * sizeof(T));
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// state-major refinement: an item (one state of one block of utterances) lists at most kSmMaxEntries utterances and at
// most kSmMaxChunks chunks (32 consecutive frames of ONE utterance, the last chunk of an utterance may be short);
// sgmm_set_corpus ends a block where one more utterance would break either bound (see there)
constexpr uint32_t kSmMaxEntries = 256;
constexpr uint32_t kSmBlockFrames = 65536;
constexpr uint32_t kSmMaxChunks = kSmBlockFrames / 32 + kSmMaxEntries;
constexpr uint32_t kSmTargetChunks = 2048;   // a block keeps growing past kSmBlockFrames until its busiest state has this many chunks

struct sgmm_pack_row {   // one row of a packed GEMM task (sgmm_set_corpus)
  uint64_t cell;     // first cell of this frame's 16-frame tile in the ragged emission matrix + (t & 15)  (sgmm_cell_rel)
  uint32_t frame;    // absolute frame index
  uint32_t colmap;   // first entry of this row's segment in the column map (union column -> own column, 0xFFFF: none)
};

struct ViterbiClass {  // utterances that share one launch (same lanes-per-position shape + smem size)
  int C = 0;                 // positions per lane
  int W = 1;                 // warps per utterance: 1 = k_viterbi, 4 = k_viterbi_mw (long utterances)
  size_t smem = 0;           // dynamic shared memory per CTA
  std::vector<uint32_t> utts;
  DevBuf<uint32_t> d_utts;
};

// Everything sgmm_set_model / sgmm_mstep derive from the model parameters (K5).  Kept apart from the context so that a new
// model is built into a spare set and swapped in only when every step succeeded.
struct SgmmModel {
  // exact f64 tables, dim-pair-major: tabT[ip * Dtot + j] = {mean.x, mean.y} etc.
  uint32_t dim = 0, npair = 0, n_states = 0, Dtot = 0, max_dens = 0;
  std::vector<uint32_t> h_state_off;
  // the parameters as sgmm_set_model takes them (state-major, row-major f64): the device table builder's input
  DevBuf<double> d_mean_src, d_inv_src;
  DevBuf<double2> d_meanT, d_invT;
  DevBuf<double> d_mean_tail, d_inv_tail, d_norm, d_logw;
  // row-major copies for one-thread-per-cell refinement: meanR[D][dimP] (dimP = dim rounded up to 4, rows 32-byte
  // aligned), the DISTINCT variance rows invU[V][dimP] and var_of[D] (pooled models: V = n_states or 1)
  uint32_t dimP = 0, n_var_rows = 0;
  DevBuf<double> d_meanR, d_invU;
  DevBuf<uint32_t> d_var_of;
  DevBuf<uint32_t> d_state_off;
  // f32 tables of the FFMA / tensor scorers: [D][4*dv4] mean and h = 1/(2 var), c = norm - log w, per-state guard constants
  uint32_t dv4 = 0;
  DevBuf<float4> d_mu32, d_h32, d_sguard;
  DevBuf<float4> d_mu32T;          // [dv4][D]: the mean table density-minor (two-stage rescan)
  DevBuf<float> d_c32;
  // tensor-core scorer (k_score_tc.cu): 3xTF32 operands, K-major, KP = padded K
  uint32_t tc_KP = 0, tc_ksteps = 0, tc_Dpad = 0;
  int tc_full = 0;                 // 1: K = x^2 | x | 1;  0: variances pooled per state, K = x | 1
  uint32_t tc_tile_rows = 128;     // densities per B tile: 256 selects the N = 256 GEMM kernel (k_tc_gemm256)
  bool tc_ok = false;
  DevBuf<float> d_Bimg;
  DevBuf<uint32_t> d_tile_first;    // [n_states + 1] first B tile of every state
  DevBuf<uint2> d_tile_info;        // per tile {offset into d_Bimg in 16-byte units, densities N (multiple of 16)}
  DevBuf<float4> d_sguard_tc;      // per state {max ||g||, max ||h||, max |c|, -}
  DevBuf<uint32_t> d_col_off;
  float tc_eps = 0.f;
  // fp16 operand form of the wide GEMM (build_f16_image / k_tc_prep16)
  bool tc_f16_ok = false;
  uint32_t tc_ks16 = 0;
  // packed remainder (option "tc_rem"): K = 16 q + r with 0 < 3 r <= 16 — the r trailing terms of the three products
  // hi.hi, lo.hi, hi.lo share ONE k-step ([Ahi r | Alo r | Ahi r] . [Bhi r | Bhi r | Blo r]) instead of three mostly
  // empty ones: 3 q + 1 MMAs per unit instead of 3 (q + 1) (7 instead of 9 at d = 33).  0 = not packed.
  uint32_t tc_rem16 = 0;
  float tc_eps16 = 0.f;
  DevBuf<__half> d_Bimg16;
  DevBuf<uint2> d_tile_info16;
  DevBuf<float4> d_sguard16;
  bool pooled_states = false;     // every state's densities share one inverse-variance row
  // topology-only facts kept for sgmm_mstep (the M-step changes parameters, not the layout)
  std::vector<uint32_t> h_col_off, h_tile_first;
  size_t img_elems = 0, img16_elems = 0;   // sizes of d_Bimg (floats) / d_Bimg16 (halves)
  std::vector<uint32_t> h_var_of;          // variance row of every density (consecutive-duplicate folding)

  void release_all() {
    d_mean_src.release(); d_inv_src.release(); d_meanT.release(); d_invT.release(); d_mean_tail.release();
    d_inv_tail.release(); d_norm.release(); d_logw.release(); d_meanR.release(); d_invU.release(); d_var_of.release();
    d_state_off.release(); d_mu32.release(); d_h32.release(); d_sguard.release(); d_mu32T.release(); d_c32.release();
    d_Bimg.release(); d_tile_first.release(); d_tile_info.release(); d_sguard_tc.release(); d_col_off.release();
    d_Bimg16.release(); d_tile_info16.release(); d_sguard16.release();
  }
};

struct sgmm_ctx : SgmmModel {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  uint64_t launches = 0;
  // stage stopwatch: a stage may be bracketed several times per call (chunk loops); every bracket gets its own event
  // pair out of a ring, sgmm_stage_ms adds them up (ring full: the pairs are folded into stage_ms after a sync)
  static constexpr int kStageRing = 8;
  float stage_ms[SGMM_T_COUNT] = {0};
  int stage_n[SGMM_T_COUNT] = {0};
  cudaEvent_t ev[2 * SGMM_T_COUNT * kStageRing] = {nullptr};
  cudaEvent_t tev[2] = {nullptr, nullptr};
  // side streams so that independent launches (the Viterbi shape classes) overlap
  static constexpr int kAux = 16;
  cudaStream_t aux[kAux] = {nullptr};
  cudaEvent_t ev_fork = nullptr, ev_join[kAux] = {nullptr};

  // ---- scorer state that is not part of the model tables
  int scorer = 0;   // 0 auto, 1 exact f64, 2 f32 FFMA + certified refinement
  DevBuf<unsigned long long> d_fallbacks;
  int tc_wide = -1;                // option "tc_wide" (read by sgmm_set_model): -1 auto, 0 never, 1 whenever it fits
  bool tc_A_valid = false;
  DevBuf<float> d_Ahi, d_Alo;
  DevBuf<float2> d_xnorm;
  DevBuf<uint32_t> d_ci1;
  DevBuf<uint4> d_summask;         // sum scores on the tensor path: [cell][4] {mask lo, mask hi, sub-part minimum, width}
  DevBuf<float> d_e32;             // tolerance mode: the f32 emission matrix (cell-indexed like d_E)
  DevBuf<unsigned long long> d_rescan;   // [0] = count, [1..] = cell ids
  size_t rescan_cap = 0;
  int host_tables = 0;             // option "host_tables": 1 = build the K5 tables on the host (the checker of the device builder)
  SgmmModel spare;                 // sgmm_set_model / sgmm_mstep build here, then swap: a failed call leaves the model untouched
  DevBuf<int> d_model_flags;       // device builder: [0] non-finite fp16-image entries
  // device-resident M-step (k_model.cu): variances [D][dim], weights [D], per-state totals, global-pooling scratch
  DevBuf<double> d_mstep_tmp;
  bool var_valid = false;          // d_mstep_tmp holds the current model's variances / weights (after sgmm_mstep)
  bool stats_valid = false;        // d_acc holds the statistics of the last sgmm_accumulate* call ...
  uint32_t stats_model_D = 0;      // ... for a model with this many densities

  // ---- features
  uint64_t n_frames = 0;
  uint32_t feat_dim = 0;
  DevBuf<float> d_x;
  // sgmm_process_features (k_features.cu): raw file frames, utterance offsets, {mean, stddev}
  DevBuf<float> d_raw;
  DevBuf<uint64_t> d_feat_off;
  DevBuf<double> d_feat_norm;
  // option "async_upload": sgmm_set_features returns once the chunked copies are queued on copy_stream; the GEMM
  // scorer consumes the chunks as they land (k_score_tc.cu), every other consumer waits for the last one
  static constexpr int kFeatChunks = 16;
  bool async_upload = false;
  int feat_chunks = 16;                       // option "upload_chunks" (1..kFeatChunks)
  int feat_pending = 0;                       // chunks of the upload in flight (0: features resident)
  uint64_t feat_chunk_end[kFeatChunks] = {0};   // exclusive end frame of every chunk
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t feat_ev[kFeatChunks] = {nullptr};

  // ---- corpus
  uint32_t n_utts = 0, max_aut_state = 0, max_uniq = 0;   // max_uniq: most distinct states in one automaton
  std::vector<uint64_t> h_frame_off;
  std::vector<uint32_t> h_aut_off;
  DevBuf<uint64_t> d_frame_off, d_E_off;
  DevBuf<uint32_t> d_aut, d_aut_off, d_uniq, d_uniq_off, d_nuniq_pad;
  DevBuf<uint16_t> d_pos2uniq;
  DevBuf<uint4> d_tasks128, d_tasks256;
  uint64_t n_tasks128 = 0, n_tasks256 = 0;
  std::vector<uint64_t> h_t128_end, h_t256_end;   // exclusive end frame of every task (ascending)
  bool tc_pack = true;             // option "tc_pack": pack short utterance tails into shared 128-row GEMM halves
  size_t n_packs = 0;
  // state-major refinement (k_tc_refine_sm): work items (block of utterances, state) and, per item, the
  // utterances of the block that visit the state
  int tc_refine = 0;               // option "tc_refine": 0 automatic, 1 frame-major kernel, 2 state-major kernel
  int tol = 0;                     // option "tolerance" (experiment, default off): Viterbi consumes the tensor filter's f32 scores; no
                                   // refinement; only the aligned cells are re-scored in f64 (K3).  NOT bit-identical to the reference.
  bool tol_active = false;         // the last alignment pass ran in tolerance mode (its emission matrix is f32: d_e32)
  int sum_tc = 1;                  // option "sum_tc": sum scores of the gathered pass through the tensor filter's candidate masks (0: k_sum_pruned)
  bool tc_sum_pass = false;        // set around sgmm_launch_emission_tc by the sum-score launcher
  int tc_dense_sm = 1;             // option "tc_dense_sm": the dense pass is refined by the state-major kernel (0: frame-major gather)
  int tc_pairs = 1;                // option "tc_pairs": the filter names the TWO candidate columns of a near-tie (refined in place) instead of "uncertain"
  int tc_civ = 1;                  // option "tc_civ": state-major refinement reads the inverse variances from the constant bank (0: shared memory)
  int tc_lq = 1;                   // option "tc_lq": state-major refinement queues an item's near-ties as one run (0: cell by cell)
  int tc_rem = 1;                  // option "tc_rem" (read by sgmm_set_model): pack the K remainder of the fp16 form
  int tc_f16 = -1;                 // option "tc_f16": -1 / 1 use the fp16 operand form when available, 0 never
  uint64_t tc_A_key = 0;           // shape / form the resident A rows were prepared for
  int tc_A_form = 0;               // which form the resident A rows (d_Ahi / d_Alo / d_xnorm) are in: 0 tf32, 1 fp16
  int tc_epi16 = 0;                // option "tc_epi16": sixteen epilogue warps per GEMM unit (k_tc_gemm256<.., E16>)
  int tc_rescan = 0;               // option "tc_rescan": 0 two-stage (f32 prefilter, f64 candidates), 1 full f64 scan
  size_t n_sm_items = 0;
  DevBuf<uint4> d_sm_items;        // {state, first entry, entries, chunks}
  DevBuf<uint4> d_sm_ent_a;        // {E_off[u] lo, hi, ncol | column << 16, chunks before}
  DevBuf<uint2> d_sm_ent_b;        // {first frame (< 2^32: the launcher checks), frames}
  DevBuf<uint32_t> d_sm_ivoff;     // per item: offset of its state's inverse-variance row (k_sm_ivoff, every refinement launch)
  DevBuf<uint4> d_dsm_items, d_dsm_ent_a;   // the dense pass's items of the state-major refinement (built per launch)
  DevBuf<uint2> d_dsm_ent_b;
  DevBuf<uint4> d_pack_desc;
  DevBuf<uint32_t> d_pack_uniq;
  DevBuf<sgmm_pack_row> d_pack_rows;
  DevBuf<uint16_t> d_pack_colmap;
  int tc_debug = 0;                // option "tc_debug": timing experiments (k_score_tc.cu), never set in production
  int tc_halves = 0;               // option "tc_halves": 0 auto, 1 / 2 = frame halves (128 rows each) per GEMM CTA
  DevBuf<uint4> d_tasks, d_tasks64;   // emission warp-tasks {utt, t0, 0, 0}: 32-frame and 64-frame granules
  uint64_t n_tasks = 0, n_tasks64 = 0, E_elems = 0;
  DevBuf<double> d_E;             // ragged emission matrix, utterance u: [ceil(T_u / 16)][nuniq_pad_u][16] (sgmm_cell_rel)
  DevBuf<uint16_t> d_Amin;        // argmin density of every emission cell (same indexing as d_E)
  DevBuf<uint16_t> d_fcol;        // per frame: emission column of the aligned position (written by the backtrace)
  uint64_t model_version = 1, E_model_version = 0, tol_model_version = 0;
  uint64_t serial = 0;            // unique per context of this process (sgmm_create): keys caches that outlive a context
  bool use_cache = true;          // option "rescore_cache"
  bool max_approx = true;         // option "max_approx": score() = min_score (1) or sum_score (0), Mixtures.cpp:737-744
  bool E_max_approx = true;       // which of the two the resident emission matrix holds
  bool sum_prune = true;          // option "sum_prune": sum scores through the f32-pruned kernel (k_sum_pruned)
  bool ali_has_cols = false;      // resident alignment came from sgmm_align (d_fcol valid)
  std::vector<ViterbiClass> vclasses;
  long long rec_run_frames = 0;          // option "rec_run_frames": frames of dense scores per search launch (0 = ~4 GB of scores)
  int rec_reg = 1;                       // option "rec_reg": small lexicons (<= 256 slots, <= 32 words) are searched by k_recognize_reg
  int vit_mw = SGMM_VIT_MW_MIN_FRAMES;   // option "vit_mw" (read by sgmm_set_corpus): utterances of at least this many frames take k_viterbi_mw; 0 = never
  DevBuf<uint32_t> d_all_states;  // 0..n_states-1 (dense scoring)

  // ---- tdp
  double tdp_loop = 0, tdp_forward = 0, tdp_skip = 0;
  double beam = __builtin_huge_val();   // sgmm_set_beam: pruning threshold of the destination-convention aligner
  uint32_t silence_state = 0;

  // ---- alignment + per-frame E-step results
  bool have_alignment = false;
  DevBuf<sgmm_alignment_item> d_ali;
  DevBuf<double> d_costs;
  DevBuf<uint32_t> d_fdens;       // global density id per frame
  DevBuf<double> d_fscore;

  // ---- accumulation
  DevBuf<uint32_t> d_keys[2], d_vals[2], d_hist, d_cnt_u32, d_seg_off, d_item_off, d_soft_keys;
  DevBuf<double> d_partial, d_acc;   // d_acc = [count D | sum_x D*d | sum_xx D*d | score_sum 1]
  DevBuf<double> d_score_part;
  DevBuf<double> d_dense_scores;
  DevBuf<uint32_t> d_dense_argmin;

  // ---- recognition (k_recognize.cu): lexicon tables + per-call buffers
  uint32_t lex_words = 0, lex_live = 0, lex_silence = 0, lex_max_state = 0;
  DevBuf<uint32_t> d_lex_info, d_lex_state, d_lex_base;   // slot table in processing order (uint4), in table order (uint4), word-end slots
  DevBuf<uint64_t> d_rec_frame_off, d_rec_tb_off;
  DevBuf<uint32_t> d_rec_tb, d_rec_words, d_rec_nwords, d_rec_perm, d_rec_woff, d_rec_compact;
  DevBuf<double> d_rec_score;

  // ---- NCCL (dlopen'ed lazily)
  void* nccl_comm = nullptr;
  int rank = 0, world = 1;
  bool allreduce = true;          // option "allreduce": 0 leaves the per-rank partial sums un-reduced (checks)
};

#define SGMM_CHECK(ctx, call)                                                             \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess) {                                                              \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                   ":" + std::to_string(__LINE__) + ")";                                  \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

#define SGMM_FAIL(ctx, msg) do { (ctx)->err = (msg); return 1; } while (0)

#define SGMM_LAUNCH_CHECK(ctx)                     \
  do {                                             \
    (ctx)->launches++;                             \
    SGMM_CHECK(ctx, cudaGetLastError());           \
  } while (0)

// CUDA-event brackets on the context's stream; sgmm_stage_ms sums every bracket of a stage since the call began
inline void sgmm_stage_fold(sgmm_ctx* c, int idx) {
  for (int i = 0; i < c->stage_n[idx]; i++) {
    cudaEvent_t a = c->ev[2 * (idx * sgmm_ctx::kStageRing + i)], b = c->ev[2 * (idx * sgmm_ctx::kStageRing + i) + 1];
    float v = 0.f;
    if (cudaEventSynchronize(b) == cudaSuccess && cudaEventElapsedTime(&v, a, b) == cudaSuccess) c->stage_ms[idx] += v;
  }
  c->stage_n[idx] = 0;
}
inline void sgmm_stage_begin(sgmm_ctx* c, int idx) {
  if (c->stage_n[idx] == sgmm_ctx::kStageRing) sgmm_stage_fold(c, idx);
  cudaEventRecord(c->ev[2 * (idx * sgmm_ctx::kStageRing + c->stage_n[idx])], c->stream);
}
inline void sgmm_stage_end(sgmm_ctx* c, int idx) {
  cudaEventRecord(c->ev[2 * (idx * sgmm_ctx::kStageRing + c->stage_n[idx]) + 1], c->stream);
  c->stage_n[idx]++;
}
inline void sgmm_stage_reset(sgmm_ctx* c) {
  for (int i = 0; i < SGMM_T_COUNT; i++) { c->stage_ms[i] = 0.f; c->stage_n[i] = 0; }
}
struct StageTimer {
  sgmm_ctx* c; int idx;
  StageTimer(sgmm_ctx* c, int idx) : c(c), idx(idx) { sgmm_stage_begin(c, idx); }
  ~StageTimer() { sgmm_stage_end(c, idx); }
};

// kernels / launchers (defined in k_*.cu)
int sgmm_launch_emission(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                         uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin);
int sgmm_launch_emission_ffma(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                              uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin);
int sgmm_launch_emission_tc(sgmm_ctx* c, bool dense, uint64_t dense_f0, uint64_t dense_nframes, double* d_scores,
                           uint32_t* d_argmin);
int sgmm_effective_scorer(const sgmm_ctx* c);
int sgmm_launch_viterbi(sgmm_ctx* c, int tdp_convention);
int sgmm_launch_rec_compact(sgmm_ctx* c, uint32_t n_utts);
int sgmm_build_lexicon(sgmm_ctx* c, uint32_t n_words, const uint32_t* word_off, const uint32_t* word_states,
                       uint32_t silence_word);
int sgmm_launch_recognize(sgmm_ctx* c, uint32_t u0, uint32_t n, uint64_t chunk_f0, double word_penalty,
                          double am_threshold);
// make the compute stream wait for a queued feature upload (option "async_upload"); no-op when resident
inline int sgmm_wait_features(sgmm_ctx* c) {
  if (c->feat_pending) {
    SGMM_CHECK(c, cudaStreamWaitEvent(c->stream, c->feat_ev[c->feat_pending - 1], 0));
    c->feat_pending = 0;
  }
  return 0;
}
int sgmm_launch_rescore(sgmm_ctx* c, bool first_pass);
int sgmm_launch_rescore_cached(sgmm_ctx* c);
int sgmm_launch_rescore_verdict(sgmm_ctx* c);   // tolerance mode: winners from the filter's verdicts of the alignment pass
int sgmm_launch_accumulate(sgmm_ctx* c);
int sgmm_sort_frames(sgmm_ctx* c, const uint32_t* keys, uint32_t n_keys, const uint32_t** sorted);
// non-max-approx branch (k_soft.cu)
int sgmm_launch_emission_sum(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                             uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin);
int sgmm_launch_sum_aligned(sgmm_ctx* c);
int sgmm_launch_sum_aligned_pruned(sgmm_ctx* c, bool* handled);
// f32-pruned sum scores (k_score_tc.cu); *handled = false when the model shape has no f32 tables
int sgmm_launch_emission_sum_pruned(sgmm_ctx* c, bool dense, uint64_t dense_f0, uint64_t dense_nframes, double* d_scores,
                                    uint32_t* d_argmin, bool* handled);
int sgmm_launch_accumulate_soft(sgmm_ctx* c);
int sgmm_launch_score_sum(sgmm_ctx* c);
int sgmm_nccl_allreduce(sgmm_ctx* c, double* d_buf, size_t n);

// Exact f64 density score in the reference's operation order (Mixtures.cpp:645-690) with the frame's
// features read from global memory (f32 -> f64 is exact).  Shared by the refinement kernels.
struct ExactTables {
  const double2* meanT;
  const double2* invT;
  const double* mean_tail;
  const double* inv_tail;
  const double* norm;
  const double* logw;
  uint32_t dim, npair, Dtot;
  const double* meanR;     // row-major twins (see sgmm_ctx)
  const double* invU;
  const uint32_t* var_of;
  uint32_t dimP;
};
#ifdef __CUDACC__
__device__ __forceinline__ double sgmm_exact_score(const ExactTables& m, const float* __restrict__ xf, uint32_t k) {
  double l0 = 0.0, l1 = 0.0;
  const double2* mp = m.meanT + k;
  const double2* vp = m.invT + k;
  for (uint32_t ip = 0; ip < m.npair; ip++) {
    const double2 mu = __ldg(mp), iv = __ldg(vp);
    double a = __dsub_rn((double)__ldg(xf + 2 * ip), mu.x);
    double b = __dsub_rn((double)__ldg(xf + 2 * ip + 1), mu.y);
    a = __dmul_rn(a, a); b = __dmul_rn(b, b);
    a = __dmul_rn(a, iv.x); b = __dmul_rn(b, iv.y);
    l0 = __dadd_rn(l0, a); l1 = __dadd_rn(l1, b);
    mp += m.Dtot; vp += m.Dtot;
  }
  double dist = __dadd_rn(l0, l1);
  if (m.dim & 1) {
    const double c = __dsub_rn((double)__ldg(xf + m.dim - 1), __ldg(m.mean_tail + k));
    dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), __ldg(m.inv_tail + k)));
  }
  const double s = __dadd_rn(__ldg(m.norm + k), __dmul_rn(dist, 0.5));
  return __dsub_rn(s, __ldg(m.logw + k));
}
// Same value, same operation order, from the row-major tables: one thread walks ONE density row (contiguous
// 16-byte loads; the variance row is shared by every density of a pooled state and stays in L1).
__device__ __forceinline__ double sgmm_exact_score_rm(const ExactTables& m, const float* __restrict__ xf, uint32_t k) {
  const uint32_t v = __ldg(m.var_of + k);
  const double2* mp = reinterpret_cast<const double2*>(m.meanR + (size_t)k * m.dimP);
  const double2* vp = reinterpret_cast<const double2*>(m.invU + (size_t)v * m.dimP);
  double l0 = 0.0, l1 = 0.0;
  for (uint32_t ip = 0; ip < m.npair; ip++) {
    const double2 mu = __ldg(mp + ip), iv = __ldg(vp + ip);
    double a = __dsub_rn((double)__ldg(xf + 2 * ip), mu.x);
    double b = __dsub_rn((double)__ldg(xf + 2 * ip + 1), mu.y);
    a = __dmul_rn(a, a); b = __dmul_rn(b, b);
    a = __dmul_rn(a, iv.x); b = __dmul_rn(b, iv.y);
    l0 = __dadd_rn(l0, a); l1 = __dadd_rn(l1, b);
  }
  double dist = __dadd_rn(l0, l1);
  if (m.dim & 1) {
    const double mu = __ldg(m.meanR + (size_t)k * m.dimP + m.dim - 1);
    const double iv = __ldg(m.invU + (size_t)v * m.dimP + m.dim - 1);
    const double c = __dsub_rn((double)__ldg(xf + m.dim - 1), mu);
    dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), iv));
  }
  const double s = __dadd_rn(__ldg(m.norm + k), __dmul_rn(dist, 0.5));
  return __dsub_rn(s, __ldg(m.logw + k));
}
#endif
