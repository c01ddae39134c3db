// Host-side mirror of the reference's C++ interfaces for the GMM acoustic-model hot path
// (kkromberg/SpeechRecognition, src/sietill).  Class names, method names, argument meaning and error
// behaviour follow the reference's headers (cited per class); every compute method forwards to the CUDA
// path through the C-ABI of include/sietill_b200.h.  There is no CPU implementation of scoring, alignment
// or accumulation in here: without a B200 those methods fail loudly (std::runtime_error carrying
// sgmm_last_error).  What stays on the host is what the reference also keeps outside its hot loop: the
// M-step (finalize / split / eliminate, O(D*d) per iteration) and the byte-compatible file formats.
//
// Differences from the reference, all deliberate:
//   * DensityIdx is 32 bit (reference: 16, Types.hpp:17) so that BASELINE config 5 (288 064 densities) fits;
//   * Aligner / Trainer gain BATCHED entry points (align_corpus) — one utterance at a time cannot fill a GPU;
//     the per-utterance methods keep their reference signatures and run as a one-utterance batch;
//   * Configuration is a plain key/value map (the JSON loader is outside the hot path).
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <iosfwd>
#include <iterator>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "sietill_b200.h"

namespace sietill_b200 {

// ---- Types.hpp:14-40 ------------------------------------------------------------------------------
typedef size_t SegmentIdx;
typedef size_t WordIdx;
typedef uint16_t StateIdx;
typedef uint32_t DensityIdx;

struct MixtureDensity {
  DensityIdx mean_idx;
  DensityIdx var_idx;
  MixtureDensity(DensityIdx mean_idx, DensityIdx var_idx) : mean_idx(mean_idx), var_idx(var_idx) {}
};
typedef std::vector<MixtureDensity> Mixture;

struct AlignmentItem {  // same 8 bytes as sgmm_alignment_item
  uint16_t count;
  StateIdx state;
  float weight;
  AlignmentItem() : count(0u), state(0u), weight(0.0f) {}
  AlignmentItem(uint16_t count, StateIdx state, float weight) : count(count), state(state), weight(weight) {}
};
static_assert(sizeof(AlignmentItem) == sizeof(sgmm_alignment_item), "AlignmentItem must stay 8 bytes");
typedef std::vector<AlignmentItem> Alignment;
typedef std::vector<WordIdx>::const_iterator WordIter;

// ---- Iter.hpp:12-104: a pointer that advances `size` elements per step ----------------------------
template <typename T>
class Iter2D {
 public:
  typedef std::random_access_iterator_tag iterator_category;
  typedef T* value_type;
  typedef ptrdiff_t difference_type;
  typedef T** pointer;
  typedef T*& reference;

  size_t size;
  Iter2D(T* pos, size_t size) : size(size), pos_(pos) {}
  T* operator*() const { return pos_; }
  T* operator[](difference_type i) const { return pos_ + i * (difference_type)size; }
  Iter2D& operator++() { pos_ += size; return *this; }
  Iter2D operator++(int) { Iter2D r(*this); pos_ += size; return r; }
  Iter2D& operator--() { pos_ -= size; return *this; }
  Iter2D& operator+=(difference_type d) { pos_ += d * (difference_type)size; return *this; }
  Iter2D& operator-=(difference_type d) { pos_ -= d * (difference_type)size; return *this; }
  Iter2D operator+(difference_type d) const { return Iter2D(pos_ + d * (difference_type)size, size); }
  Iter2D operator-(difference_type d) const { return Iter2D(pos_ - d * (difference_type)size, size); }
  difference_type operator-(Iter2D const& o) const { return (pos_ - o.pos_) / (difference_type)size; }
  bool operator==(Iter2D const& o) const { return pos_ == o.pos_ && size == o.size; }
  bool operator!=(Iter2D const& o) const { return !(*this == o); }
  bool operator<(Iter2D const& o) const { return pos_ < o.pos_; }

 private:
  T* pos_;
};
typedef Iter2D<const float> FeatureIter;
typedef Iter2D<AlignmentItem> AlignmentIter;
typedef Iter2D<const AlignmentItem> ConstAlignmentIter;

// ---- Config.hpp: key/value lookups with defaults (the reference's Parameter<T>(config) idiom) ------
class Configuration {
 public:
  Configuration() {}
  Configuration& set(std::string const& key, std::string const& value) { values_[key] = value; return *this; }
  Configuration& set(std::string const& key, double value);
  Configuration& set(std::string const& key, unsigned value) { return set(key, std::to_string(value)); }
  Configuration& set(std::string const& key, bool value) { return set(key, std::string(value ? "true" : "false")); }
  bool has(std::string const& key) const { return values_.count(key) != 0; }
  std::string get_string(std::string const& key, std::string const& dflt) const;
  double get_double(std::string const& key, double dflt) const;     // throws std::runtime_error on a non-number
  unsigned get_uint(std::string const& key, unsigned dflt) const;
  bool get_bool(std::string const& key, bool dflt) const;

 private:
  std::map<std::string, std::string> values_;
};

// ---- MarkovAutomaton.hpp:17-67 ---------------------------------------------------------------------
struct MarkovAutomaton {
  std::vector<StateIdx> states;
  MarkovAutomaton() {}
  MarkovAutomaton(StateIdx start, uint16_t num, uint16_t repetitions);
  StateIdx first_state() const { return states.front(); }
  StateIdx last_state() const { return states.back(); }
  size_t num_states() const { return states.size(); }
  StateIdx max_state() const;
  StateIdx& operator[](size_t i) { return states[i]; }
  StateIdx operator[](size_t i) const { return states[i]; }
  static MarkovAutomaton concat(std::vector<MarkovAutomaton const*> automata);
};

// ---- Lexicon.hpp:16-33 ------------------------------------------------------------------------------
class Lexicon {
 public:
  WordIdx add_word(std::string const& orth, uint16_t num_states, uint16_t state_repetitions, bool silence = false);
  MarkovAutomaton const& get_silence_automaton() const { return automata_.at(silence_); }
  MarkovAutomaton const& get_automaton_for_word(WordIdx w) const { return automata_.at(w); }
  StateIdx num_states() const { return next_state_; }
  WordIdx num_words() const { return automata_.size(); }
  WordIdx silence_idx() const { return silence_; }
  WordIdx operator[](std::string const& orth) const;

 private:
  std::vector<std::string> orth_;
  std::vector<MarkovAutomaton> automata_;
  WordIdx silence_ = 0;
  StateIdx next_state_ = 0;
};
Lexicon build_sietill_lexicon();   // Lexicon.cpp:70-85: silence (1 state) + 11 German digit words, 106 states

// ---- Corpus.hpp:53-81: the accessors Trainer / Recognizer use, over caller-owned arrays -------------
class Corpus {
 public:
  // features: row-major float[n_frames * dim]; frame_off[U + 1] in FRAMES; words: concatenated word ids,
  // word_off[U + 1].  Nothing is copied except the two offset tables.
  Corpus(const float* features, size_t dim, std::vector<size_t> frame_off, std::vector<WordIdx> words,
         std::vector<size_t> word_off);
  size_t get_corpus_size() const { return frame_off_.size() - 1; }
  size_t get_total_frame_count() const { return frame_off_.back(); }
  size_t get_max_seq_length() const;
  size_t get_features_per_timeframe() const { return dim_; }
  std::pair<WordIter, WordIter> get_word_sequence(SegmentIdx s) const;
  std::pair<FeatureIter, FeatureIter> get_feature_sequence(SegmentIdx s) const;
  std::pair<FeatureIter, FeatureIter> get_all_features() const;
  std::pair<size_t, size_t> get_feature_offsets(SegmentIdx s) const;   // in floats, as the reference's
  std::vector<size_t> const& frame_offsets() const { return frame_off_; }

 private:
  const float* features_;
  size_t dim_;
  std::vector<size_t> frame_off_;
  std::vector<WordIdx> words_;
  std::vector<size_t> word_off_;
};

// ---- TdpModel.hpp:12-30 -----------------------------------------------------------------------------
class TdpModel {
 public:
  const StateIdx silence_state;
  TdpModel(Configuration const& config, StateIdx silence_state);   // "tdp-loop" / "tdp-forward" / "tdp-skip"
  double score(StateIdx to, size_t jump) const;                     // TdpModel.cpp:19-29
  double loop() const { return tdp_loop_; }
  double forward() const { return tdp_forward_; }
  double skip() const { return tdp_skip_; }

 private:
  const double tdp_loop_, tdp_forward_, tdp_skip_;
};

// ---- FeatureScorer.hpp:12-16 ------------------------------------------------------------------------
class FeatureScorer {
 public:
  virtual ~FeatureScorer() {}
  virtual void prepare_sequence(FeatureIter const& start, FeatureIter const& end) = 0;
  virtual double score(FeatureIter const& iter, StateIdx state_idx) const = 0;
};

// One GPU context (sgmm_ctx) with RAII and exceptions; shared by MixtureModel and Aligner.
class Device {
 public:
  explicit Device(int device_index = 0);
  ~Device();
  Device(Device const&) = delete;
  Device& operator=(Device const&) = delete;
  sgmm_ctx* ctx() const { return ctx_; }
  int index() const { return index_; }
  void check(int rc, const char* what) const;   // throws std::runtime_error(what + sgmm_last_error)

 private:
  sgmm_ctx* ctx_ = nullptr;
  int index_ = 0;
};

// ---- SignalAnalysis.hpp:17-118, the load-time half: process_features on the GPU ---------------------------
// Same configuration keys and defaults as the reference (SignalAnalysis.cpp:46-56); the normalization file is the
// reference's (two blobs of n_features_total doubles: mean, stddev, :363-375).  process_features replaces the
// per-segment call of Corpus::read (Corpus.cpp:101) by ONE batched device pass over all segments; the processed
// matrix stays resident on `dev` as its feature matrix (sgmm_process_features).
class SignalAnalysis {
 public:
  const bool energy_max_norm;
  const size_t n_features_in_file, n_features_first, n_features_second, n_features_total, deriv_step;

  explicit SignalAnalysis(Configuration const& config);
  void read_normalization_file(std::istream& in);                       // SignalAnalysis.cpp:371-375
  void set_normalization(std::vector<double> const& mean, std::vector<double> const& stddev);
  // raw: [n_frames][n_features_in_file]; frame_off: n_segments + 1 segment boundaries (in frames);
  // out: [n_frames][n_features_total] or nullptr (keep the result on the device only)
  void process_features(Device& dev, const float* raw, size_t n_frames, std::vector<size_t> const& frame_off,
                        float* out) const;

 private:
  bool apply_mean_var_normalization_ = false;
  std::vector<double> mean_, stddev_;
};

// ---- Mixtures.hpp:16-93 -----------------------------------------------------------------------------
class MixtureModel : public FeatureScorer {
 public:
  enum VarianceModel { GLOBAL_POOLING, MIXTURE_POOLING, NO_POOLING };

  const size_t dimension;
  const VarianceModel var_model;

  // config keys: "load-mixtures-from" (+ "action" == "recognize"), as Mixtures.cpp:158-193; "device" (GPU index);
  // "device-mstep" (default true): accumulate() leaves the statistics on the GPU and finalize() runs there
  // (sgmm_mstep) — the host copies of the accumulators / parameters are fetched lazily, only when a
  // host-side method needs them (split, eliminate, write, the per-sequence scoring API); false = the statistics come
  // home after every accumulate() and finalize() is the host code below.
  MixtureModel(Configuration const& config, size_t dimension, size_t num_mixtures, VarianceModel var_model,
               bool max_approx);
  ~MixtureModel();

  void reset_accumulators();                                             // Mixtures.cpp:237-245
  // E-step on the GPU (K3 + K4); max_approx == false is the posterior-weighted branch (sgmm_accumulate_soft).
  void accumulate(ConstAlignmentIter alignment_begin, ConstAlignmentIter alignment_end, FeatureIter feature_begin,
                  FeatureIter feature_end, bool first_pass, bool max_approx = true);
  void finalize();                                                       // Mixtures.cpp:374-461 (device, or host f64)
  void split(size_t min_obs);                                            // Mixtures.cpp:465-543
  void eliminate(double min_obs);                                        // Mixtures.cpp:547-576
  size_t num_densities() const;                                          // Mixtures.cpp:580-582

  std::pair<double, DensityIdx> min_score(FeatureIter const& iter, StateIdx mixture_idx) const;  // :696-713
  double sum_score(FeatureIter const& iter, StateIdx mixture_idx, std::vector<double>* weights) const;  // :719-728
  void prepare_sequence(FeatureIter const& start, FeatureIter const& end) override;  // scores the whole sequence (K1)
  double score(FeatureIter const& iter, StateIdx mixture_idx) const override;        // lookup, or a 1-frame K1 call

  void read(std::istream& in);                                           // .mix v2, Mixtures.cpp:748-843
  void write(std::ostream& out) const;                                   // Mixtures.cpp:845-878

  // ---- additions for the batched path
  Device& device() const { return *device_; }
  size_t num_mixtures() const { return mixtures_.size(); }
  // uploads the parameter tables if they changed since the last upload (sgmm_set_model)
  void sync_device() const;
  // sum over the corpus of min_score(x_f, aligned state).first — Trainer::calc_am_score's numerator
  // (Training.cpp:585-612) as a by-product of the last accumulate() with first_pass == false.
  double last_score_sum() const { return last_score_sum_; }
  // the GPU keeps the corpus features resident between calls; identified by (pointer, frame count)
  void set_corpus_features(FeatureIter begin, FeatureIter end) const;
  // attach one rank of a multi-GPU job: accumulate() then all-reduces with ONE ncclAllReduce
  void comm_init(int rank, int world_size, const void* nccl_unique_id, size_t id_bytes);

  // raw tables (tests; reference layout, Mixtures.hpp:72-87)
  struct Tables {
    std::vector<double> means, mean_acc, mean_w, mean_wlog, mean_wacc, vars, vars_inv, var_acc, var_wacc, norm;
    std::vector<uint64_t> mean_refs, var_refs;
    std::vector<uint32_t> mix_offsets, mean_idx, var_idx;   // mixtures_ flattened, state-major
  };
  Tables export_tables() const;
  void import_tables(Tables const& t);

 private:
  friend class Aligner;
  friend class Trainer;
  friend class Recognizer;
  void calculate_variance(DensityIdx var_idx, const double* mean);
  void finalize_host();
  MixtureDensity create_mixture_density(DensityIdx var_index);
  void mark_dirty() { ++version_; }
  // device-resident M-step: which side holds the current accumulators / parameters
  bool device_mstep_ = true;
  mutable bool stats_on_device_ = false;    // the accumulators of the last accumulate() are on the GPU only
  mutable bool params_on_device_ = false;   // the parameters of the last finalize() are on the GPU only
  void ensure_host() const;                 // fetch whatever is device-only into the host arrays

  bool max_approx_;
  std::vector<double> means_, mean_accumulators_, mean_weights_, mean_weights_log_, mean_weight_accumulators_;
  std::vector<size_t> mean_refs_;
  std::vector<double> vars_, vars_inv_, var_accumulators_, var_weight_accumulators_;
  std::vector<size_t> var_refs_;
  std::vector<double> norm_;
  std::vector<Mixture> mixtures_;
  std::vector<MixtureDensity> stale_first_;   // last erased entry of an emptied mixture (see finalize)
  double minimal_variance_value_;

  std::unique_ptr<Device> device_;
  uint64_t version_ = 1;
  mutable uint64_t uploaded_version_ = 0;
  mutable const float* resident_feat_ = nullptr;
  mutable size_t resident_frames_ = 0;
  mutable uint64_t resident_feat_print_ = 0;             // sampled fingerprint of the resident feature block
  mutable const AlignmentItem* resident_ali_ = nullptr;   // host range whose content is resident on the GPU
  mutable size_t resident_ali_frames_ = 0;
  mutable uint64_t resident_ali_sum_ = 0;                 // checksum of that range when it was made resident
  double last_score_sum_ = 0.0;
  // prepare_sequence cache
  mutable const float* seq_begin_ = nullptr;
  mutable size_t seq_frames_ = 0;
  mutable uint64_t seq_version_ = 0;
  mutable std::vector<double> seq_scores_;
  mutable std::vector<double> seq_sum_scores_;   // sum scores of the prepared sequence (models with max_approx == false)
  mutable std::vector<uint32_t> seq_argmin_;
  mutable std::unique_ptr<Device> scratch_;          // one-off sequences (per-utterance API)
  mutable uint64_t scratch_version_ = 0;
  Device& scratch() const;
};

// ---- Alignment.hpp:17-66 ----------------------------------------------------------------------------
class Aligner {
 public:
  Aligner(MixtureModel const& mixtures, TdpModel const& tdp_model, size_t max_aligns);

  // Alignment.cpp:50-144; runs as a one-utterance batch on the GPU (K1 gathered + K2).
  double align_sequence_full(FeatureIter feature_begin, FeatureIter feature_end, MarkovAutomaton const& reference,
                             AlignmentIter align_begin, AlignmentIter align_end);
  // Alignment.cpp:149-288: destination-state TDP convention, beam of pruning_threshold after every frame
  // (:227-239), backtrace from the highest position alive in the last frame.  +inf = exact search.
  double align_sequence_pruned(FeatureIter feature_begin, FeatureIter feature_end, MarkovAutomaton const& reference,
                               AlignmentIter align_begin, AlignmentIter align_end, double pruning_threshold);

  // Batched: every utterance of the corpus in one call (the loop of Training.cpp:168-183).  `alignment`
  // covers all frames (num_max_aligns == 1); costs (optional) receives the per-utterance path costs.
  void align_corpus(Corpus const& corpus, std::vector<MarkovAutomaton> const& automata, Alignment& alignment,
                    bool destination_tdp = false, std::vector<double>* costs = nullptr,
                    double pruning_threshold = std::numeric_limits<double>::infinity());

 private:
  double align_one(FeatureIter fb, FeatureIter fe, MarkovAutomaton const& ref, AlignmentIter ab, AlignmentIter ae,
                   int convention, double beam);
  MixtureModel const& mixtures_;
  TdpModel const& tdp_model_;
  size_t max_aligns_;
  uint64_t corpus_key_ = 0;            // content hash (frame offsets + automata) of the resident corpus layout
  bool corpus_key_valid_ = false;
};

void write_alignment(std::ostream& out, Alignment const& alignment, size_t max_aligns);   // Alignment.cpp:303-308
void read_alignment(std::istream& in, Alignment& alignment, size_t& max_aligns);          // Alignment.cpp:310-317

// ---- Recognizer.hpp:22-73 ----------------------------------------------------------------------------
struct EDAccumulator {   // Util.hpp: edit-distance counts
  size_t total_count = 0, substitute_count = 0, insert_count = 0, delete_count = 0;
  EDAccumulator& operator+=(EDAccumulator const& o);
};
EDAccumulator editDistance(WordIter ref_begin, WordIter ref_end, WordIter hyp_begin, WordIter hyp_end);

class Recognizer {
 public:
  // config keys and defaults as Recognizer.cpp:31-34: "am-threshold" 20, "word-penalty" 10, "pruned-search" true
  // (the unpruned search is a stub in the reference, :234-240, and is not offered), "max-recognition-runs" 1000.
  Recognizer(Configuration const& config, Lexicon const& lexicon, MixtureModel const& scorer, TdpModel const& tdp_model);

  // Recognizer.cpp:37-99: recognises the first max-recognition-runs segments in ONE batched GPU call (the reference's
  // OpenMP loop), then accumulates the edit distances.  Results are kept in the public fields below.
  void recognize(Corpus const& corpus);
  // Recognizer.cpp:103-232 for one sequence (a one-utterance batch on the scratch device context).
  void recognizeSequence_pruned(FeatureIter feature_begin, FeatureIter feature_end, std::vector<WordIdx>& output);

  std::vector<std::vector<WordIdx>> hypotheses;   // per segment, after recognize()
  EDAccumulator errors;
  size_t ref_total = 0, sentence_errors = 0;
  double wer = 0.0, ser = 0.0, seconds = 0.0;

 private:
  void upload_lexicon(Device& dev) const;
  const double am_threshold_, word_penalty_;
  const bool pruned_search_;
  const size_t max_recognition_runs_;
  Lexicon const& lexicon_;
  MixtureModel const& scorer_;
  TdpModel const& tdp_model_;
};

// ---- Training.hpp:17-110 ----------------------------------------------------------------------------
class Trainer {
 public:
  // config keys and defaults as Training.cpp:24-39: "min-obs", "num-splits", "num-aligns", "num-estimates",
  // "num-max-aligns" (must be 1), "pruning-threshold", "mixture-path", "alignment-path",
  // "training-stats-path", "realign", "alignment-pruning".
  Trainer(Configuration const& config, Lexicon const& lexicon, MixtureModel& mixtures, TdpModel const& tdp_model,
          bool max_approx);

  // Training.cpp:51-235.  Without `initial`, the reference's initial segmentation is run (Training.cpp:78-99: the
  // coordinate-descent approximation when "approx-linear-segmentation" is true — the default — else the exact
  // 3-segment dynamic programme); `initial` (all frames, one item each) replaces it.
  void train(Corpus const& corpus, const Alignment* initial = nullptr);
  // Training.cpp:350-424 / :429-510: speech / silence boundaries on the first feature component, then the linear
  // mapping of frames to automaton positions (:513-546).  Host code as in the reference.
  std::pair<size_t, size_t> linear_segmentation_running_sums(MarkovAutomaton const& automaton, FeatureIter feature_begin,
                                                             FeatureIter feature_end, AlignmentIter align_begin,
                                                             AlignmentIter align_end) const;
  std::pair<size_t, size_t> linear_segmentation_approximation(MarkovAutomaton const& automaton, FeatureIter feature_begin,
                                                              FeatureIter feature_end, AlignmentIter align_begin,
                                                              AlignmentIter align_end) const;
  void linear_segmentation_alignment_mapping(MarkovAutomaton const& automaton, std::pair<size_t, size_t> const& boundaries,
                                             AlignmentIter const& align_begin, AlignmentIter const& align_end) const;
  MarkovAutomaton build_segment_automaton(WordIter segment_begin, WordIter segment_end) const;   // :239-253
  MixtureModel const& get_mixtures() const { return mixtures_; }

  struct StatLine { int split, align, estimate; double am_score; };   // one line of the training-stats file
  std::vector<StatLine> const& stats() const { return stats_; }
  Alignment const& alignment() const { return alignment_; }
  // Timer buckets of Training.cpp:45-49,230-234 (seconds, host wall clock)
  double align_seconds = 0, estimate_seconds = 0, score_seconds = 0, io_seconds = 0;

 private:
  double calc_am_score(Corpus const& corpus, Alignment const& alignment);   // Training.cpp:585-612
  const unsigned min_obs_, num_splits_, num_aligns_, num_estimates_, num_max_aligns_;
  const double pruning_threshold_;
  const std::string mixture_path_, alignment_path_, training_stats_path_;
  const bool max_approx_, realign_, alignment_pruning_, approx_linear_segmentation_;
  Lexicon const& lexicon_;
  MixtureModel& mixtures_;
  TdpModel const& tdp_model_;
  Aligner aligner_;
  std::vector<StatLine> stats_;
  Alignment alignment_;
};

// 64-bit checksum of an alignment range (guards the "already resident on the GPU" shortcut against the
// caller editing the host copy in place).
uint64_t alignment_checksum(const AlignmentItem* items, size_t n);

// Contiguous utterance ranges balanced by frame count: rank r owns utterances
// [bounds[r], bounds[r + 1]).  Deterministic; every rank computes the same table.
std::vector<size_t> shard_utterances(std::vector<size_t> const& frame_off, int world_size);

}  // namespace sietill_b200
