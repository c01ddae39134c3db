// Trainer: the EM driver of the reference (src/sietill/Training.cpp:51-235) with the per-utterance alignment
// loop (:168-183) replaced by ONE batched GPU call per pass, and calc_am_score (:585-612) by a device
// reduction over the resident alignment.  Order of operations, stats lines and file names are the reference's.
#include <chrono>
#include <fstream>
#include <sstream>

#include "sietill_host.hpp"

namespace sietill_b200 {

namespace {
struct Stopwatch {
  double& acc;
  std::chrono::steady_clock::time_point t0;
  explicit Stopwatch(double& acc) : acc(acc), t0(std::chrono::steady_clock::now()) {}
  ~Stopwatch() { acc += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};
}  // namespace

Trainer::Trainer(Configuration const& config, Lexicon const& lexicon, MixtureModel& mixtures, TdpModel const& tdp_model,
                 bool max_approx)
    : min_obs_(config.get_uint("min-obs", 1u)), num_splits_(config.get_uint("num-splits", 1u)),
      num_aligns_(config.get_uint("num-aligns", 1u)), num_estimates_(config.get_uint("num-estimates", 1u)),
      num_max_aligns_(config.get_uint("num-max-aligns", 1u)),
      pruning_threshold_(config.get_double("pruning-threshold", 50.0)),
      mixture_path_(config.get_string("mixture-path", "")), alignment_path_(config.get_string("alignment-path", "")),
      training_stats_path_(config.get_string("training-stats-path", "")), max_approx_(max_approx),
      realign_(config.get_bool("realign", true)), alignment_pruning_(config.get_bool("alignment-pruning", true)),
      approx_linear_segmentation_(config.get_bool("approx-linear-segmentation", true)),
      lexicon_(lexicon), mixtures_(mixtures), tdp_model_(tdp_model), aligner_(mixtures, tdp_model, num_max_aligns_) {
}

// Training.cpp:239-253: sil w1 sil w2 ... sil
MarkovAutomaton Trainer::build_segment_automaton(WordIter segment_begin, WordIter segment_end) const {
  std::vector<MarkovAutomaton const*> parts;
  for (WordIter w = segment_begin; w != segment_end; ++w) {
    parts.push_back(&lexicon_.get_silence_automaton());
    parts.push_back(&lexicon_.get_automaton_for_word(*w));
  }
  parts.push_back(&lexicon_.get_silence_automaton());
  return MarkovAutomaton::concat(parts);
}

// Training.cpp:585-612: average of score(x_f, aligned state) with the CURRENT model.  The alignment is the
// one resident on the device (made so by the accumulate / align_corpus call that precedes every call site).
double Trainer::calc_am_score(Corpus const& corpus, Alignment const& alignment) {
  mixtures_.sync_device();
  Device& dev = mixtures_.device();
  if (mixtures_.resident_ali_ != alignment.data() || mixtures_.resident_ali_frames_ != alignment.size()) {
    std::pair<FeatureIter, FeatureIter> all = corpus.get_all_features();
    mixtures_.set_corpus_features(all.first, all.second);
    dev.check(sgmm_set_alignment(dev.ctx(), reinterpret_cast<const sgmm_alignment_item*>(alignment.data()), alignment.size()),
              "sgmm_set_alignment");
    mixtures_.resident_ali_ = alignment.data();
    mixtures_.resident_ali_frames_ = alignment.size();
    mixtures_.resident_ali_sum_ = alignment_checksum(alignment.data(), alignment.size());
  }
  double sum = 0.0;
  dev.check(sgmm_am_score_sum(dev.ctx(), &sum), "sgmm_am_score_sum");
  // with a communicator attached the sum is over every rank's shard: divide by the frames behind it
  int64_t frames = 0;
  dev.check(sgmm_get_counter(dev.ctx(), "frames_global", &frames), "sgmm_get_counter(frames_global)");
  (void)corpus;
  return sum / (double)frames;
}

void Trainer::train(Corpus const& corpus, const Alignment* initial) {
  if (num_max_aligns_ != 1) throw std::runtime_error("Trainer: num-max-aligns must be 1");
  const size_t U = corpus.get_corpus_size(), total_frames = corpus.get_total_frame_count();
  std::vector<MarkovAutomaton> segment_automata;
  segment_automata.reserve(U);
  for (SegmentIdx s = 0; s < U; s++) {
    std::pair<WordIter, WordIter> words = corpus.get_word_sequence(s);
    segment_automata.push_back(build_segment_automaton(words.first, words.second));
  }
  stats_.clear();
  std::ofstream stats_out;
  if (!training_stats_path_.empty()) {
    stats_out.open(training_stats_path_.c_str(), std::ios_base::out | std::ios_base::trunc);
    if (!stats_out.good()) throw std::runtime_error("Error opening stat file: " + training_stats_path_);
  }
  auto log_stat = [&](int i, int j, int k, double am) {
    stats_.push_back(StatLine{i, j, k, am});
    if (stats_out.is_open()) stats_out << i << " " << j << " " << k << " " << am << std::endl;
  };
  auto write_mix = [&](std::string const& name) {
    if (mixture_path_.empty()) return;   // the reference would write into the working directory; we stay silent
    Stopwatch sw(io_seconds);
    std::ofstream out((mixture_path_ + name).c_str(), std::ios_base::out | std::ios_base::trunc | std::ios_base::binary);
    if (!out.good()) throw std::runtime_error("Could not open " + mixture_path_ + name);
    mixtures_.write(out);
  };

  // ---- initial segmentation
  if (initial) {
    if (initial->size() != total_frames) throw std::runtime_error("Trainer: initial alignment length differs from the corpus");
    alignment_ = *initial;
  } else {
    alignment_.assign(total_frames, AlignmentItem());
    Stopwatch sw(align_seconds);
    for (SegmentIdx s = 0; s < U; s++) {      // Training.cpp:78-99
      std::pair<FeatureIter, FeatureIter> f = corpus.get_feature_sequence(s);
      AlignmentIter ab(alignment_.data() + corpus.frame_offsets()[s], 1), ae(alignment_.data() + corpus.frame_offsets()[s + 1], 1);
      if (approx_linear_segmentation_) linear_segmentation_approximation(segment_automata[s], f.first, f.second, ab, ae);
      else linear_segmentation_running_sums(segment_automata[s], f.first, f.second, ab, ae);
    }
  }

  std::pair<FeatureIter, FeatureIter> features = corpus.get_all_features();
  ConstAlignmentIter ali_begin(alignment_.data(), 1), ali_end(alignment_.data() + alignment_.size(), 1);
  {
    Stopwatch sw(estimate_seconds);
    mixtures_.accumulate(ali_begin, ali_end, features.first, features.second, true);
    mixtures_.finalize();
  }
  double am;
  { Stopwatch sw(score_seconds); am = calc_am_score(corpus, alignment_); }
  log_stat(-1, 0, 0, am);
  write_mix("lin.mix");

  for (unsigned i = 0; i <= num_splits_; i++) {
    if (i > 0) {
      {
        Stopwatch sw(estimate_seconds);
        mixtures_.split(2u * min_obs_);
        mixtures_.accumulate(ali_begin, ali_end, features.first, features.second, false, max_approx_);
        mixtures_.finalize();
        mixtures_.eliminate(min_obs_);
        mixtures_.accumulate(ali_begin, ali_end, features.first, features.second, false, max_approx_);
        mixtures_.finalize();
      }
      { Stopwatch sw(score_seconds); am = calc_am_score(corpus, alignment_); }
      log_stat((int)i, -1, 0, am);
    }
    for (unsigned j = 0; j < num_aligns_; j++) {
      if (realign_) {
        {
          Stopwatch sw(align_seconds);
          aligner_.align_corpus(corpus, segment_automata, alignment_, alignment_pruning_, nullptr, pruning_threshold_);
          ali_begin = ConstAlignmentIter(alignment_.data(), 1);
          ali_end = ConstAlignmentIter(alignment_.data() + alignment_.size(), 1);
        }
        if (!alignment_path_.empty()) {
          Stopwatch sw(io_seconds);
          std::ostringstream name;
          name << alignment_path_ << i << "-" << j << ".dump";
          std::ofstream out(name.str().c_str(), std::ios_base::out | std::ios_base::trunc | std::ios_base::binary);
          write_alignment(out, alignment_, num_max_aligns_);
        }
      }
      const unsigned n_est = i == 0 ? 1 : num_estimates_;   // a single Gaussian does not move by re-estimating (:197)
      for (unsigned k = 0; k < n_est; k++) {
        {
          Stopwatch sw(estimate_seconds);
          mixtures_.accumulate(ali_begin, ali_end, features.first, features.second, false, max_approx_);
          mixtures_.finalize();
        }
        { Stopwatch sw(score_seconds); am = calc_am_score(corpus, alignment_); }
        log_stat((int)i, (int)j, (int)k, am);
      }
    }
    std::ostringstream name;
    name << i << ".mix";
    write_mix(name.str());
  }
}

// ---- initial segmentation (Training.cpp:350-556) ------------------------------------------------------
namespace {

// f32 running sums of the first feature component and of its square (Training.cpp:364-381, :438-455)
void energy_prefix_sums(FeatureIter begin, FeatureIter end, std::vector<float>& sum, std::vector<float>& sq) {
  const size_t N = (size_t)(end - begin);
  sum.assign(N, 0.0f); sq.assign(N, 0.0f);
  size_t n = 0;
  for (FeatureIter it = begin; it != end; ++it, ++n) {
    const float c0 = **it;
    sum[n] = c0; sq[n] = c0 * c0;
    if (n > 0) { sum[n] += sum[n - 1]; sq[n] += sq[n - 1]; }
  }
}

// within-segment scatter of frames [b, e] (Training.cpp:549-557): f32 differences, the rest in f64
double segment_scatter(std::vector<float> const& sum, std::vector<float> const& sq, size_t b, size_t e) {
  const double d = sum[e] - sum[b - 1];
  double score = sq[e] - sq[b - 1];
  score -= (d * d) / (e - b + 1);
  return score;
}

}  // namespace

std::pair<size_t, size_t> Trainer::linear_segmentation_approximation(MarkovAutomaton const& automaton, FeatureIter feature_begin,
                                                                     FeatureIter feature_end, AlignmentIter align_begin,
                                                                     AlignmentIter align_end) const {
  const size_t N = (size_t)(feature_end - feature_begin);
  std::vector<float> sum, sq;
  energy_prefix_sums(feature_begin, feature_end, sum, sq);
  std::pair<size_t, size_t> b(N / 2 - 1, N / 2);                  // :458-459
  for (int sweep = 0; sweep < 3; sweep++) {                       // three coordinate-descent sweeps (:461)
    float best = 1e10f;                                           // a float, as in the reference: every stored best is rounded
    size_t pick = b.first;
    for (size_t n = 1; n + 1 < b.second; n++) {                   // n = 1 .. second - 2 (:469-471)
      const double c = segment_scatter(sum, sq, 1, n) + segment_scatter(sum, sq, n + 1, b.second);
      if (c < best) { pick = n; best = (float)c; }
    }
    b.first = pick;
    best = 1e10f; pick = b.second;
    for (size_t n = b.first + 1; n + 1 < N; n++) {                // n = first + 1 .. N - 2 (:488-490)
      const double c = segment_scatter(sum, sq, b.first + 1, n) + segment_scatter(sum, sq, n + 1, N - 1);
      if (c < best) { pick = n; best = (float)c; }
    }
    b.second = pick;
  }
  linear_segmentation_alignment_mapping(automaton, b, align_begin, align_end);
  return b;
}

std::pair<size_t, size_t> Trainer::linear_segmentation_running_sums(MarkovAutomaton const& automaton, FeatureIter feature_begin,
                                                                    FeatureIter feature_end, AlignmentIter align_begin,
                                                                    AlignmentIter align_end) const {
  const size_t K = 4, N = (size_t)(feature_end - feature_begin);  // three segments: silence, speech, silence
  std::vector<float> sum, sq;
  energy_prefix_sums(feature_begin, feature_end, sum, sq);
  std::vector<std::vector<float>> cost(K, std::vector<float>(N, 1e10f));
  std::vector<std::vector<size_t>> back(K, std::vector<size_t>(N, 0));
  cost[0][0] = 0.0f;
  for (size_t k = 1; k < K; k++)
    for (size_t n = 0; n < N; n++)
      for (size_t m = 0; m < n; m++) {                            // boundary m before n; segment m+1 .. n (:399-410)
        const double d = sum[n] - sum[m];
        double local = sq[n] - sq[m];
        local -= (d * d) / (n - m);
        if (cost[k][n] > cost[k - 1][m] + local) {
          back[k][n] = m;
          cost[k][n] = (float)(cost[k - 1][m] + local);
        }
      }
  std::pair<size_t, size_t> b;
  b.second = back[K - 1][N - 1];
  b.first = back[K - 2][b.second];
  linear_segmentation_alignment_mapping(automaton, b, align_begin, align_end);
  return b;
}

// Training.cpp:513-546: frames up to the first boundary -> position 0, after the second -> the last position, between
// them linearly over ALL positions with an f32 slope; only state and count are written (weight keeps its value).
void Trainer::linear_segmentation_alignment_mapping(MarkovAutomaton const& automaton, std::pair<size_t, size_t> const& boundaries,
                                                    AlignmentIter const& align_begin, AlignmentIter const& align_end) const {
  const double per_frame = ((float)automaton.num_states()) / (boundaries.second - boundaries.first);
  size_t n = 0;
  for (AlignmentIter it = align_begin; it != align_end; ++it, ++n) {
    size_t pos;
    if (n <= boundaries.first) pos = 0;
    else if (n > boundaries.second) pos = automaton.num_states() - 1;
    else pos = (size_t)(per_frame * (n - boundaries.first - 1));
    (*it)->state = automaton[pos];
    (*it)->count = 1;
  }
}

}  // namespace sietill_b200
