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
      lexicon_(lexicon), mixtures_(mixtures), tdp_model_(tdp_model), aligner_(mixtures, tdp_model, num_max_aligns_) {
  if (!max_approx) throw std::runtime_error("Trainer: only max-approx training is on the GPU path");
  if (alignment_pruning_ && !(pruning_threshold_ >= 1e30))
    throw std::runtime_error("Trainer: the GPU aligner is the exact search; set alignment-pruning false (or "
                             "pruning-threshold >= 1e30 for the pruned aligner's conventions without a beam)");
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
  return sum / (double)corpus.get_total_frame_count();
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
    for (SegmentIdx s = 0; s < U; s++) {
      const size_t f0 = corpus.frame_offsets()[s], T = corpus.frame_offsets()[s + 1] - f0;
      const size_t S = segment_automata[s].num_states();
      for (size_t t = 0; t < T; t++) alignment_[f0 + t] = AlignmentItem(1, segment_automata[s][t * S / T], 0.0f);
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
          aligner_.align_corpus(corpus, segment_automata, alignment_, alignment_pruning_);
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

}  // namespace sietill_b200
