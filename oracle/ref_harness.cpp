// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI harness around the UNMODIFIED reference classes of /root/reference/src/sietill
// (MixtureModel, Aligner, TdpModel).  It is compiled by oracle/Makefile together with the
// reference's own .cpp files *where they lie* (never copied) into oracle/_ref/libsietill_ref.so.
// Only tests/, __graft_entry__.smoke() and bench.py's reference/cpu_baseline legs may load it.
//
// The reference keeps its parameter tables private; the harness reads them through the
// usual `#define private public` test trick (layout is unchanged, so objects compiled without
// the define link fine).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <numeric>
#include <sstream>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>
#include <emmintrin.h>

#define private public
#define protected public
#include "Mixtures.hpp"
#include "Alignment.hpp"
#include "TdpModel.hpp"
#include "Training.hpp"
#include "Lexicon.hpp"
#include "Recognizer.hpp"
#include "SignalAnalysis.hpp"
#undef private
#undef protected

namespace {

struct RefModel {
  std::unique_ptr<Configuration> config;
  std::unique_ptr<MixtureModel>  mixtures;
  std::unique_ptr<TdpModel>      tdp;
};

struct Silencer {  // the reference prints a banner from its constructor (Mixtures.cpp:165,195)
  std::streambuf* old_out;
  std::streambuf* old_err;
  std::ostringstream sink;
  Silencer() : old_out(std::cout.rdbuf(sink.rdbuf())), old_err(std::cerr.rdbuf(sink.rdbuf())) {}
  ~Silencer() { std::cout.rdbuf(old_out); std::cerr.rdbuf(old_err); }
};

}  // namespace

extern "C" {

// config_json: path of a JSON file holding at least "verbosity":"noLog", the tdp-* keys and —
// to load a .mix — "action":"recognize" + "load-mixtures-from" (Mixtures.cpp:171-174).
// pooling: 0 none, 1 mixture, 2 global (SieTill.cpp:93-102).
void* ref_model_create(const char* config_json, uint64_t dim, uint64_t n_mixtures, int pooling,
                       int max_approx, uint16_t silence_state) {
  Silencer quiet;
  RefModel* m = new RefModel();
  m->config.reset(new Configuration(std::string(config_json)));
  MixtureModel::VarianceModel vm = pooling == 0 ? MixtureModel::NO_POOLING
                                 : pooling == 1 ? MixtureModel::MIXTURE_POOLING
                                                : MixtureModel::GLOBAL_POOLING;
  m->mixtures.reset(new MixtureModel(*m->config, dim, n_mixtures, vm, max_approx != 0));
  m->tdp.reset(new TdpModel(*m->config, silence_state));
  return m;
}

void ref_model_destroy(void* h) { delete static_cast<RefModel*>(h); }

// sizes: [0]=dim [1]=n_mixtures [2]=M (means ever created) [3]=V [4]=total live densities
void ref_model_sizes(void* h, uint64_t* out) {
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  out[0] = mm.dimension;
  out[1] = mm.mixtures_.size();
  out[2] = mm.mean_refs_.size();
  out[3] = mm.var_refs_.size();
  uint64_t total = 0;
  for (auto const& mix : mm.mixtures_) total += mix.size();
  out[4] = total;
}

// Copies every table of MixtureModel (Mixtures.hpp:72-87).  Any pointer may be NULL.
void ref_model_export(void* h, double* means, double* mean_acc, double* mean_w, double* mean_wlog,
                      double* mean_wacc, uint64_t* mean_refs, double* vars, double* vars_inv,
                      double* var_acc, double* var_wacc, uint64_t* var_refs, double* norm,
                      uint32_t* mix_offsets /*n_mix+1*/, uint32_t* mean_idx, uint32_t* var_idx) {
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  auto cp = [](double* dst, std::vector<double> const& src) {
    if (dst && !src.empty()) std::memcpy(dst, src.data(), src.size() * sizeof(double));
  };
  cp(means, mm.means_); cp(mean_acc, mm.mean_accumulators_); cp(mean_w, mm.mean_weights_);
  cp(mean_wlog, mm.mean_weights_log_); cp(mean_wacc, mm.mean_weight_accumulators_);
  cp(vars, mm.vars_); cp(vars_inv, mm.vars_inv_); cp(var_acc, mm.var_accumulators_);
  cp(var_wacc, mm.var_weight_accumulators_); cp(norm, mm.norm_);
  if (mean_refs) for (size_t i = 0; i < mm.mean_refs_.size(); i++) mean_refs[i] = mm.mean_refs_[i];
  if (var_refs)  for (size_t i = 0; i < mm.var_refs_.size();  i++) var_refs[i]  = mm.var_refs_[i];
  uint32_t pos = 0;
  for (size_t m = 0; m < mm.mixtures_.size(); m++) {
    if (mix_offsets) mix_offsets[m] = pos;
    for (size_t d = 0; d < mm.mixtures_[m].size(); d++, pos++) {
      if (mean_idx) mean_idx[pos] = mm.mixtures_[m][d].mean_idx;
      if (var_idx)  var_idx[pos]  = mm.mixtures_[m][d].var_idx;
    }
  }
  if (mix_offsets) mix_offsets[mm.mixtures_.size()] = pos;
}

int ref_model_write(void* h, const char* path) {
  std::ofstream out(path, std::ios_base::out | std::ios_base::trunc | std::ios_base::binary);
  if (!out.good()) return 1;
  static_cast<RefModel*>(h)->mixtures->write(out);
  return out.good() ? 0 : 1;
}

// E-step exactly as Trainer::train calls it (Training.cpp:114-118,200): `states[f]` is the aligned
// state of frame f; the harness wraps it in AlignmentItems with max_aligns = 1.
void ref_accumulate(void* h, const uint16_t* states, const float* x, uint64_t n_frames,
                    int first_pass, int max_approx) {
  Silencer quiet;
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  std::vector<AlignmentItem> ali(n_frames);
  for (uint64_t f = 0; f < n_frames; f++) ali[f] = AlignmentItem(1, states[f], 1.0f);
  mm.accumulate(ConstAlignmentIter(ali.data(), 1), ConstAlignmentIter(ali.data() + n_frames, 1),
                FeatureIter(x, mm.dimension), FeatureIter(x + n_frames * mm.dimension, mm.dimension),
                first_pass != 0, max_approx != 0);
}

void ref_finalize(void* h)                    { Silencer q; static_cast<RefModel*>(h)->mixtures->finalize(); }
void ref_split(void* h, uint64_t min_obs)     { Silencer q; static_cast<RefModel*>(h)->mixtures->split(min_obs); }
void ref_eliminate(void* h, double min_obs)   { Silencer q; static_cast<RefModel*>(h)->mixtures->eliminate(min_obs); }
uint64_t ref_num_densities(void* h)           { return static_cast<RefModel*>(h)->mixtures->num_densities(); }

// min_score for every (frame, listed state): scores[f*n_states+j], argmin likewise (Mixtures.cpp:696-713).
// x must have 2 floats of slack after the last frame (the reference over-reads, Mixtures.cpp:653).
void ref_min_score(void* h, const float* x, uint64_t n_frames, const uint16_t* states, uint64_t n_states,
                   double* scores, uint16_t* argmin) {
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  for (uint64_t f = 0; f < n_frames; f++) {
    FeatureIter it(x + f * mm.dimension, mm.dimension);
    for (uint64_t j = 0; j < n_states; j++) {
      std::pair<double, DensityIdx> r = mm.min_score(it, states[j]);
      scores[f * n_states + j] = r.first;
      if (argmin) argmin[f * n_states + j] = r.second;
    }
  }
}

// FeatureScorer::score (max-approx or sum, per the model's max_approx_; Mixtures.cpp:737-744)
void ref_score(void* h, const float* x, uint64_t n_frames, const uint16_t* states, uint64_t n_states,
               double* scores) {
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  for (uint64_t f = 0; f < n_frames; f++) {
    FeatureIter it(x + f * mm.dimension, mm.dimension);
    for (uint64_t j = 0; j < n_states; j++) scores[f * n_states + j] = mm.score(it, states[j]);
  }
}

// Aligner::align_sequence_full (Alignment.cpp:50-144) on one utterance.
double ref_align_full(void* h, const float* x, uint64_t T, const uint16_t* automaton, uint64_t S,
                      uint16_t* out_states) {
  RefModel* m = static_cast<RefModel*>(h);
  MarkovAutomaton ref;
  ref.states.assign(automaton, automaton + S);
  std::vector<AlignmentItem> ali(T);
  Aligner aligner(*m->mixtures, *m->tdp, 1);
  size_t d = m->mixtures->dimension;
  double cost = aligner.align_sequence_full(FeatureIter(x, d), FeatureIter(x + T * d, d), ref,
                                            AlignmentIter(ali.data(), 1), AlignmentIter(ali.data() + T, 1));
  for (uint64_t t = 0; t < T; t++) out_states[t] = ali[t].state;
  return cost;
}

double ref_align_pruned(void* h, const float* x, uint64_t T, const uint16_t* automaton, uint64_t S,
                        double threshold, uint16_t* out_states) {
  RefModel* m = static_cast<RefModel*>(h);
  MarkovAutomaton ref;
  ref.states.assign(automaton, automaton + S);
  std::vector<AlignmentItem> ali(T);
  Aligner aligner(*m->mixtures, *m->tdp, 1);
  size_t d = m->mixtures->dimension;
  double cost = aligner.align_sequence_pruned(FeatureIter(x, d), FeatureIter(x + T * d, d), ref,
                                              AlignmentIter(ali.data(), 1), AlignmentIter(ali.data() + T, 1),
                                              threshold);
  for (uint64_t t = 0; t < T; t++) out_states[t] = ali[t].state;
  return cost;
}

// Batched form used for CPU-baseline timing: the loop of Training.cpp:168-183 with alignment-pruning false.
// frame_off[u]..frame_off[u+1] are frame indices, aut_off likewise into `automata`.  Returns sum of costs.
double ref_align_batch(void* h, const float* x, uint64_t U, const uint64_t* frame_off,
                       const uint16_t* automata, const uint32_t* aut_off, uint16_t* out_states,
                       double* costs) {
  RefModel* m = static_cast<RefModel*>(h);
  Aligner aligner(*m->mixtures, *m->tdp, 1);
  size_t d = m->mixtures->dimension;
  double total = 0.0;
  std::vector<AlignmentItem> ali;
  for (uint64_t u = 0; u < U; u++) {
    uint64_t f0 = frame_off[u], f1 = frame_off[u + 1];
    MarkovAutomaton ref;
    ref.states.assign(automata + aut_off[u], automata + aut_off[u + 1]);
    ali.assign(f1 - f0, AlignmentItem());
    double c = aligner.align_sequence_full(FeatureIter(x + f0 * d, d), FeatureIter(x + f1 * d, d), ref,
                                           AlignmentIter(ali.data(), 1),
                                           AlignmentIter(ali.data() + (f1 - f0), 1));
    for (uint64_t t = 0; t < f1 - f0; t++) out_states[f0 + t] = ali[t].state;
    if (costs) costs[u] = c;
    total += c;
  }
  return total;
}

// Trainer::calc_am_score restated over a flat state array (Training.cpp:585-612): same call, same order.
double ref_am_score(void* h, const uint16_t* states, const float* x, uint64_t n_frames) {
  MixtureModel& mm = *static_cast<RefModel*>(h)->mixtures;
  double total = 0.0;
  for (uint64_t f = 0; f < n_frames; f++)
    total += mm.score(FeatureIter(x + f * mm.dimension, mm.dimension), states[f]);
  return total / n_frames;
}

double ref_tdp_score(void* h, uint16_t state, uint64_t jump) {
  return static_cast<RefModel*>(h)->tdp->score(state, jump);
}

// Recognizer::recognizeSequence_pruned (Recognizer.cpp:103-232) over the built-in sietill lexicon
// (Lexicon.cpp:70-85; 106 states).  am-threshold / word-penalty come from the config file.
// Returns the number of recognised words (written to out_words, at most max_words).
uint64_t ref_recognize(void* h, const float* x, uint64_t T, uint32_t* out_words, uint64_t max_words) {
  RefModel* m = static_cast<RefModel*>(h);
  Lexicon lexicon = build_sietill_lexicon();
  Recognizer rec(*m->config, lexicon, *m->mixtures, *m->tdp);
  std::vector<WordIdx> out;
  size_t d = m->mixtures->dimension;
  rec.recognizeSequence_pruned(FeatureIter(x, d), FeatureIter(x + T * d, d), out);
  for (size_t i = 0; i < out.size() && i < max_words; i++) out_words[i] = out[i];
  return out.size();
}

// Recognizer::editDistance (Recognizer.cpp:332-388): out = {total, substitutions, insertions, deletions}
void ref_edit_distance(void* h, const uint16_t* ref, uint64_t n_ref, const uint16_t* hyp, uint64_t n_hyp, uint64_t* out) {
  RefModel* m = static_cast<RefModel*>(h);
  Lexicon lexicon = build_sietill_lexicon();
  Recognizer rec(*m->config, lexicon, *m->mixtures, *m->tdp);
  std::vector<WordIdx> r(ref, ref + n_ref), hy(hyp, hyp + n_hyp);
  EDAccumulator e = rec.editDistance(r.begin(), r.end(), hy.begin(), hy.end());
  out[0] = e.total_count; out[1] = e.substitute_count; out[2] = e.insert_count; out[3] = e.delete_count;
}

// Trainer::linear_segmentation_approximation / linear_segmentation_running_sums (Training.cpp:350-546) on one utterance
void ref_linear_segmentation(void* h, int approx, const float* x, uint64_t T, const uint16_t* automaton, uint64_t S,
                             uint16_t* states, uint64_t* bounds) {
  RefModel* m = static_cast<RefModel*>(h);
  Lexicon lexicon = build_sietill_lexicon();
  Trainer trainer(*m->config, lexicon, *m->mixtures, *m->tdp, true);
  MarkovAutomaton aut;
  for (uint64_t i = 0; i < S; i++) aut.states.push_back(automaton[i]);
  std::vector<AlignmentItem> ali(T);
  size_t d = m->mixtures->dimension;
  FeatureIter fb(x, d), fe(x + T * d, d);
  AlignmentIter ab(&ali[0], 1), ae(&ali[0] + T, 1);
  std::pair<size_t, size_t> b = approx ? trainer.linear_segmentation_approximation(aut, fb, fe, ab, ae)
                                       : trainer.linear_segmentation_running_sums(aut, fb, fe, ab, ae);
  for (uint64_t t = 0; t < T; t++) states[t] = ali[t].state;
  bounds[0] = b.first; bounds[1] = b.second;
}


// SignalAnalysis::process_features (SignalAnalysis.cpp:379-399) on one utterance.  config_json carries
// "n-features-file" / "n-features-first" / "n-features-second" / "deriv-step" / "energy-max-norm"; mean / stddev
// (may be NULL) are what read_normalization_file (:371-375) would have loaded.  Returns n_features_total.
uint64_t ref_process_features(const char* config_json, const double* mean, const double* stddev, const float* raw,
                              uint64_t T, float* out) {
  Silencer quiet;
  Configuration config{std::string(config_json)};
  SignalAnalysis sa(config);
  if (mean) {
    sa.mean_.assign(mean, mean + sa.n_features_total);
    sa.stddev_.assign(stddev, stddev + sa.n_features_total);
    sa.apply_mean_var_normalization_ = true;
  }
  std::vector<float> f(raw, raw + T * sa.n_features_in_file);
  sa.process_features(f);
  std::copy(f.begin(), f.end(), out);
  return sa.n_features_total;
}

}  // extern "C"
