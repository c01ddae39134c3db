// Configuration, MarkovAutomaton, Lexicon, Corpus view, TdpModel, utterance sharding — the small host
// types either side of the hot path (reference: Config.hpp, MarkovAutomaton.hpp, Lexicon.cpp, Corpus.cpp,
// TdpModel.cpp under src/sietill).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <sstream>

#include "sietill_host.hpp"

namespace sietill_b200 {

// ---- Configuration ------------------------------------------------------------------------------------
Configuration& Configuration::set(std::string const& key, double value) {
  std::ostringstream ss;
  ss.precision(17);
  ss << value;
  values_[key] = ss.str();
  return *this;
}
std::string Configuration::get_string(std::string const& key, std::string const& dflt) const {
  auto it = values_.find(key);
  return it == values_.end() ? dflt : it->second;
}
double Configuration::get_double(std::string const& key, double dflt) const {
  auto it = values_.find(key);
  if (it == values_.end()) return dflt;
  char* end = nullptr;
  const double v = std::strtod(it->second.c_str(), &end);
  if (end == it->second.c_str() || *end != 0) throw std::runtime_error("Configuration: " + key + " is not a number");
  return v;
}
unsigned Configuration::get_uint(std::string const& key, unsigned dflt) const {
  auto it = values_.find(key);
  if (it == values_.end()) return dflt;
  char* end = nullptr;
  const unsigned long v = std::strtoul(it->second.c_str(), &end, 10);
  if (end == it->second.c_str() || *end != 0) throw std::runtime_error("Configuration: " + key + " is not an unsigned integer");
  return (unsigned)v;
}
bool Configuration::get_bool(std::string const& key, bool dflt) const {
  auto it = values_.find(key);
  if (it == values_.end()) return dflt;
  if (it->second == "true" || it->second == "1") return true;
  if (it->second == "false" || it->second == "0") return false;
  throw std::runtime_error("Configuration: " + key + " is not a boolean");
}

// ---- MarkovAutomaton (MarkovAutomaton.hpp:22-66) ----------------------------------------------------------
MarkovAutomaton::MarkovAutomaton(StateIdx start, uint16_t num, uint16_t repetitions) {
  states.reserve((size_t)num * repetitions);
  for (StateIdx s = start; s < start + num; s++) states.insert(states.end(), repetitions, s);
}
StateIdx MarkovAutomaton::max_state() const {
  StateIdx m = std::numeric_limits<StateIdx>::min();
  for (StateIdx s : states) m = std::max(m, s);
  return m;
}
MarkovAutomaton MarkovAutomaton::concat(std::vector<MarkovAutomaton const*> automata) {
  MarkovAutomaton r;
  for (MarkovAutomaton const* a : automata) r.states.insert(r.states.end(), a->states.begin(), a->states.end());
  return r;
}

// ---- Lexicon (Lexicon.cpp:11-85): linear whole-word HMMs with contiguous state ids ----------------------------
WordIdx Lexicon::add_word(std::string const& orth, uint16_t num_states, uint16_t state_repetitions, bool silence) {
  const WordIdx w = automata_.size();
  orth_.push_back(orth);
  automata_.push_back(MarkovAutomaton(next_state_, num_states, state_repetitions));
  next_state_ = (StateIdx)(next_state_ + num_states);
  if (silence) silence_ = w;
  return w;
}
WordIdx Lexicon::operator[](std::string const& orth) const {
  for (WordIdx w = 0; w < orth_.size(); w++)
    if (orth_[w] == orth) return w;
  throw std::runtime_error("Lexicon: unknown word " + orth);
}
Lexicon build_sietill_lexicon() {
  // the German digits task: one silence state, eleven whole-word models of 9 or 12 states, each state
  // twice in the automaton (Lexicon.cpp:70-85) -> 106 states
  struct Entry { const char* orth; uint16_t states; };
  const Entry words[] = {{"eins", 9}, {"zwei", 9}, {"drei", 9}, {"vier", 9}, {"fuenf", 12}, {"sechs", 9},
                         {"sieben", 12}, {"acht", 9}, {"neun", 9}, {"null", 9}, {"zwo", 9}};
  Lexicon lex;
  lex.add_word("[silence]", 1, 1, true);
  for (Entry const& e : words) lex.add_word(e.orth, e.states, 2);
  return lex;
}

// ---- Corpus view --------------------------------------------------------------------------------------
Corpus::Corpus(const float* features, size_t dim, std::vector<size_t> frame_off, std::vector<WordIdx> words,
               std::vector<size_t> word_off)
    : features_(features), dim_(dim), frame_off_(std::move(frame_off)), words_(std::move(words)),
      word_off_(std::move(word_off)) {
  if (frame_off_.empty() || frame_off_[0] != 0) throw std::runtime_error("Corpus: frame offsets must start at 0");
  if (word_off_.size() != frame_off_.size()) throw std::runtime_error("Corpus: word and frame offset tables differ in length");
}
size_t Corpus::get_max_seq_length() const {
  size_t m = 0;
  for (size_t s = 0; s + 1 < frame_off_.size(); s++) m = std::max(m, frame_off_[s + 1] - frame_off_[s]);
  return m;
}
std::pair<WordIter, WordIter> Corpus::get_word_sequence(SegmentIdx s) const {
  return std::make_pair(words_.begin() + (ptrdiff_t)word_off_.at(s), words_.begin() + (ptrdiff_t)word_off_.at(s + 1));
}
std::pair<FeatureIter, FeatureIter> Corpus::get_feature_sequence(SegmentIdx s) const {
  return std::make_pair(FeatureIter(features_ + frame_off_.at(s) * dim_, dim_),
                        FeatureIter(features_ + frame_off_.at(s + 1) * dim_, dim_));
}
std::pair<FeatureIter, FeatureIter> Corpus::get_all_features() const {
  return std::make_pair(FeatureIter(features_, dim_), FeatureIter(features_ + frame_off_.back() * dim_, dim_));
}
std::pair<size_t, size_t> Corpus::get_feature_offsets(SegmentIdx s) const {
  return std::make_pair(frame_off_.at(s) * dim_, frame_off_.at(s + 1) * dim_);
}

// ---- TdpModel (TdpModel.cpp:9-29) -----------------------------------------------------------------------
TdpModel::TdpModel(Configuration const& config, StateIdx silence_state)
    : silence_state(silence_state), tdp_loop_(config.get_double("tdp-loop", 0.0)),
      tdp_forward_(config.get_double("tdp-forward", 0.0)), tdp_skip_(config.get_double("tdp-skip", 0.0)) {}

double TdpModel::score(StateIdx to, size_t jump) const {
  if (to == silence_state) return tdp_forward_;   // silence pays the forward penalty whatever the jump
  switch (jump) {
    case 0: return tdp_loop_;
    case 1: return tdp_forward_;
    case 2: return tdp_skip_;
    default: return std::numeric_limits<double>::infinity();
  }
}

// ---- sharding --------------------------------------------------------------------------
// bounds[r] = first utterance whose start frame is >= r * total / world (cut points rounded to utterance
// boundaries); contiguous, deterministic, every utterance in exactly one shard.
std::vector<size_t> shard_utterances(std::vector<size_t> const& frame_off, int world_size) {
  if (world_size < 1) throw std::runtime_error("shard_utterances: world_size must be >= 1");
  const size_t U = frame_off.size() - 1, total = frame_off.back();
  std::vector<size_t> bounds((size_t)world_size + 1, U);
  bounds[0] = 0;
  for (int r = 1; r < world_size; r++) {
    const double target = (double)total * r / world_size;
    size_t u = (size_t)(std::lower_bound(frame_off.begin(), frame_off.end(), (size_t)std::ceil(target)) - frame_off.begin());
    // choose the closer of the two utterance boundaries around the target
    if (u > 0 && u <= U && target - (double)frame_off[u - 1] < (double)frame_off[std::min(u, U)] - target) u--;
    bounds[r] = std::max(bounds[r - 1], std::min(u, U));
  }
  return bounds;
}

// ---- SignalAnalysis (load-time feature processing, SignalAnalysis.cpp:46-56,371-399) ---------------------------
SignalAnalysis::SignalAnalysis(Configuration const& config)
    : energy_max_norm(config.get_bool("energy-max-norm", true)),
      n_features_in_file(config.get_uint("n-features-file", 12u)),
      n_features_first(config.get_uint("n-features-first", 12u)),
      n_features_second(config.get_uint("n-features-second", 1u)),
      n_features_total(n_features_in_file + n_features_first + n_features_second),
      deriv_step(config.get_uint("deriv-step", 3u)) {}

void SignalAnalysis::read_normalization_file(std::istream& in) {
  std::vector<double> mean(n_features_total), stddev(n_features_total);
  in.read(reinterpret_cast<char*>(mean.data()), (std::streamsize)(sizeof(double) * mean.size()));
  in.read(reinterpret_cast<char*>(stddev.data()), (std::streamsize)(sizeof(double) * stddev.size()));
  if (!in) throw std::runtime_error("SignalAnalysis: normalization file too short");
  set_normalization(mean, stddev);
}

void SignalAnalysis::set_normalization(std::vector<double> const& mean, std::vector<double> const& stddev) {
  if (mean.size() != n_features_total || stddev.size() != n_features_total)
    throw std::runtime_error("SignalAnalysis: normalization vectors must have n_features_total entries");
  mean_ = mean;
  stddev_ = stddev;
  apply_mean_var_normalization_ = true;
}

void SignalAnalysis::process_features(Device& dev, const float* raw, size_t n_frames, std::vector<size_t> const& frame_off,
                                      float* out) const {
  if (frame_off.size() < 2) throw std::runtime_error("SignalAnalysis: no segments");
  std::vector<uint64_t> off(frame_off.begin(), frame_off.end());
  dev.check(sgmm_process_features(dev.ctx(), raw, n_frames, (uint32_t)n_features_in_file, (uint32_t)n_features_first,
                                  (uint32_t)n_features_second, (uint32_t)deriv_step, (uint32_t)(off.size() - 1), off.data(),
                                  apply_mean_var_normalization_ ? mean_.data() : nullptr,
                                  apply_mean_var_normalization_ ? stddev_.data() : nullptr, energy_max_norm ? 1 : 0, out),
            "sgmm_process_features");
}

}  // namespace sietill_b200
