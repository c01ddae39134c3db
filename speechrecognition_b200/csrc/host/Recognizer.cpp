// Host mirror of the reference's Recognizer (Recognizer.hpp:53-130, Recognizer.cpp:37-232,332-388) over the C-ABI:
// the word-loop search itself runs on the GPU (sgmm_recognize, k_recognize.cu), batched over segments.
#include <chrono>
#include <stdexcept>

#include "sietill_host.hpp"

namespace sietill_b200 {

EDAccumulator& EDAccumulator::operator+=(EDAccumulator const& o) {
  total_count += o.total_count; substitute_count += o.substitute_count;
  insert_count += o.insert_count; delete_count += o.delete_count;
  return *this;
}

// Recognizer.cpp:332-388.  Levenshtein alignment that prefers, on equal cost: match, substitution, insertion, deletion
// (each step taken only on strict improvement, :356-381).  Two rows are swapped per hypothesis word and only column 0
// of the NEW current row is incremented (:346-349), so column 0 holds ceil(h / 2) instead of h — kept, because it
// decides how leading insertions are counted in the reference's WER.
EDAccumulator editDistance(WordIter ref_begin, WordIter ref_end, WordIter hyp_begin, WordIter hyp_end) {
  const size_t R = (size_t)(ref_end - ref_begin), H = (size_t)(hyp_end - hyp_begin);
  struct Cell { uint16_t total = 0, sub = 0, ins = 0, del = 0; };   // 16-bit counters as in Recognizer.hpp:18-22
  std::vector<Cell> row_a(R + 1), row_b(R + 1);
  std::vector<Cell>* cur = &row_a;
  std::vector<Cell>* prev = &row_b;
  for (size_t r = 1; r <= R; r++) { (*cur)[r] = (*cur)[r - 1]; (*cur)[r].total++; (*cur)[r].del++; }
  for (size_t h = 1; h <= H; h++) {
    std::swap(cur, prev);
    (*cur)[0].total++; (*cur)[0].ins++;
    for (size_t r = 1; r <= R; r++) {
      uint16_t best = 0xFFFF;
      Cell pick;
      const Cell& diag = (*prev)[r - 1];
      if (diag.total < best && ref_begin[r - 1] == hyp_begin[h - 1]) { pick = diag; best = pick.total; }
      if (diag.total + 1 < best) { pick = diag; pick.total++; pick.sub++; best = pick.total; }
      const Cell& up = (*prev)[r];
      if (up.total + 1 < best) { pick = up; pick.total++; pick.ins++; best = pick.total; }
      const Cell& left = (*cur)[r - 1];
      if (left.total + 1 < best) { pick = left; pick.total++; pick.del++; best = pick.total; }
      if (best != 0xFFFF) (*cur)[r] = pick;     // (an untouched cell keeps its old content, as in the reference)
    }
  }
  EDAccumulator out;
  out.total_count = (*cur)[R].total; out.substitute_count = (*cur)[R].sub;
  out.insert_count = (*cur)[R].ins; out.delete_count = (*cur)[R].del;
  return out;
}

Recognizer::Recognizer(Configuration const& config, Lexicon const& lexicon, MixtureModel const& scorer,
                       TdpModel const& tdp_model)
    : am_threshold_(config.get_double("am-threshold", 20.0)),
      word_penalty_(config.get_double("word-penalty", 10.0)),
      pruned_search_(config.get_bool("pruned-search", true)),
      max_recognition_runs_(config.get_uint("max-recognition-runs", 1000)),
      lexicon_(lexicon), scorer_(scorer), tdp_model_(tdp_model) {
  if (!pruned_search_)
    throw std::runtime_error("Recognizer: pruned-search=false selects recognizeSequence, a stub in the reference "
                             "(Recognizer.cpp:234-240); only the pruned search exists");
}

void Recognizer::upload_lexicon(Device& dev) const {
  std::vector<uint32_t> off(1, 0), states;
  for (WordIdx w = 0; w < lexicon_.num_words(); w++) {
    MarkovAutomaton const& a = lexicon_.get_automaton_for_word(w);
    for (size_t i = 0; i < a.num_states(); i++) states.push_back(a[i]);
    off.push_back((uint32_t)states.size());
  }
  dev.check(sgmm_set_lexicon(dev.ctx(), (uint32_t)lexicon_.num_words(), off.data(), states.data(),
                             (uint32_t)lexicon_.silence_idx()), "sgmm_set_lexicon");
  dev.check(sgmm_set_tdp(dev.ctx(), tdp_model_.loop(), tdp_model_.forward(), tdp_model_.skip(), tdp_model_.silence_state),
            "sgmm_set_tdp");
}

void Recognizer::recognizeSequence_pruned(FeatureIter feature_begin, FeatureIter feature_end, std::vector<WordIdx>& output) {
  const size_t T = (size_t)(feature_end - feature_begin);
  output.clear();
  if (T == 0) return;
  if (feature_begin.size != scorer_.dimension) throw std::runtime_error("Recognizer: feature dimension differs from the model's");
  Device& dev = scorer_.scratch();
  dev.check(sgmm_set_features(dev.ctx(), *feature_begin, T, (uint32_t)scorer_.dimension), "sgmm_set_features");
  upload_lexicon(dev);
  const uint64_t fo[2] = {0, T};
  std::vector<uint32_t> words(T + 1);
  uint32_t off[2] = {0, 0};
  dev.check(sgmm_recognize(dev.ctx(), 1, fo, word_penalty_, am_threshold_, words.data(), words.size(), off, nullptr),
            "sgmm_recognize");
  for (uint32_t i = 0; i < off[1]; i++) output.push_back((WordIdx)words[i]);
}

void Recognizer::recognize(Corpus const& corpus) {
  const auto t0 = std::chrono::steady_clock::now();
  const size_t n = std::min(corpus.get_corpus_size(), max_recognition_runs_);
  hypotheses.assign(n, std::vector<WordIdx>());
  errors = EDAccumulator(); ref_total = 0; sentence_errors = 0; wer = ser = 0.0;
  if (n == 0) return;
  // the first n segments are one contiguous block of the corpus' feature array (Corpus.cpp:107-110)
  std::vector<size_t> const& fo = corpus.frame_offsets();
  std::pair<FeatureIter, FeatureIter> all = corpus.get_all_features();
  FeatureIter begin = all.first, end = all.first;
  end += (ptrdiff_t)fo[n];
  scorer_.sync_device();
  scorer_.set_corpus_features(begin, end);
  Device& dev = scorer_.device();
  upload_lexicon(dev);
  std::vector<uint64_t> fo64(fo.begin(), fo.begin() + n + 1);
  std::vector<uint32_t> words(fo[n] + 1), off(n + 1);
  dev.check(sgmm_recognize(dev.ctx(), (uint32_t)n, fo64.data(), word_penalty_, am_threshold_, words.data(), words.size(),
                           off.data(), nullptr), "sgmm_recognize");
  for (size_t s = 0; s < n; s++) {
    for (uint32_t i = off[s]; i < off[s + 1]; i++) hypotheses[s].push_back((WordIdx)words[i]);
    std::pair<WordIter, WordIter> ref = corpus.get_word_sequence((SegmentIdx)s);
    EDAccumulator ed = editDistance(ref.first, ref.second, hypotheses[s].cbegin(), hypotheses[s].cend());
    errors += ed;
    ref_total += (size_t)(ref.second - ref.first);
    if (ed.total_count > 0) sentence_errors++;
  }
  wer = ref_total ? 100.0 * (double)errors.total_count / (double)ref_total : 0.0;
  ser = 100.0 * (double)sentence_errors / (double)n;
  seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace sietill_b200
