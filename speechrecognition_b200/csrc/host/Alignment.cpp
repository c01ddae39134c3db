// Aligner: the reference's per-utterance signatures (src/sietill/Alignment.hpp:46-55) plus the batched
// entry the training driver uses; both run K1 (gathered scoring) + K2 (Viterbi + backtrace) on the GPU.
#include <istream>
#include <limits>
#include <ostream>

#include "sietill_host.hpp"

namespace sietill_b200 {

Aligner::Aligner(MixtureModel const& mixtures, TdpModel const& tdp_model, size_t max_aligns)
    : mixtures_(mixtures), tdp_model_(tdp_model), max_aligns_(max_aligns) {
  // the reference stores num_max_aligns items per frame but only ever fills slot 0 (Alignment.cpp:131-133)
  if (max_aligns != 1) throw std::runtime_error("Aligner: num-max-aligns must be 1");
}

namespace {
void upload_corpus(Device& dev, std::vector<size_t> const& frame_off, std::vector<MarkovAutomaton const*> const& automata) {
  const size_t U = automata.size();
  std::vector<uint64_t> fo(frame_off.begin(), frame_off.end());
  std::vector<uint32_t> aut, ao(U + 1, 0);
  for (size_t u = 0; u < U; u++) {
    aut.insert(aut.end(), automata[u]->states.begin(), automata[u]->states.end());
    ao[u + 1] = (uint32_t)aut.size();
  }
  dev.check(sgmm_set_corpus(dev.ctx(), (uint32_t)U, fo.data(), aut.data(), ao.data()), "sgmm_set_corpus");
}
}  // namespace

double Aligner::align_one(FeatureIter fb, FeatureIter fe, MarkovAutomaton const& ref, AlignmentIter ab, AlignmentIter ae,
                          int convention, double beam) {
  const size_t T = (size_t)(fe - fb);
  if (ab.size != 1 || (size_t)(ae - ab) != T) throw std::runtime_error("Aligner: alignment range does not match the features");
  if (fb.size != mixtures_.dimension) throw std::runtime_error("Aligner: feature dimension differs from the model's");
  if (T == 0 || ref.num_states() == 0) throw std::runtime_error("Aligner: empty sequence or automaton");
  Device& dev = mixtures_.scratch();
  dev.check(sgmm_set_features(dev.ctx(), *fb, T, (uint32_t)mixtures_.dimension), "sgmm_set_features");
  upload_corpus(dev, {0, T}, {&ref});
  dev.check(sgmm_set_tdp(dev.ctx(), tdp_model_.loop(), tdp_model_.forward(), tdp_model_.skip(), tdp_model_.silence_state),
            "sgmm_set_tdp");
  dev.check(sgmm_set_beam(dev.ctx(), beam), "sgmm_set_beam");
  double cost = 0.0;
  dev.check(sgmm_align(dev.ctx(), convention, reinterpret_cast<sgmm_alignment_item*>(*ab), &cost), "sgmm_align");
  if (*ab == mixtures_.resident_ali_) mixtures_.resident_ali_ = nullptr;   // host copy changed under the token
  return cost;
}

double Aligner::align_sequence_full(FeatureIter feature_begin, FeatureIter feature_end, MarkovAutomaton const& reference,
                                    AlignmentIter align_begin, AlignmentIter align_end) {
  return align_one(feature_begin, feature_end, reference, align_begin, align_end, SGMM_TDP_SOURCE,
                   std::numeric_limits<double>::infinity());
}

double Aligner::align_sequence_pruned(FeatureIter feature_begin, FeatureIter feature_end, MarkovAutomaton const& reference,
                                      AlignmentIter align_begin, AlignmentIter align_end, double pruning_threshold) {
  return align_one(feature_begin, feature_end, reference, align_begin, align_end, SGMM_TDP_DESTINATION, pruning_threshold);
}

void Aligner::align_corpus(Corpus const& corpus, std::vector<MarkovAutomaton> const& automata, Alignment& alignment,
                           bool destination_tdp, std::vector<double>* costs, double pruning_threshold) {
  const size_t U = corpus.get_corpus_size();
  if (automata.size() != U) throw std::runtime_error("Aligner: one automaton per segment expected");
  mixtures_.sync_device();
  std::pair<FeatureIter, FeatureIter> all = corpus.get_all_features();
  mixtures_.set_corpus_features(all.first, all.second);
  Device& dev = mixtures_.device();
  // the resident corpus layout is keyed on its CONTENT (frame offsets + automata), not on the Corpus address: changed
  // automata or a recycled address must not reuse stale tables
  uint64_t key = 0xCBF29CE484222325ull ^ U;
  for (size_t v : corpus.frame_offsets()) key = (key ^ (uint64_t)v) * 0x100000001B3ull;
  for (size_t u = 0; u < U; u++) {
    key = (key ^ (uint64_t)automata[u].states.size()) * 0x100000001B3ull;
    for (StateIdx st : automata[u].states) key = (key ^ (uint64_t)st) * 0x100000001B3ull;
  }
  if (!corpus_key_valid_ || corpus_key_ != key) {
    std::vector<MarkovAutomaton const*> ptrs(U);
    for (size_t u = 0; u < U; u++) ptrs[u] = &automata[u];
    upload_corpus(dev, corpus.frame_offsets(), ptrs);
    corpus_key_ = key;
    corpus_key_valid_ = true;
  }
  dev.check(sgmm_set_tdp(dev.ctx(), tdp_model_.loop(), tdp_model_.forward(), tdp_model_.skip(), tdp_model_.silence_state),
            "sgmm_set_tdp");
  dev.check(sgmm_set_beam(dev.ctx(), destination_tdp ? pruning_threshold : std::numeric_limits<double>::infinity()),
            "sgmm_set_beam");
  alignment.resize(corpus.get_total_frame_count());
  if (costs) costs->resize(U);
  dev.check(sgmm_align(dev.ctx(), destination_tdp ? SGMM_TDP_DESTINATION : SGMM_TDP_SOURCE,
                       reinterpret_cast<sgmm_alignment_item*>(alignment.data()), costs ? costs->data() : nullptr),
            "sgmm_align");
  // the alignment just written is also the one resident on the device: the next accumulate() on this very
  // range skips the upload and reuses the scores of this pass (same model, same cells)
  mixtures_.resident_ali_ = alignment.data();
  mixtures_.resident_ali_frames_ = alignment.size();
  mixtures_.resident_ali_sum_ = alignment_checksum(alignment.data(), alignment.size());
}

// Alignment.cpp:303-317: u64 max_aligns, u64 frames, then the raw 8-byte items
void write_alignment(std::ostream& out, Alignment const& alignment, size_t max_aligns) {
  const uint64_t ma = max_aligns, n = alignment.size() / max_aligns;
  out.write(reinterpret_cast<const char*>(&ma), sizeof(ma));
  out.write(reinterpret_cast<const char*>(&n), sizeof(n));
  out.write(reinterpret_cast<const char*>(alignment.data()), (std::streamsize)(alignment.size() * sizeof(AlignmentItem)));
}
void read_alignment(std::istream& in, Alignment& alignment, size_t& max_aligns) {
  uint64_t ma = 0, n = 0;
  in.read(reinterpret_cast<char*>(&ma), sizeof(ma));
  in.read(reinterpret_cast<char*>(&n), sizeof(n));
  if (!in.good()) throw std::runtime_error("read_alignment: truncated header");
  max_aligns = ma;
  alignment.resize(ma * n);
  in.read(reinterpret_cast<char*>(alignment.data()), (std::streamsize)(alignment.size() * sizeof(AlignmentItem)));
  if ((size_t)in.gcount() != alignment.size() * sizeof(AlignmentItem)) throw std::runtime_error("read_alignment: truncated body");
}

}  // namespace sietill_b200
