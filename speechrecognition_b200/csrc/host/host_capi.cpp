// Flat C shim over the C++ host classes so that the parity tests (Python, ctypes) can drive MixtureModel /
// Aligner / Trainer exactly as a C++ caller would.  Not part of the drop-in boundary (that is
// include/sietill_b200.h); table order matches the checkers' export format so the tests can diff directly.
#include <cstring>
#include <fstream>

#include "sietill_host.hpp"

using namespace sietill_b200;

namespace {
thread_local std::string g_err;
int g_device_mstep = 1;   // default of the "device-mstep" key for models built through this shim (tests run both)

struct HostModel {
  Configuration config;
  std::unique_ptr<MixtureModel> model;
};

MixtureModel::VarianceModel pooling_of(int p) {   // 0 none, 1 mixture, 2 global (the tests' numbering)
  return p == 0 ? MixtureModel::NO_POOLING : p == 1 ? MixtureModel::MIXTURE_POOLING : MixtureModel::GLOBAL_POOLING;
}

template <typename F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (std::exception const& e) {
    g_err = e.what();
    return 1;
  }
}

template <typename T>
void copy_out(T* dst, std::vector<T> const& src) {
  if (dst && !src.empty()) memcpy(dst, src.data(), src.size() * sizeof(T));
}
}  // namespace

extern "C" {

const char* sthost_last_error() { return g_err.c_str(); }
void sthost_set_device_mstep(int on) { g_device_mstep = on; }

void* sthost_model_create(uint32_t dim, uint32_t n_mix, int pooling, int max_approx, int host_only, int device,
                          const char* load_from) {
  HostModel* h = new HostModel();
  const int rc = guarded([&] {
    h->config.set("host-only", host_only != 0).set("device", (unsigned)device).set("device-mstep", g_device_mstep != 0);
    if (load_from && *load_from) h->config.set("load-mixtures-from", std::string(load_from)).set("action", std::string("recognize"));
    h->model.reset(new MixtureModel(h->config, dim, n_mix, pooling_of(pooling), max_approx != 0));
  });
  if (rc) { delete h; return nullptr; }
  return h;
}
void sthost_model_destroy(void* h) { delete static_cast<HostModel*>(h); }

int sthost_model_sizes(void* h, uint64_t* out) {
  return guarded([&] {
    MixtureModel::Tables t = static_cast<HostModel*>(h)->model->export_tables();
    out[0] = static_cast<HostModel*>(h)->model->dimension;
    out[1] = t.mix_offsets.size() - 1; out[2] = t.mean_w.size(); out[3] = t.var_wacc.size(); out[4] = t.mean_idx.size();
  });
}

int sthost_model_export(void* h, double* means, double* mean_acc, double* mean_w, double* mean_wlog, double* mean_wacc,
                        uint64_t* mean_refs, double* vars, double* vars_inv, double* var_acc, double* var_wacc,
                        uint64_t* var_refs, double* norm, uint32_t* mix_offsets, uint32_t* mean_idx, uint32_t* var_idx) {
  return guarded([&] {
    MixtureModel::Tables t = static_cast<HostModel*>(h)->model->export_tables();
    copy_out(means, t.means); copy_out(mean_acc, t.mean_acc); copy_out(mean_w, t.mean_w); copy_out(mean_wlog, t.mean_wlog);
    copy_out(mean_wacc, t.mean_wacc); copy_out(mean_refs, t.mean_refs); copy_out(vars, t.vars); copy_out(vars_inv, t.vars_inv);
    copy_out(var_acc, t.var_acc); copy_out(var_wacc, t.var_wacc); copy_out(var_refs, t.var_refs); copy_out(norm, t.norm);
    copy_out(mix_offsets, t.mix_offsets); copy_out(mean_idx, t.mean_idx); copy_out(var_idx, t.var_idx);
  });
}

int sthost_model_import(void* h, uint32_t M, uint32_t V, const double* means, const double* mean_acc, const double* mean_w,
                        const double* mean_wlog, const double* mean_wacc, const uint64_t* mean_refs, const double* vars,
                        const double* vars_inv, const double* var_acc, const double* var_wacc, const uint64_t* var_refs,
                        const double* norm, const uint32_t* mix_offsets, const uint32_t* mean_idx, const uint32_t* var_idx) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    const size_t d = m.dimension, N = m.num_mixtures();
    MixtureModel::Tables t;
    t.means.assign(means, means + M * d); t.mean_acc.assign(mean_acc, mean_acc + M * d);
    t.mean_w.assign(mean_w, mean_w + M); t.mean_wlog.assign(mean_wlog, mean_wlog + M);
    t.mean_wacc.assign(mean_wacc, mean_wacc + M); t.mean_refs.assign(mean_refs, mean_refs + M);
    t.vars.assign(vars, vars + V * d); t.vars_inv.assign(vars_inv, vars_inv + V * d);
    t.var_acc.assign(var_acc, var_acc + V * d); t.var_wacc.assign(var_wacc, var_wacc + V);
    t.var_refs.assign(var_refs, var_refs + V); t.norm.assign(norm, norm + V);
    t.mix_offsets.assign(mix_offsets, mix_offsets + N + 1);
    t.mean_idx.assign(mean_idx, mean_idx + mix_offsets[N]); t.var_idx.assign(var_idx, var_idx + mix_offsets[N]);
    m.import_tables(t);
  });
}

int sthost_model_finalize(void* h) { return guarded([&] { static_cast<HostModel*>(h)->model->finalize(); }); }
int sthost_model_split(void* h, uint64_t min_obs) { return guarded([&] { static_cast<HostModel*>(h)->model->split(min_obs); }); }
int sthost_model_eliminate(void* h, double min_obs) { return guarded([&] { static_cast<HostModel*>(h)->model->eliminate(min_obs); }); }
uint64_t sthost_model_num_densities(void* h) { return static_cast<HostModel*>(h)->model->num_densities(); }

int sthost_model_write(void* h, const char* path) {
  return guarded([&] {
    std::ofstream out(path, std::ios::binary | std::ios::trunc);
    if (!out.good()) throw std::runtime_error(std::string("cannot open ") + path);
    static_cast<HostModel*>(h)->model->write(out);
  });
}
int sthost_model_read(void* h, const char* path) {
  return guarded([&] {
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) throw std::runtime_error(std::string("cannot open ") + path);
    static_cast<HostModel*>(h)->model->read(in);
  });
}

// MixtureModel::accumulate over a whole corpus
int sthost_model_accumulate(void* h, const sgmm_alignment_item* ali, const float* x, uint64_t n_frames, int first_pass,
                            double* score_sum) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    const AlignmentItem* a = reinterpret_cast<const AlignmentItem*>(ali);
    m.accumulate(ConstAlignmentIter(a, 1), ConstAlignmentIter(a + n_frames, 1), FeatureIter(x, m.dimension),
                 FeatureIter(x + n_frames * m.dimension, m.dimension), first_pass != 0, true);
    if (score_sum) *score_sum = m.last_score_sum();
  });
}

// MixtureModel::accumulate(..., first_pass = false, max_approx = false): the posterior-weighted E-step
int sthost_model_accumulate_soft(void* h, const sgmm_alignment_item* ali, const float* x, uint64_t n_frames,
                                 double* score_sum) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    const AlignmentItem* a = reinterpret_cast<const AlignmentItem*>(ali);
    m.accumulate(ConstAlignmentIter(a, 1), ConstAlignmentIter(a + n_frames, 1), FeatureIter(x, m.dimension),
                 FeatureIter(x + n_frames * m.dimension, m.dimension), false, false);
    if (score_sum) *score_sum = m.last_score_sum();
  });
}

// FeatureScorer::score per (frame, state) — sum_score for a model built with max_approx == false
int sthost_model_scores(void* h, const float* x, uint64_t T, int prepared, double* scores) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    const size_t N = m.num_mixtures();
    FeatureIter b(x, m.dimension), e(x + T * m.dimension, m.dimension);
    if (prepared) m.prepare_sequence(b, e);
    for (uint64_t t = 0; t < T; t++)
      for (size_t s = 0; s < N; s++) scores[t * N + s] = m.score(b + (ptrdiff_t)t, (StateIdx)s);
  });
}

// FeatureScorer idiom: prepare_sequence, then min_score per (frame, state); prepared == 0 skips the
// prepare call so that every lookup goes through the one-frame path
int sthost_model_score_sequence(void* h, const float* x, uint64_t T, int prepared, double* scores, uint32_t* argmin) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    const size_t N = m.num_mixtures();
    FeatureIter b(x, m.dimension), e(x + T * m.dimension, m.dimension);
    if (prepared) m.prepare_sequence(b, e);
    for (uint64_t t = 0; t < T; t++)
      for (size_t s = 0; s < N; s++) {
        std::pair<double, DensityIdx> r = m.min_score(b + (ptrdiff_t)t, (StateIdx)s);
        scores[t * N + s] = r.first;
        argmin[t * N + s] = r.second;
        if (m.score(b + (ptrdiff_t)t, (StateIdx)s) != r.first) throw std::runtime_error("score() != min_score().first");
      }
  });
}

// Aligner::align_sequence_full (pruned == 0) / align_sequence_pruned with the given threshold
int sthost_align_sequence(void* h, double loop, double forward, double skip, uint32_t silence, const float* x, uint64_t T,
                          const uint16_t* automaton, uint64_t S, int pruned, double threshold,
                          sgmm_alignment_item* out, double* cost) {
  return guarded([&] {
    MixtureModel& m = *static_cast<HostModel*>(h)->model;
    Configuration cfg;
    cfg.set("tdp-loop", loop).set("tdp-forward", forward).set("tdp-skip", skip);
    TdpModel tdp(cfg, (StateIdx)silence);
    Aligner aligner(m, tdp, 1);
    MarkovAutomaton ref;
    ref.states.assign(automaton, automaton + S);
    AlignmentItem* a = reinterpret_cast<AlignmentItem*>(out);
    FeatureIter fb(x, m.dimension), fe(x + T * m.dimension, m.dimension);
    const double c = pruned ? aligner.align_sequence_pruned(fb, fe, ref, AlignmentIter(a, 1), AlignmentIter(a + T, 1), threshold)
                            : aligner.align_sequence_full(fb, fe, ref, AlignmentIter(a, 1), AlignmentIter(a + T, 1));
    if (cost) *cost = c;
  });
}

double sthost_tdp_score(double loop, double forward, double skip, uint32_t silence, uint32_t state, uint64_t jump) {
  Configuration cfg;
  cfg.set("tdp-loop", loop).set("tdp-forward", forward).set("tdp-skip", skip);
  return TdpModel(cfg, (StateIdx)silence).score((StateIdx)state, jump);
}

// Trainer::train on the sietill digits lexicon.  words are lexicon word ids (1..11).  stats: up to
// stats_cap rows of {split, align, estimate, am_score}; *n_stats receives the number of rows produced.
// model_out receives a handle (sthost_model_*) to the trained model.
int sthost_train(uint32_t dim, int pooling, int device, const float* x, uint64_t n_utts, const uint64_t* frame_off,
                 const uint64_t* words, const uint64_t* word_off, double loop, double forward, double skip,
                 uint32_t min_obs, uint32_t num_splits, uint32_t num_aligns, uint32_t num_estimates, int alignment_pruning,
                 double pruning_threshold, int max_approx, const sgmm_alignment_item* initial, const char* mixture_path,
                 const char* alignment_path,
                 double* stats, uint64_t stats_cap, uint64_t* n_stats, sgmm_alignment_item* final_alignment,
                 double* timers /* [4]: align, estimate, score, io seconds */, void** model_out) {
  HostModel* h = new HostModel();
  const int rc = guarded([&] {
    Lexicon lex = build_sietill_lexicon();
    std::vector<size_t> fo(frame_off, frame_off + n_utts + 1), wo(word_off, word_off + n_utts + 1);
    std::vector<WordIdx> w(words, words + word_off[n_utts]);
    Corpus corpus(x, dim, fo, w, wo);
    Configuration& cfg = h->config;
    cfg.set("device-mstep", g_device_mstep != 0);
    cfg.set("device", (unsigned)device).set("tdp-loop", loop).set("tdp-forward", forward).set("tdp-skip", skip);
    cfg.set("min-obs", min_obs).set("num-splits", num_splits).set("num-aligns", num_aligns).set("num-estimates", num_estimates);
    cfg.set("alignment-pruning", alignment_pruning != 0).set("pruning-threshold", pruning_threshold);
    if (mixture_path && *mixture_path) cfg.set("mixture-path", std::string(mixture_path));
    if (alignment_path && *alignment_path) cfg.set("alignment-path", std::string(alignment_path));
    TdpModel tdp(cfg, lex.get_silence_automaton()[0]);
    h->model.reset(new MixtureModel(cfg, dim, lex.num_states(), pooling_of(pooling), max_approx != 0));
    Trainer trainer(cfg, lex, *h->model, tdp, max_approx != 0);
    Alignment init;
    if (initial) {
      init.resize(fo.back());
      memcpy(init.data(), initial, init.size() * sizeof(AlignmentItem));
    }
    trainer.train(corpus, initial ? &init : nullptr);
    const size_t n = trainer.stats().size();
    if (n_stats) *n_stats = n;
    for (size_t i = 0; i < n && i < stats_cap; i++) {
      stats[4 * i + 0] = trainer.stats()[i].split; stats[4 * i + 1] = trainer.stats()[i].align;
      stats[4 * i + 2] = trainer.stats()[i].estimate; stats[4 * i + 3] = trainer.stats()[i].am_score;
    }
    if (final_alignment) memcpy(final_alignment, trainer.alignment().data(), trainer.alignment().size() * sizeof(AlignmentItem));
    if (timers) { timers[0] = trainer.align_seconds; timers[1] = trainer.estimate_seconds; timers[2] = trainer.score_seconds; timers[3] = trainer.io_seconds; }
  });
  if (rc) { delete h; return 1; }
  if (model_out) *model_out = h; else delete h;
  return 0;
}

int sthost_shard_utterances(const uint64_t* frame_off, uint64_t n_utts, int world_size, uint64_t* bounds) {
  return guarded([&] {
    std::vector<size_t> fo(frame_off, frame_off + n_utts + 1);
    std::vector<size_t> b = shard_utterances(fo, world_size);
    for (size_t i = 0; i < b.size(); i++) bounds[i] = b[i];
  });
}

// build_segment_automaton over the sietill lexicon: returns the number of positions written (<= cap)
uint64_t sthost_segment_automaton(const uint64_t* words, uint64_t n_words, uint16_t* out, uint64_t cap) {
  Lexicon lex = build_sietill_lexicon();
  std::vector<MarkovAutomaton const*> parts;
  for (uint64_t i = 0; i < n_words; i++) {
    parts.push_back(&lex.get_silence_automaton());
    parts.push_back(&lex.get_automaton_for_word(words[i]));
  }
  parts.push_back(&lex.get_silence_automaton());
  MarkovAutomaton a = MarkovAutomaton::concat(parts);
  for (size_t i = 0; i < a.num_states() && i < cap; i++) out[i] = a[i];
  return a.num_states();
}

// Recognizer::editDistance (Recognizer.cpp:332-388): out = {total, substitutions, insertions, deletions}
void sthost_edit_distance(const uint64_t* ref, uint64_t n_ref, const uint64_t* hyp, uint64_t n_hyp, uint64_t* out) {
  std::vector<WordIdx> r(ref, ref + n_ref), h(hyp, hyp + n_hyp);
  EDAccumulator e = editDistance(r.cbegin(), r.cend(), h.cbegin(), h.cend());
  out[0] = e.total_count; out[1] = e.substitute_count; out[2] = e.insert_count; out[3] = e.delete_count;
}

// Recognizer::recognize (Recognizer.cpp:37-99) over the sietill digits lexicon with a model handle from
// sthost_model_* / sthost_train.  hyp_words / hyp_off: recognised word ids per segment (hyp_off[n_utts + 1]);
// stats = {WER %, SER %, total errors, substitutions, insertions, deletions, reference words, seconds}.
int sthost_recognize(void* model, const float* x, uint64_t n_utts, const uint64_t* frame_off, const uint64_t* words,
                     const uint64_t* word_off, double loop, double forward, double skip, double word_penalty,
                     double am_threshold, uint64_t* hyp_words, uint64_t hyp_cap, uint64_t* hyp_off, double* stats) {
  HostModel* h = static_cast<HostModel*>(model);
  return guarded([&] {
    Lexicon lex = build_sietill_lexicon();
    std::vector<size_t> fo(frame_off, frame_off + n_utts + 1), wo(word_off, word_off + n_utts + 1);
    std::vector<WordIdx> w(words, words + word_off[n_utts]);
    Corpus corpus(x, h->model->dimension, fo, w, wo);
    Configuration cfg = h->config;
    cfg.set("tdp-loop", loop).set("tdp-forward", forward).set("tdp-skip", skip);
    cfg.set("am-threshold", am_threshold).set("word-penalty", word_penalty);
    cfg.set("max-recognition-runs", (unsigned)n_utts);
    TdpModel tdp(cfg, lex.get_silence_automaton()[0]);
    Recognizer rec(cfg, lex, *h->model, tdp);
    rec.recognize(corpus);
    uint64_t n = 0;
    hyp_off[0] = 0;
    for (size_t s = 0; s < rec.hypotheses.size(); s++) {
      for (WordIdx v : rec.hypotheses[s]) {
        if (n >= hyp_cap) throw std::runtime_error("sthost_recognize: hypothesis buffer too small");
        hyp_words[n++] = v;
      }
      hyp_off[s + 1] = n;
    }
    if (stats) {
      stats[0] = rec.wer; stats[1] = rec.ser; stats[2] = (double)rec.errors.total_count;
      stats[3] = (double)rec.errors.substitute_count; stats[4] = (double)rec.errors.insert_count;
      stats[5] = (double)rec.errors.delete_count; stats[6] = (double)rec.ref_total; stats[7] = rec.seconds;
    }
    // the single-sequence entry point must agree with the batch (first segment)
    if (n_utts > 0) {
      std::vector<WordIdx> one;
      std::pair<FeatureIter, FeatureIter> f = corpus.get_feature_sequence(0);
      rec.recognizeSequence_pruned(f.first, f.second, one);
      if (one != rec.hypotheses[0]) throw std::runtime_error("sthost_recognize: recognizeSequence_pruned differs from the batch");
    }
  });
}

// Trainer::linear_segmentation_approximation (approx != 0) / linear_segmentation_running_sums over one utterance
// (CPU only: a host-only model stands in for the Trainer's scorer).  states[T] receives the aligned states,
// bounds[2] the two boundaries.
int sthost_linear_segmentation(int approx, const float* x, uint64_t T, uint32_t dim, const uint16_t* automaton, uint64_t S,
                               uint16_t* states, uint64_t* bounds) {
  return guarded([&] {
    Lexicon lex = build_sietill_lexicon();
    Configuration cfg;
    cfg.set("host-only", true).set("alignment-pruning", false);
    MixtureModel model(cfg, dim, lex.num_states(), MixtureModel::NO_POOLING, true);
    TdpModel tdp(cfg, lex.get_silence_automaton()[0]);
    Trainer trainer(cfg, lex, model, tdp, true);
    MarkovAutomaton aut;
    aut.states.assign(automaton, automaton + S);
    Alignment ali(T, AlignmentItem());
    FeatureIter fb(x, dim), fe(x + T * dim, dim);
    AlignmentIter ab(ali.data(), 1), ae(ali.data() + T, 1);
    std::pair<size_t, size_t> b = approx ? trainer.linear_segmentation_approximation(aut, fb, fe, ab, ae)
                                         : trainer.linear_segmentation_running_sums(aut, fb, fe, ab, ae);
    for (uint64_t t = 0; t < T; t++) states[t] = ali[t].state;
    bounds[0] = b.first; bounds[1] = b.second;
  });
}


// SignalAnalysis (n-features-file/first/second, deriv-step, energy-max-norm as in the reference's configuration) with an
// optional normalization FILE (the reference's format): process all segments on the GPU of `model`'s context.
int sthost_process_features(void* model, uint32_t n_file, uint32_t n_first, uint32_t n_second, uint32_t deriv_step,
                            int energy_max_norm, const char* normalization_path, const float* raw, uint64_t n_frames,
                            uint64_t n_utts, const uint64_t* frame_off, float* out) {
  HostModel* h = static_cast<HostModel*>(model);
  return guarded([&] {
    Configuration cfg;
    cfg.set("n-features-file", n_file).set("n-features-first", n_first).set("n-features-second", n_second);
    cfg.set("deriv-step", deriv_step).set("energy-max-norm", energy_max_norm != 0);
    SignalAnalysis sa(cfg);
    if (normalization_path && *normalization_path) {
      std::ifstream in(normalization_path, std::ios::binary);
      if (!in.good()) throw std::runtime_error(std::string("cannot open ") + normalization_path);
      sa.read_normalization_file(in);
    }
    std::vector<size_t> fo(frame_off, frame_off + n_utts + 1);
    sa.process_features(h->model->device(), raw, n_frames, fo, out);
  });
}

}  // extern "C"
