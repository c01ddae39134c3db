// MixtureModel: host-side model storage + M-step, GPU-side E-step and scoring.
// Semantics follow the reference's src/sietill/Mixtures.cpp (line numbers cited per method); the code is a
// fresh implementation around the C-ABI of include/sietill_b200.h.
#include <cmath>
#include <cstring>
#include <fstream>
#include <istream>
#include <ostream>

#include "sietill_host.hpp"

namespace sietill_b200 {

namespace {
const char kMagic[8] = {'M', 'I', 'X', 'S', 'E', 'T', 0, 0};   // Mixtures.cpp:153
const uint32_t kVersion = 2u;                                  // Mixtures.cpp:154

[[noreturn]] void fail(std::string const& msg) { throw std::runtime_error("MixtureModel: " + msg); }

template <typename T>
void get(std::istream& in, T& v, const char* what) {
  in.read(reinterpret_cast<char*>(&v), sizeof(T));
  if (in.gcount() != (std::streamsize)sizeof(T)) fail(std::string("Error reading ") + what);
}
template <typename T>
void put(std::ostream& out, T const& v) { out.write(reinterpret_cast<const char*>(&v), sizeof(T)); }

// one "accumulator block" of a .mix file: u32 n, then n x {u32 dim, f64[dim] sums, f64 weight}
void read_block(std::istream& in, size_t dim, std::vector<size_t>& refs, std::vector<double>& sums,
                std::vector<double>& weight) {
  uint32_t n;
  get(in, n, "size");
  refs.assign(n, 0);
  sums.assign((size_t)n * dim, 0.0);
  weight.assign(n, 0.0);
  for (size_t i = 0; i < n; i++) {
    uint32_t d;
    get(in, d, "dimension");
    if (d != dim) fail("Invalid dimension");
    in.read(reinterpret_cast<char*>(&sums[i * dim]), sizeof(double) * dim);
    if ((size_t)in.gcount() != sizeof(double) * dim) fail("Error reading features");
    get(in, weight[i], "weight");
  }
}
void write_block(std::ostream& out, uint32_t live, uint32_t dim, std::vector<size_t> const& refs,
                 std::vector<double> const& sums, std::vector<double> const& weight) {
  put(out, live);
  for (size_t i = 0; i < refs.size(); i++) {
    if (refs[i] == 0) continue;
    put(out, dim);
    out.write(reinterpret_cast<const char*>(&sums[i * dim]), sizeof(double) * dim);
    put(out, weight[i]);
  }
}
uint32_t compact_map(std::vector<size_t> const& refs, std::vector<uint32_t>& map) {
  map.assign(refs.size(), 0);
  uint32_t n = 0;
  for (size_t i = 0; i < refs.size(); i++)
    if (refs[i] > 0) map[i] = n++;
  return n;
}
}  // namespace

uint64_t alignment_checksum(const AlignmentItem* items, size_t n) {
  const uint64_t* w = reinterpret_cast<const uint64_t*>(items);
  uint64_t h[4] = {0x9E3779B97F4A7C15ull, 0xC2B2AE3D27D4EB4Full, 0x165667B19E3779F9ull, 0x27D4EB2F165667C5ull};
  size_t i = 0;
  for (; i + 4 <= n; i += 4)
    for (int k = 0; k < 4; k++) h[k] = (h[k] ^ w[i + k]) * 0x100000001B3ull + (h[k] >> 29);
  for (; i < n; i++) h[0] = (h[0] ^ w[i]) * 0x100000001B3ull + (h[0] >> 29);
  return h[0] ^ (h[1] * 3) ^ (h[2] * 5) ^ (h[3] * 7) ^ n;
}

// ---- Device -----------------------------------------------------------------------------------------
Device::Device(int device_index) : index_(device_index) {
  if (sgmm_create(device_index, &ctx_) != 0)
    throw std::runtime_error(std::string("sgmm_create: ") + sgmm_last_error(nullptr));
}
Device::~Device() { sgmm_destroy(ctx_); }
void Device::check(int rc, const char* what) const {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + sgmm_last_error(ctx_));
}

// ---- construction (Mixtures.cpp:158-233) -----------------------------------------------------------------
MixtureModel::MixtureModel(Configuration const& config, size_t dimension, size_t num_mixtures,
                           VarianceModel var_model, bool max_approx)
    : dimension(dimension), var_model(var_model), max_approx_(max_approx), mixtures_(num_mixtures),
      stale_first_(num_mixtures, MixtureDensity(0, 0)), minimal_variance_value_(1e-4) {
  const std::string from = config.get_string("load-mixtures-from", "");
  const std::string action = config.get_string("action", "");
  // "host-only": M-step and file formats without a GPU (CPU tests); every compute method then throws
  if (!config.get_bool("host-only", false)) device_.reset(new Device((int)config.get_uint("device", 0)));
  device_mstep_ = device_ && config.get_bool("device-mstep", true);
  if (!from.empty() && action == "recognize") {
    std::ifstream in(from.c_str(), std::ios::binary);
    if (!in.good()) fail("cannot open " + from);
    read(in);
  } else {
    for (size_t m = 0; m < num_mixtures; m++) {
      // one density per mixture; with GLOBAL_POOLING they all share variance slot 0 (:176-190)
      const DensityIdx var_index = var_model != GLOBAL_POOLING ? (DensityIdx)var_refs_.size() : 0;
      mixtures_[m].push_back(create_mixture_density(var_index));
    }
  }
}

MixtureModel::~MixtureModel() {}

// Mixtures.cpp:207-233: a new mean slot always; a new variance slot unless var_index names an existing one.
MixtureDensity MixtureModel::create_mixture_density(DensityIdx var_index) {
  const DensityIdx mean_index = (DensityIdx)mean_refs_.size();
  mean_refs_.push_back(1);
  mean_weights_.push_back(0.0);
  mean_weights_log_.push_back(0.0);
  mean_weight_accumulators_.push_back(0.0);
  means_.resize(means_.size() + dimension, 0.0);
  mean_accumulators_.resize(mean_accumulators_.size() + dimension, 0.0);
  if (var_index >= var_refs_.size()) {
    norm_.push_back(0.0);
    var_weight_accumulators_.push_back(0.0);
    var_refs_.push_back(1);
    vars_.resize(vars_.size() + dimension, 0.0);
    vars_inv_.resize(vars_inv_.size() + dimension, 0.0);
    var_accumulators_.resize(var_accumulators_.size() + dimension, minimal_variance_value_);
  }
  return MixtureDensity(mean_index, var_index);
}

// Mixtures.cpp:237-245: the second-order sums restart at 1e-4, not 0.
void MixtureModel::reset_accumulators() {
  stats_on_device_ = false;
  std::fill(mean_accumulators_.begin(), mean_accumulators_.end(), 0.0);
  std::fill(mean_weight_accumulators_.begin(), mean_weight_accumulators_.end(), 0.0);
  std::fill(var_accumulators_.begin(), var_accumulators_.end(), minimal_variance_value_);
  std::fill(var_weight_accumulators_.begin(), var_weight_accumulators_.end(), 0.0);
}

// ---- device plumbing ----------------------------------------------------------------------------------
namespace {
void pack_tables(size_t dim, std::vector<Mixture> const& mixtures, std::vector<double> const& means,
                 std::vector<double> const& vars_inv, std::vector<double> const& norm,
                 std::vector<double> const& logw, std::vector<uint32_t>& off, std::vector<double>& pm,
                 std::vector<double>& pv, std::vector<double>& pn, std::vector<double>& pw) {
  off.assign(mixtures.size() + 1, 0);
  for (size_t m = 0; m < mixtures.size(); m++) off[m + 1] = off[m] + (uint32_t)mixtures[m].size();
  const size_t D = off.back();
  pm.resize(D * dim); pv.resize(D * dim); pn.resize(D); pw.resize(D);
  size_t j = 0;
  for (Mixture const& mix : mixtures)
    for (MixtureDensity const& md : mix) {
      memcpy(&pm[j * dim], &means[(size_t)md.mean_idx * dim], dim * sizeof(double));
      memcpy(&pv[j * dim], &vars_inv[(size_t)md.var_idx * dim], dim * sizeof(double));
      pn[j] = norm[md.var_idx];
      pw[j] = logw[md.mean_idx];
      j++;
    }
}
}  // namespace

void MixtureModel::sync_device() const {
  if (!device_) fail("built host-only: no GPU context (there is no CPU scoring path)");
  if (uploaded_version_ == version_) return;
  ensure_host();
  std::vector<uint32_t> off;
  std::vector<double> pm, pv, pn, pw;
  pack_tables(dimension, mixtures_, means_, vars_inv_, norm_, mean_weights_log_, off, pm, pv, pn, pw);
  device_->check(sgmm_set_model(device_->ctx(), (uint32_t)dimension, (uint32_t)mixtures_.size(), off.data(), pm.data(),
                                pv.data(), pn.data(), pw.data()),
                 "sgmm_set_model");
  // FeatureScorer::score on this context = min_score or sum_score, as the constructor's max_approx says (:737-744)
  device_->check(sgmm_set_option(device_->ctx(), "max_approx", max_approx_ ? 1 : 0), "sgmm_set_option");
  uploaded_version_ = version_;
}

Device& MixtureModel::scratch() const {
  if (!device_) fail("built host-only: no GPU context (there is no CPU scoring path)");
  if (!scratch_) scratch_.reset(new Device(device_->index()));   // same GPU as the main context
  if (scratch_version_ != version_) {
    ensure_host();
    std::vector<uint32_t> off;
    std::vector<double> pm, pv, pn, pw;
    pack_tables(dimension, mixtures_, means_, vars_inv_, norm_, mean_weights_log_, off, pm, pv, pn, pw);
    scratch_->check(sgmm_set_model(scratch_->ctx(), (uint32_t)dimension, (uint32_t)mixtures_.size(), off.data(),
                                   pm.data(), pv.data(), pn.data(), pw.data()),
                    "sgmm_set_model");
    scratch_version_ = version_;
  }
  return *scratch_;
}

void MixtureModel::set_corpus_features(FeatureIter begin, FeatureIter end) const {
  if (!device_) fail("built host-only: no GPU context");
  if (begin.size != dimension) fail("feature dimension differs from the model's");
  const size_t n = (size_t)(end - begin);
  // identity = (pointer, length, fingerprint of ~2048 sampled words): guards against a recycled address
  uint64_t print = 0xCBF29CE484222325ull ^ n;
  {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(*begin);
    const size_t words = n * dimension, step = words / 2048 + 1;
    for (size_t i = 0; i < words; i += step) print = (print ^ w[i]) * 0x100000001B3ull;
    if (words) print = (print ^ w[words - 1]) * 0x100000001B3ull;
  }
  if (*begin == resident_feat_ && n == resident_frames_ && print == resident_feat_print_) return;
  device_->check(sgmm_set_features(device_->ctx(), *begin, n, (uint32_t)dimension), "sgmm_set_features");
  resident_feat_ = *begin;
  resident_frames_ = n;
  resident_feat_print_ = print;
  resident_ali_ = nullptr;   // sgmm_set_features drops the resident alignment
}

void MixtureModel::comm_init(int rank, int world_size, const void* id, size_t id_bytes) {
  if (!device_) fail("built host-only: no GPU context");
  device_->check(sgmm_comm_init(device_->ctx(), rank, world_size, id, id_bytes), "sgmm_comm_init");
}

// ---- E-step (Mixtures.cpp:278-372) on the GPU ------------------------------------------------------------
void MixtureModel::accumulate(ConstAlignmentIter alignment_begin, ConstAlignmentIter alignment_end,
                              FeatureIter feature_begin, FeatureIter feature_end, bool first_pass, bool max_approx) {
  if (alignment_begin.size != 1) fail("accumulate: num-max-aligns must be 1");
  const size_t n = (size_t)(feature_end - feature_begin);
  if ((size_t)(alignment_end - alignment_begin) != n) fail("accumulate: alignment and feature ranges differ in length");
  sync_device();
  set_corpus_features(feature_begin, feature_end);
  sgmm_ctx* c = device_->ctx();
  const uint64_t sum = alignment_checksum(*alignment_begin, n);
  if (*alignment_begin != resident_ali_ || n != resident_ali_frames_ || sum != resident_ali_sum_) {
    device_->check(sgmm_set_alignment(c, reinterpret_cast<const sgmm_alignment_item*>(*alignment_begin), n),
                   "sgmm_set_alignment");
    resident_ali_ = *alignment_begin;
    resident_ali_frames_ = n;
    resident_ali_sum_ = sum;
  }
  double score_sum = 0.0;
  if (!first_pass && !max_approx)   // posterior-weighted branch, :300-357 (score_sum = sum of sum scores)
    device_->check(sgmm_accumulate_soft(c, nullptr, nullptr, nullptr, &score_sum), "sgmm_accumulate_soft");
  else
    device_->check(sgmm_accumulate(c, first_pass ? 1 : 0, nullptr, nullptr, nullptr, &score_sum), "sgmm_accumulate");
  last_score_sum_ = score_sum;
  // the statistics stay on the GPU (one all-reduced block); finalize() consumes them there, the host arrays follow on demand
  stats_on_device_ = true;
  if (!device_mstep_) ensure_host();
}

// Whatever the GPU holds that the host arrays do not: the accumulators of the last accumulate() and / or the parameters
// of the last device finalize().  The per-density statistics are folded into the reference's slots (:333-357); the
// variance sums start from the 1e-4 seed exactly once, whatever the number of GPUs behind the all-reduce.
void MixtureModel::ensure_host() const {
  if (!stats_on_device_ && !params_on_device_) return;
  MixtureModel& self = *const_cast<MixtureModel*>(this);
  sgmm_ctx* c = device_->ctx();
  std::vector<uint32_t> off(mixtures_.size() + 1, 0);
  for (size_t m = 0; m < mixtures_.size(); m++) off[m + 1] = off[m] + (uint32_t)mixtures_[m].size();
  const size_t D = off.back(), d = dimension;
  if (stats_on_device_) {
    std::vector<double> count(D), sum_x(D * d), sum_xx(D * d);
    device_->check(sgmm_get_statistics(c, count.data(), sum_x.data(), sum_xx.data(), nullptr), "sgmm_get_statistics");
    self.reset_accumulators();
    size_t j = 0;
    for (Mixture const& mix : mixtures_)
      for (MixtureDensity const& md : mix) {
        self.mean_weight_accumulators_[md.mean_idx] += count[j];
        self.var_weight_accumulators_[md.var_idx] += count[j];
        double* ma = &self.mean_accumulators_[(size_t)md.mean_idx * d];
        double* va = &self.var_accumulators_[(size_t)md.var_idx * d];
        for (size_t i = 0; i < d; i++) ma[i] += sum_x[j * d + i];
        for (size_t i = 0; i < d; i++) va[i] += sum_xx[j * d + i];
        j++;
      }
    stats_on_device_ = false;
  }
  if (params_on_device_) {   // exactly what the device scores with (its log differs from glibc's by <= 1 ulp)
    std::vector<double> mean(D * d), inv(D * d), var(D * d), norm(D), logw(D), w(D);
    device_->check(sgmm_get_model(c, mean.data(), inv.data(), norm.data(), logw.data(), var.data(), w.data()), "sgmm_get_model");
    size_t j = 0;
    for (Mixture const& mix : mixtures_)
      for (MixtureDensity const& md : mix) {
        memcpy(&self.means_[(size_t)md.mean_idx * d], &mean[j * d], d * sizeof(double));
        memcpy(&self.vars_[(size_t)md.var_idx * d], &var[j * d], d * sizeof(double));
        memcpy(&self.vars_inv_[(size_t)md.var_idx * d], &inv[j * d], d * sizeof(double));
        self.norm_[md.var_idx] = norm[j];
        self.mean_weights_[md.mean_idx] = w[j];
        self.mean_weights_log_[md.mean_idx] = logw[j];
        j++;
      }
    // The reference recomputes the variance slot of a mixture that eliminate() emptied from a 0/0 pooled mean
    // (Mixtures.cpp:425, see finalize_host): no density uses it, the device never builds it; the host copy follows suit.
    if (var_model == MIXTURE_POOLING) {
      const std::vector<double> nan_mean(d, std::nan(""));
      for (size_t m = 0; m < mixtures_.size(); m++)
        if (mixtures_[m].empty()) self.calculate_variance(stale_first_[m].var_idx, nan_mean.data());
    }
    params_on_device_ = false;
  }
}

// ---- M-step (host, f64) ------------------------------------------------------------------------------------
// Mixtures.cpp:251-275: sigma^2 = E[x^2] - mean^2, its inverse, norm = (d log 2pi + sum log sigma^2) / 2
void MixtureModel::calculate_variance(DensityIdx var_idx, const double* mean) {
  const size_t d = dimension;
  double* v = &vars_[(size_t)var_idx * d];
  double* inv = &vars_inv_[(size_t)var_idx * d];
  const double* acc = &var_accumulators_[(size_t)var_idx * d];
  const double w = var_weight_accumulators_[var_idx];
  for (size_t i = 0; i < d; i++) v[i] = acc[i] / w;
  for (size_t i = 0; i < d; i++) v[i] = v[i] - mean[i] * mean[i];
  for (size_t i = 0; i < d; i++) inv[i] = 1 / v[i];
  double n = d * std::log(2 * M_PI);
  for (size_t i = 0; i < d; i++) n = n + std::log(v[i]);
  norm_[var_idx] = n / 2;
}

// Mixtures.cpp:374-461.  With the statistics still on the GPU the M-step runs there (sgmm_mstep: the same operation
// order; every derived table is rebuilt by device kernels) and the model never leaves the device.
void MixtureModel::finalize() {
  if (stats_on_device_ && device_mstep_) {
    const int pooling = var_model == NO_POOLING ? 0 : var_model == MIXTURE_POOLING ? 1 : 2;
    device_->check(sgmm_mstep(device_->ctx(), pooling, minimal_variance_value_), "sgmm_mstep");
    device_->check(sgmm_set_option(device_->ctx(), "max_approx", max_approx_ ? 1 : 0), "sgmm_set_option");
    mark_dirty();
    uploaded_version_ = version_;
    params_on_device_ = true;
    return;
  }
  ensure_host();
  finalize_host();
  mark_dirty();
}

void MixtureModel::finalize_host() {
  const size_t d = dimension;
  double total_observations = 0.0;
  std::vector<double> pooled(d);
  for (size_t m = 0; m < mixtures_.size(); m++) {
    Mixture const& mix = mixtures_[m];
    double total_mix = 0.0;
    for (MixtureDensity const& md : mix) {
      total_mix += mean_weight_accumulators_[md.mean_idx];
      for (size_t i = 0; i < d; i++)
        means_[(size_t)md.mean_idx * d + i] =
            mean_accumulators_[(size_t)md.mean_idx * d + i] / mean_weight_accumulators_[md.mean_idx];
      if (var_model == NO_POOLING) calculate_variance(md.var_idx, &means_[(size_t)md.mean_idx * d]);
    }
    for (MixtureDensity const& md : mix) {
      mean_weights_[md.mean_idx] = mean_weight_accumulators_[md.mean_idx] / total_mix;
      mean_weights_log_[md.mean_idx] = std::log(mean_weights_[md.mean_idx]);
    }
    if (var_model == MIXTURE_POOLING) {   // :408-427: variance around the mixture's pooled mean
      std::fill(pooled.begin(), pooled.end(), 0.0);
      for (MixtureDensity const& md : mix)
        for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] + mean_accumulators_[(size_t)md.mean_idx * d + i];
      for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] / total_mix;
      // The reference indexes mixtures_[m][0] even when eliminate() emptied the mixture (:425): the vector's
      // storage still holds the last erased entry, whose variance slot is then recomputed from a 0/0 mean.
      const MixtureDensity first = mix.empty() ? stale_first_[m] : mix[0];
      calculate_variance(first.var_idx, pooled.data());
    }
    total_observations += total_mix;
  }
  if (var_model == GLOBAL_POOLING) {      // :432-450: one variance around the global mean
    std::fill(pooled.begin(), pooled.end(), 0.0);
    for (Mixture const& mix : mixtures_)
      for (MixtureDensity const& md : mix)
        for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] + mean_accumulators_[(size_t)md.mean_idx * d + i];
    for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] / total_observations;
    calculate_variance(0, pooled.data());
  }
}

// Mixtures.cpp:465-543: mixtures ascending, densities descending; the clone goes to the end of the mixture
// with mean - sigma (the original moves to mean + sigma); weight, log weight and count are copied, not
// halved; with NO_POOLING the variance slot is cloned too.
void MixtureModel::split(size_t min_obs) {
  ensure_host();
  const size_t d = dimension;
  for (Mixture& mix : mixtures_) {
    for (size_t k = mix.size(); k-- > 0;) {
      const MixtureDensity orig = mix[k];
      if (!(mean_weight_accumulators_[orig.mean_idx] >= (double)min_obs)) continue;
      const DensityIdx new_var = var_model == NO_POOLING ? (DensityIdx)var_refs_.size() : orig.var_idx;
      const MixtureDensity clone = create_mixture_density(new_var);
      mean_weights_[clone.mean_idx] = mean_weights_[orig.mean_idx];
      mean_weights_log_[clone.mean_idx] = mean_weights_log_[orig.mean_idx];
      mean_weight_accumulators_[clone.mean_idx] = mean_weight_accumulators_[orig.mean_idx];
      for (size_t i = 0; i < d; i++) {
        const double shift = std::sqrt(vars_[(size_t)orig.var_idx * d + i]);
        const double mu = means_[(size_t)orig.mean_idx * d + i];
        means_[(size_t)orig.mean_idx * d + i] = mu + shift;
        means_[(size_t)clone.mean_idx * d + i] = mu - shift;
      }
      if (var_model == NO_POOLING) {
        var_weight_accumulators_[clone.var_idx] = var_weight_accumulators_[orig.var_idx];
        for (size_t i = 0; i < d; i++) {
          var_accumulators_[(size_t)clone.var_idx * d + i] = var_accumulators_[(size_t)orig.var_idx * d + i];
          vars_[(size_t)clone.var_idx * d + i] = vars_[(size_t)orig.var_idx * d + i];
          vars_inv_[(size_t)clone.var_idx * d + i] = vars_inv_[(size_t)orig.var_idx * d + i];
        }
        norm_[clone.var_idx] = norm_[orig.var_idx];
      }
      mix.push_back(clone);
    }
  }
  mark_dirty();
}

// Mixtures.cpp:547-576: drop densities seen fewer than min_obs times (descending inside a mixture).
void MixtureModel::eliminate(double min_obs) {
  ensure_host();
  for (size_t m = 0; m < mixtures_.size(); m++) {
    Mixture& mix = mixtures_[m];
    for (size_t k = mix.size(); k-- > 0;) {
      const MixtureDensity md = mix[k];
      if (mean_weight_accumulators_[md.mean_idx] < min_obs) {
        if (mix.size() == 1) stale_first_[m] = md;
        mix.erase(mix.begin() + (ptrdiff_t)k);
        mean_refs_[md.mean_idx] = 0;
        if (var_model == NO_POOLING) var_refs_[md.var_idx] = 0;
      }
    }
  }
  mark_dirty();
}

size_t MixtureModel::num_densities() const {   // Mixtures.cpp:580-582: live means
  size_t n = 0;
  for (size_t r : mean_refs_) n += r != 0;
  return n;
}

// ---- scoring (Mixtures.cpp:696-744) through K1 ---------------------------------------------------------------
void MixtureModel::prepare_sequence(FeatureIter const& start, FeatureIter const& end) {
  if (start.size != dimension) fail("feature dimension differs from the model's");
  Device& s = scratch();
  const size_t T = (size_t)(end - start), N = mixtures_.size();
  seq_scores_.resize(T * N);
  seq_argmin_.resize(T * N);
  if (T) {
    s.check(sgmm_set_features(s.ctx(), *start, T, (uint32_t)dimension), "sgmm_set_features");
    s.check(sgmm_score_all(s.ctx(), 0, T, seq_scores_.data(), seq_argmin_.data()), "sgmm_score_all");
  }
  seq_sum_scores_.clear();
  if (!max_approx_ && T) {   // score() of this model is sum_score: a second batched pass with the other semantics
    seq_sum_scores_.resize(T * N);
    s.check(sgmm_set_option(s.ctx(), "max_approx", 0), "sgmm_set_option");
    const int rc = sgmm_score_all(s.ctx(), 0, T, seq_sum_scores_.data(), nullptr);
    sgmm_set_option(s.ctx(), "max_approx", 1);
    s.check(rc, "sgmm_score_all");
  }
  seq_begin_ = *start;
  seq_frames_ = T;
  seq_version_ = version_;
}

std::pair<double, DensityIdx> MixtureModel::min_score(FeatureIter const& iter, StateIdx mixture_idx) const {
  if (mixture_idx >= mixtures_.size()) fail("min_score: mixture index out of range");
  const size_t N = mixtures_.size();
  if (seq_begin_ && seq_version_ == version_ && *iter >= seq_begin_ &&
      *iter < seq_begin_ + seq_frames_ * dimension) {
    const size_t t = (size_t)(*iter - seq_begin_) / dimension;
    return std::make_pair(seq_scores_[t * N + mixture_idx], (DensityIdx)seq_argmin_[t * N + mixture_idx]);
  }
  // not prepared: one frame through K1 (correct, slow — call prepare_sequence first)
  Device& s = scratch();
  std::vector<double> sc(N);
  std::vector<uint32_t> am(N);
  s.check(sgmm_set_features(s.ctx(), *iter, 1, (uint32_t)dimension), "sgmm_set_features");
  s.check(sgmm_score_all(s.ctx(), 0, 1, sc.data(), am.data()), "sgmm_score_all");
  return std::make_pair(sc[mixture_idx], (DensityIdx)am[mixture_idx]);
}

// Mixtures.cpp:719-728; `weights` is ignored there as well
double MixtureModel::sum_score(FeatureIter const& iter, StateIdx mixture_idx, std::vector<double>*) const {
  if (mixture_idx >= mixtures_.size()) fail("sum_score: mixture index out of range");
  const size_t N = mixtures_.size();
  if (seq_begin_ && seq_version_ == version_ && !seq_sum_scores_.empty() && *iter >= seq_begin_ &&
      *iter < seq_begin_ + seq_frames_ * dimension) {
    const size_t t = (size_t)(*iter - seq_begin_) / dimension;
    return seq_sum_scores_[t * N + mixture_idx];
  }
  Device& s = scratch();   // not prepared: one frame through the sum kernel
  std::vector<double> sc(N);
  s.check(sgmm_set_features(s.ctx(), *iter, 1, (uint32_t)dimension), "sgmm_set_features");
  s.check(sgmm_set_option(s.ctx(), "max_approx", 0), "sgmm_set_option");
  const int rc = sgmm_score_all(s.ctx(), 0, 1, sc.data(), nullptr);
  sgmm_set_option(s.ctx(), "max_approx", 1);
  s.check(rc, "sgmm_score_all");
  return sc[mixture_idx];
}

double MixtureModel::score(FeatureIter const& iter, StateIdx mixture_idx) const {   // :737-744
  if (!max_approx_) return sum_score(iter, mixture_idx, nullptr);
  return min_score(iter, mixture_idx).first;
}

// ---- .mix v2 (Mixtures.cpp:748-878) ----------------------------------------------------------------------------
void MixtureModel::read(std::istream& in) {
  char magic[8];
  in.read(magic, sizeof(magic));
  stats_on_device_ = params_on_device_ = false;   // the file replaces everything
  if (in.gcount() != (std::streamsize)sizeof(magic)) fail("Error reading magic header");
  if (memcmp(magic, kMagic, 8) != 0) fail("Invalid magic header");
  uint32_t version, dim;
  get(in, version, "version");
  if (version != kVersion) fail("Invalid version");
  get(in, dim, "dimension");
  if (dim != dimension) fail("Invalid dimension");

  read_block(in, dimension, mean_refs_, mean_accumulators_, mean_weight_accumulators_);
  mean_weights_.assign(mean_refs_.size(), 0.0);
  mean_weights_log_.assign(mean_refs_.size(), 0.0);
  means_.assign(mean_accumulators_.size(), 0.0);
  read_block(in, dimension, var_refs_, var_accumulators_, var_weight_accumulators_);
  vars_.assign(var_accumulators_.size(), 0.0);
  vars_inv_.assign(var_accumulators_.size(), 0.0);
  norm_.assign(var_refs_.size(), 0.0);

  uint32_t density_count;
  get(in, density_count, "density count");
  std::vector<MixtureDensity> densities;
  densities.reserve(density_count);
  for (uint32_t i = 0; i < density_count; i++) {
    uint32_t mi, vi;
    get(in, mi, "mean_idx");
    if (mi >= mean_refs_.size()) fail("Invalid mean_idx");
    get(in, vi, "var_idx");
    if (vi >= var_refs_.size()) fail("Invalid var_idx");
    mean_refs_[mi] += 1;
    var_refs_[vi] += 1;
    densities.push_back(MixtureDensity(mi, vi));
  }
  uint32_t mixture_count;
  get(in, mixture_count, "mixture count");
  mixtures_.assign(mixture_count, Mixture());
  stale_first_.assign(mixture_count, MixtureDensity(0, 0));
  for (uint32_t m = 0; m < mixture_count; m++) {
    uint32_t n;
    get(in, n, "density count for mixture");
    mixtures_[m].reserve(n);
    for (uint32_t k = 0; k < n; k++) {
      uint32_t di;
      get(in, di, "density idx");
      if (di >= densities.size()) fail("Invalid density idx");
      mixtures_[m].push_back(densities[di]);
      double w;
      get(in, w, "density weight");
      if (!(w == mean_weight_accumulators_[densities[di].mean_idx])) fail("Inconsistent density weight");
    }
  }
  finalize();   // the file holds sufficient statistics; parameters are regenerated (:842)
}

void MixtureModel::write(std::ostream& out) const {
  ensure_host();
  out.write(kMagic, sizeof(kMagic));
  put(out, kVersion);
  put(out, (uint32_t)dimension);
  std::vector<uint32_t> mean_map, var_map;
  const uint32_t n_mean = compact_map(mean_refs_, mean_map), n_var = compact_map(var_refs_, var_map);
  write_block(out, n_mean, (uint32_t)dimension, mean_refs_, mean_accumulators_, mean_weight_accumulators_);
  write_block(out, n_var, (uint32_t)dimension, var_refs_, var_accumulators_, var_weight_accumulators_);
  uint32_t D = 0;
  for (Mixture const& mix : mixtures_) D += (uint32_t)mix.size();
  put(out, D);
  for (Mixture const& mix : mixtures_)
    for (MixtureDensity const& md : mix) {
      put(out, mean_map[md.mean_idx]);
      put(out, var_map[md.var_idx]);
    }
  put(out, (uint32_t)mixtures_.size());
  uint32_t running = 0;
  for (Mixture const& mix : mixtures_) {
    put(out, (uint32_t)mix.size());
    for (MixtureDensity const& md : mix) {
      put(out, running++);
      put(out, mean_weight_accumulators_[md.mean_idx]);
    }
  }
}

// ---- raw tables (tests) ------------------------------------------------------------------------------------
MixtureModel::Tables MixtureModel::export_tables() const {
  ensure_host();
  Tables t;
  t.means = means_; t.mean_acc = mean_accumulators_; t.mean_w = mean_weights_; t.mean_wlog = mean_weights_log_;
  t.mean_wacc = mean_weight_accumulators_; t.vars = vars_; t.vars_inv = vars_inv_; t.var_acc = var_accumulators_;
  t.var_wacc = var_weight_accumulators_; t.norm = norm_;
  t.mean_refs.assign(mean_refs_.begin(), mean_refs_.end());
  t.var_refs.assign(var_refs_.begin(), var_refs_.end());
  t.mix_offsets.assign(mixtures_.size() + 1, 0);
  for (size_t m = 0; m < mixtures_.size(); m++) {
    t.mix_offsets[m + 1] = t.mix_offsets[m] + (uint32_t)mixtures_[m].size();
    for (MixtureDensity const& md : mixtures_[m]) { t.mean_idx.push_back(md.mean_idx); t.var_idx.push_back(md.var_idx); }
  }
  return t;
}

void MixtureModel::import_tables(Tables const& t) {
  if (t.mix_offsets.size() != mixtures_.size() + 1) fail("import_tables: wrong number of mixtures");
  stats_on_device_ = params_on_device_ = false;
  means_ = t.means; mean_accumulators_ = t.mean_acc; mean_weights_ = t.mean_w; mean_weights_log_ = t.mean_wlog;
  mean_weight_accumulators_ = t.mean_wacc; vars_ = t.vars; vars_inv_ = t.vars_inv; var_accumulators_ = t.var_acc;
  var_weight_accumulators_ = t.var_wacc; norm_ = t.norm;
  mean_refs_.assign(t.mean_refs.begin(), t.mean_refs.end());
  var_refs_.assign(t.var_refs.begin(), t.var_refs.end());
  for (size_t m = 0; m < mixtures_.size(); m++) {
    mixtures_[m].clear();
    for (uint32_t p = t.mix_offsets[m]; p < t.mix_offsets[m + 1]; p++)
      mixtures_[m].push_back(MixtureDensity(t.mean_idx[p], t.var_idx[p]));
  }
  mark_dirty();
}

}  // namespace sietill_b200
