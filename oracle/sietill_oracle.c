/* TEST INFRASTRUCTURE — the oracle.  Not part of the product path; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Plain-C restatement of the reference's GMM acoustic-model path (kkromberg/SpeechRecognition,
 * src/sietill).  Every function cites the reference file:line it follows; the arithmetic keeps the
 * reference's operation ORDER so that results are bit-identical to the reference's f64 math
 * (compiled with -ffp-contract=off: the reference's SSE2 code has no FMA either).
 *
 * Parity pinning: the reference has no tests or golden vectors for this path, so
 * this restatement is pinned against outputs of the reference itself — oracle/_ref/libsietill_ref.so,
 * built from the unmodified sources — by tests/test_oracle_vs_reference.py (live, in the build
 * container) and by the committed fixtures under tests/golden/ (made by tools/make_golden.py from
 * oracle/_ref).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_NO_POOLING      0
#define ORC_MIXTURE_POOLING 1
#define ORC_GLOBAL_POOLING  2

typedef struct {
  uint32_t dim, n_mix, pooling;
  uint32_t n_means, n_vars, cap_means, cap_vars;
  /* Mixtures.hpp:72-87 */
  double *means, *mean_acc, *mean_w, *mean_wlog, *mean_wacc;
  uint64_t *mean_refs;
  double *vars, *vars_inv, *var_acc, *var_wacc, *norm;
  uint64_t *var_refs;
  /* mixtures_[m] = ordered list of (mean_idx, var_idx) */
  uint32_t *mix_n, *mix_cap;
  uint32_t **mix_mean, **mix_var;
} orc_model;

static const double kMinimalVariance = 1e-4; /* Mixtures.cpp:166 */

static void grow_means(orc_model *m, uint32_t need) {
  if (need <= m->cap_means) return;
  uint32_t cap = m->cap_means ? m->cap_means : 16;
  while (cap < need) cap *= 2;
  size_t d = m->dim;
  m->means     = realloc(m->means,     cap * d * sizeof(double));
  m->mean_acc  = realloc(m->mean_acc,  cap * d * sizeof(double));
  m->mean_w    = realloc(m->mean_w,    cap * sizeof(double));
  m->mean_wlog = realloc(m->mean_wlog, cap * sizeof(double));
  m->mean_wacc = realloc(m->mean_wacc, cap * sizeof(double));
  m->mean_refs = realloc(m->mean_refs, cap * sizeof(uint64_t));
  m->cap_means = cap;
}

static void grow_vars(orc_model *m, uint32_t need) {
  if (need <= m->cap_vars) return;
  uint32_t cap = m->cap_vars ? m->cap_vars : 16;
  while (cap < need) cap *= 2;
  size_t d = m->dim;
  m->vars     = realloc(m->vars,     cap * d * sizeof(double));
  m->vars_inv = realloc(m->vars_inv, cap * d * sizeof(double));
  m->var_acc  = realloc(m->var_acc,  cap * d * sizeof(double));
  m->var_wacc = realloc(m->var_wacc, cap * sizeof(double));
  m->norm     = realloc(m->norm,     cap * sizeof(double));
  m->var_refs = realloc(m->var_refs, cap * sizeof(uint64_t));
  m->cap_vars = cap;
}

/* MixtureModel::create_mixture_density, Mixtures.cpp:207-233: appends one mean slot, and one variance
 * slot unless var_index names an existing one. */
static void create_density(orc_model *m, uint32_t var_index) {
  size_t d = m->dim;
  grow_means(m, m->n_means + 1);
  uint32_t mi = m->n_means++;
  m->mean_refs[mi] = 1; m->mean_w[mi] = 0.0; m->mean_wlog[mi] = 0.0; m->mean_wacc[mi] = 0.0;
  for (size_t i = 0; i < d; i++) { m->means[mi * d + i] = 0.0; m->mean_acc[mi * d + i] = 0.0; }
  if (var_index >= m->n_vars) {
    grow_vars(m, m->n_vars + 1);
    uint32_t vi = m->n_vars++;
    m->norm[vi] = 0.0; m->var_wacc[vi] = 0.0; m->var_refs[vi] = 1;
    for (size_t i = 0; i < d; i++) {
      m->vars[vi * d + i] = 0.0; m->vars_inv[vi * d + i] = 0.0; m->var_acc[vi * d + i] = kMinimalVariance;
    }
  }
}

static void mix_push(orc_model *m, uint32_t s, uint32_t mean_idx, uint32_t var_idx) {
  if (m->mix_n[s] == m->mix_cap[s]) {
    uint32_t cap = m->mix_cap[s] ? 2 * m->mix_cap[s] : 4;
    m->mix_mean[s] = realloc(m->mix_mean[s], cap * sizeof(uint32_t));
    m->mix_var[s]  = realloc(m->mix_var[s],  cap * sizeof(uint32_t));
    m->mix_cap[s] = cap;
  }
  m->mix_mean[s][m->mix_n[s]] = mean_idx;
  m->mix_var[s][m->mix_n[s]]  = var_idx;
  m->mix_n[s]++;
}

/* MixtureModel::MixtureModel (training branch), Mixtures.cpp:158-193: one density per mixture; with
 * GLOBAL_POOLING every density shares variance 0. */
orc_model *orc_model_create(uint32_t dim, uint32_t n_mix, uint32_t pooling) {
  orc_model *m = calloc(1, sizeof(orc_model));
  m->dim = dim; m->n_mix = n_mix; m->pooling = pooling;
  m->mix_n = calloc(n_mix, sizeof(uint32_t)); m->mix_cap = calloc(n_mix, sizeof(uint32_t));
  m->mix_mean = calloc(n_mix, sizeof(uint32_t *)); m->mix_var = calloc(n_mix, sizeof(uint32_t *));
  for (uint32_t s = 0; s < n_mix; s++) {
    uint32_t mean_idx = m->n_means;
    uint32_t var_idx = (pooling != ORC_GLOBAL_POOLING) ? m->n_vars : 0;
    create_density(m, var_idx);
    mix_push(m, s, mean_idx, var_idx);
  }
  return m;
}

void orc_model_destroy(orc_model *m) {
  if (!m) return;
  free(m->means); free(m->mean_acc); free(m->mean_w); free(m->mean_wlog); free(m->mean_wacc);
  free(m->mean_refs); free(m->vars); free(m->vars_inv); free(m->var_acc); free(m->var_wacc);
  free(m->norm); free(m->var_refs);
  for (uint32_t s = 0; s < m->n_mix; s++) { free(m->mix_mean[s]); free(m->mix_var[s]); }
  free(m->mix_n); free(m->mix_cap); free(m->mix_mean); free(m->mix_var);
  free(m);
}

/* sizes: [0]=dim [1]=n_mix [2]=M [3]=V [4]=total live densities (same layout as ref_model_sizes) */
void orc_model_sizes(const orc_model *m, uint64_t *out) {
  out[0] = m->dim; out[1] = m->n_mix; out[2] = m->n_means; out[3] = m->n_vars;
  uint64_t t = 0;
  for (uint32_t s = 0; s < m->n_mix; s++) t += m->mix_n[s];
  out[4] = t;
}

#define CPY(dst, src, n) do { if ((dst) && (n)) memcpy((dst), (src), (n) * sizeof(*(src))); } while (0)

void orc_model_export(const orc_model *m, double *means, double *mean_acc, double *mean_w,
                      double *mean_wlog, double *mean_wacc, uint64_t *mean_refs, double *vars,
                      double *vars_inv, double *var_acc, double *var_wacc, uint64_t *var_refs,
                      double *norm, uint32_t *mix_offsets, uint32_t *mean_idx, uint32_t *var_idx) {
  size_t d = m->dim, M = m->n_means, V = m->n_vars;
  CPY(means, m->means, M * d); CPY(mean_acc, m->mean_acc, M * d); CPY(mean_w, m->mean_w, M);
  CPY(mean_wlog, m->mean_wlog, M); CPY(mean_wacc, m->mean_wacc, M); CPY(mean_refs, m->mean_refs, M);
  CPY(vars, m->vars, V * d); CPY(vars_inv, m->vars_inv, V * d); CPY(var_acc, m->var_acc, V * d);
  CPY(var_wacc, m->var_wacc, V); CPY(var_refs, m->var_refs, V); CPY(norm, m->norm, V);
  uint32_t pos = 0;
  for (uint32_t s = 0; s < m->n_mix; s++) {
    if (mix_offsets) mix_offsets[s] = pos;
    for (uint32_t k = 0; k < m->mix_n[s]; k++, pos++) {
      if (mean_idx) mean_idx[pos] = m->mix_mean[s][k];
      if (var_idx)  var_idx[pos]  = m->mix_var[s][k];
    }
  }
  if (mix_offsets) mix_offsets[m->n_mix] = pos;
}

/* Replace every table (used to load a model produced elsewhere, e.g. exported from the reference). */
void orc_model_import(orc_model *m, uint32_t M, uint32_t V, const double *means, const double *mean_acc,
                      const double *mean_w, const double *mean_wlog, const double *mean_wacc,
                      const uint64_t *mean_refs, const double *vars, const double *vars_inv,
                      const double *var_acc, const double *var_wacc, const uint64_t *var_refs,
                      const double *norm, const uint32_t *mix_offsets, const uint32_t *mean_idx,
                      const uint32_t *var_idx) {
  size_t d = m->dim;
  grow_means(m, M); grow_vars(m, V);
  m->n_means = M; m->n_vars = V;
  CPY(m->means, means, M * d); CPY(m->mean_acc, mean_acc, M * d); CPY(m->mean_w, mean_w, M);
  CPY(m->mean_wlog, mean_wlog, M); CPY(m->mean_wacc, mean_wacc, M); CPY(m->mean_refs, mean_refs, M);
  CPY(m->vars, vars, V * d); CPY(m->vars_inv, vars_inv, V * d); CPY(m->var_acc, var_acc, V * d);
  CPY(m->var_wacc, var_wacc, V); CPY(m->var_refs, var_refs, V); CPY(m->norm, norm, V);
  for (uint32_t s = 0; s < m->n_mix; s++) {
    m->mix_n[s] = 0;
    for (uint32_t p = mix_offsets[s]; p < mix_offsets[s + 1]; p++) mix_push(m, s, mean_idx[p], var_idx[p]);
  }
}

/* ------------------------------------------------------------------------------------------------
 * scoring
 * ---------------------------------------------------------------------------------------------- */

/* MixtureModel::density_score_sse, Mixtures.cpp:645-690.  The SSE2 loop keeps two f64 partial sums
 * (even dims in lane 0, odd dims in lane 1), adds them (add_double_sse_register, :633-639), then the
 * odd tail (:672-675); finally norm + dist/2, then minus the log weight (:679-680). */
double orc_density_score(const orc_model *m, const float *x, uint32_t state, uint32_t k) {
  const size_t d = m->dim;
  const double *mu = m->means    + (size_t)m->mix_mean[state][k] * d;
  const double *iv = m->vars_inv + (size_t)m->mix_var[state][k]  * d;
  double lane0 = 0.0, lane1 = 0.0;
  size_t even = d - d % 2;
  for (size_t i = 0; i < even; i += 2) {
    double a = (double)x[i] - mu[i], b = (double)x[i + 1] - mu[i + 1];
    a = a * a; b = b * b;
    a = a * iv[i]; b = b * iv[i + 1];
    lane0 = lane0 + a; lane1 = lane1 + b;
  }
  double dist = lane0 + lane1;
  if (d % 2 == 1) {
    double c = (double)x[d - 1] - mu[d - 1];
    dist += c * c * iv[d - 1];
  }
  double score = m->norm[m->mix_var[state][k]] + dist / 2;
  score -= m->mean_wlog[m->mix_mean[state][k]];
  return score;
}

/* MixtureModel::min_score, Mixtures.cpp:696-713: start 1e10 / index 0, strict '<' (first minimum wins). */
double orc_min_score(const orc_model *m, const float *x, uint32_t state, uint32_t *argmin) {
  uint32_t best = 0;
  double min_score = 1e10;
  for (uint32_t k = 0; k < m->mix_n[state]; k++) {
    double s = orc_density_score(m, x, state, k);
    if (s < min_score) { best = k; min_score = s; }
  }
  if (argmin) *argmin = best;
  return min_score;
}

/* MixtureModel::sum_score, Mixtures.cpp:719-728 */
double orc_sum_score(const orc_model *m, const float *x, uint32_t state) {
  double acc = 0.0;
  for (uint32_t k = 0; k < m->mix_n[state]; k++) acc += exp(-1 * orc_density_score(m, x, state, k));
  return -1 * log(acc);
}

/* MixtureModel::score, Mixtures.cpp:737-744 */
double orc_score(const orc_model *m, const float *x, uint32_t state, int max_approx) {
  return max_approx ? orc_min_score(m, x, state, NULL) : orc_sum_score(m, x, state);
}

/* min_score for every (frame, listed state) — same output layout as ref_min_score.  states==NULL means
 * all n_mix states in order. */
void orc_min_score_batch(const orc_model *m, const float *x, uint64_t n_frames, const uint32_t *states,
                         uint64_t n_states, double *scores, uint32_t *argmin) {
  if (!states) n_states = m->n_mix;
#pragma omp parallel for schedule(static)
  for (int64_t f = 0; f < (int64_t)n_frames; f++) {
    for (uint64_t j = 0; j < n_states; j++) {
      uint32_t a;
      double s = orc_min_score(m, x + (size_t)f * m->dim, states ? states[j] : (uint32_t)j, &a);
      scores[f * n_states + j] = s;
      if (argmin) argmin[f * n_states + j] = a;
    }
  }
}

/* FeatureScorer::score for every (frame, state): min_score or sum_score per max_approx (Mixtures.cpp:737-744) */
void orc_score_batch(const orc_model *m, const float *x, uint64_t n_frames, int max_approx, double *scores) {
#pragma omp parallel for schedule(static)
  for (int64_t f = 0; f < (int64_t)n_frames; f++)
    for (uint32_t j = 0; j < m->n_mix; j++)
      scores[(size_t)f * m->n_mix + j] = orc_score(m, x + (size_t)f * m->dim, j, max_approx);
}

/* Trainer::calc_am_score, Training.cpp:585-612: sequential f64 sum of score(x_f, aligned state) / frames */
double orc_am_score(const orc_model *m, const uint32_t *states, const float *x, uint64_t n_frames,
                    int max_approx) {
  double total = 0.0;
  for (uint64_t f = 0; f < n_frames; f++) total += orc_score(m, x + f * m->dim, states[f], max_approx);
  return total / n_frames;
}

/* ------------------------------------------------------------------------------------------------
 * EM: accumulate / finalize / split / eliminate
 * ---------------------------------------------------------------------------------------------- */

/* MixtureModel::reset_accumulators, Mixtures.cpp:237-245: the variance sums restart at 1e-4, not 0. */
void orc_reset_accumulators(orc_model *m) {
  size_t d = m->dim;
  for (size_t i = 0; i < (size_t)m->n_means * d; i++) m->mean_acc[i] = 0.0;
  for (size_t i = 0; i < m->n_means; i++) m->mean_wacc[i] = 0.0;
  for (size_t i = 0; i < (size_t)m->n_vars * d; i++) m->var_acc[i] = kMinimalVariance;
  for (size_t i = 0; i < m->n_vars; i++) m->var_wacc[i] = 0.0;
}

/* MixtureModel::accumulate, Mixtures.cpp:278-372.  states[f] = aligned state of frame f.
 * argmin_out (optional) receives the density chosen for each frame in the max-approx branch. */
void orc_accumulate(orc_model *m, const uint32_t *states, const float *x, uint64_t n_frames,
                    int first_pass, int max_approx, uint32_t *argmin_out) {
  size_t d = m->dim;
  orc_reset_accumulators(m);
  double *memb = NULL; size_t memb_cap = 0;
  for (uint64_t f = 0; f < n_frames; f++) {
    const float *xf = x + f * d;
    uint32_t s = states[f], n = m->mix_n[s];
    if (n > memb_cap) { memb_cap = 2 * n; memb = realloc(memb, memb_cap * sizeof(double)); }
    for (uint32_t k = 0; k < n; k++) memb[k] = 0.0;
    uint32_t best = 0;
    if (max_approx && !first_pass) orc_min_score(m, xf, s, &best);       /* :296-298 */
    if (argmin_out) argmin_out[f] = best;
    for (uint32_t k = 0; k < n; k++) {                                   /* :300-310 */
      if (first_pass) memb[0] = 1.0;
      else if (max_approx && k == best) memb[k] = 1.0;
      else if (!max_approx) memb[k] = exp(-1 * orc_density_score(m, xf, s, k));
    }
    if (!max_approx) {                                                   /* :316-324 */
      double sum = 0.0;
      for (uint32_t k = 0; k < n; k++) sum += memb[k];
      for (uint32_t k = 0; k < n; k++) memb[k] = memb[k] / sum;
    }
    for (uint32_t k = 0; k < n; k++) {                                   /* :333-357 */
      if (memb[k] < 1e-8) continue;
      uint32_t mi = m->mix_mean[s][k], vi = m->mix_var[s][k];
      m->mean_wacc[mi] += memb[k];
      m->var_wacc[vi]  += memb[k];
      double *ma = m->mean_acc + (size_t)mi * d, *va = m->var_acc + (size_t)vi * d;
      for (size_t i = 0; i < d; i++) { double y = xf[i]; ma[i] = ma[i] + memb[k] * y; }
      for (size_t i = 0; i < d; i++) { double y = xf[i]; va[i] = va[i] + memb[k] * y * y; }
    }
  }
  free(memb);
}

/* MixtureModel::calculate_variance, Mixtures.cpp:251-275: E[x^2] - mean^2, inverse, and the log-norm. */
static void calculate_variance(orc_model *m, uint32_t vi, const double *mean) {
  size_t d = m->dim;
  double *v = m->vars + (size_t)vi * d, *inv = m->vars_inv + (size_t)vi * d;
  const double *acc = m->var_acc + (size_t)vi * d;
  for (size_t i = 0; i < d; i++) v[i] = acc[i] / m->var_wacc[vi];
  for (size_t i = 0; i < d; i++) v[i] = v[i] - mean[i] * mean[i];
  for (size_t i = 0; i < d; i++) inv[i] = 1 / v[i];
  double n = d * log(2 * M_PI);
  for (size_t i = 0; i < d; i++) n = n + log(v[i]);
  m->norm[vi] = n / 2;
}

/* MixtureModel::finalize, Mixtures.cpp:374-461 */
void orc_finalize(orc_model *m) {
  size_t d = m->dim;
  double total_observations = 0.0;
  double *pooled = malloc(d * sizeof(double));
  for (uint32_t s = 0; s < m->n_mix; s++) {
    double total_mix = 0.0;
    for (uint32_t k = 0; k < m->mix_n[s]; k++) {
      uint32_t mi = m->mix_mean[s][k], vi = m->mix_var[s][k];
      total_mix += m->mean_wacc[mi];
      for (size_t i = 0; i < d; i++) m->means[(size_t)mi * d + i] = m->mean_acc[(size_t)mi * d + i] / m->mean_wacc[mi];
      if (m->pooling == ORC_NO_POOLING) calculate_variance(m, vi, m->means + (size_t)mi * d);
    }
    for (uint32_t k = 0; k < m->mix_n[s]; k++) {
      uint32_t mi = m->mix_mean[s][k];
      m->mean_w[mi] = m->mean_wacc[mi] / total_mix;
      m->mean_wlog[mi] = log(m->mean_w[mi]);
    }
    if (m->pooling == ORC_MIXTURE_POOLING) {                             /* :408-427 */
      for (size_t i = 0; i < d; i++) pooled[i] = 0.0;
      for (uint32_t k = 0; k < m->mix_n[s]; k++)
        for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] + m->mean_acc[(size_t)m->mix_mean[s][k] * d + i];
      for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] / total_mix;
      /* The reference reads mixtures_[m][0] even when eliminate() emptied the mixture (:425) — the vector's
       * storage still holds the last erased entry, so the stale variance slot is recomputed from a 0/0 mean and
       * becomes NaN.  Reproduced here: mix_var[s][0] keeps the erased entry exactly like the vector's storage. */
      if (m->mix_cap[s]) calculate_variance(m, m->mix_var[s][0], pooled);
    }
    total_observations += total_mix;
  }
  if (m->pooling == ORC_GLOBAL_POOLING) {                                /* :432-450 */
    for (size_t i = 0; i < d; i++) pooled[i] = 0.0;
    for (uint32_t s = 0; s < m->n_mix; s++)
      for (uint32_t k = 0; k < m->mix_n[s]; k++)
        for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] + m->mean_acc[(size_t)m->mix_mean[s][k] * d + i];
    for (size_t i = 0; i < d; i++) pooled[i] = pooled[i] / total_observations;
    calculate_variance(m, 0, pooled);
  }
  free(pooled);
}

/* MixtureModel::split + update_split_densities, Mixtures.cpp:465-543.  Mixtures ascending, densities
 * descending; the clone is appended to the mixture; mean_acc of the clone stays 0; weights, log
 * weights and counts are copied, not halved; with NO_POOLING the variance slot is cloned too. */
void orc_split(orc_model *m, uint64_t min_obs) {
  size_t d = m->dim;
  for (uint32_t s = 0; s < m->n_mix; s++) {
    for (int k = (int)m->mix_n[s] - 1; k >= 0; k--) {
      uint32_t mo = m->mix_mean[s][k], vo = m->mix_var[s][k];
      if (!(m->mean_wacc[mo] >= (double)min_obs)) continue;
      uint32_t ms = m->n_means;
      uint32_t vs = (m->pooling == ORC_NO_POOLING) ? m->n_vars : vo;
      create_density(m, vs);
      m->mean_w[ms] = m->mean_w[mo]; m->mean_wlog[ms] = m->mean_wlog[mo]; m->mean_wacc[ms] = m->mean_wacc[mo];
      for (size_t i = 0; i < d; i++) {
        double shift = sqrt(m->vars[(size_t)vo * d + i]);
        double plus = m->means[(size_t)mo * d + i] + shift, minus = m->means[(size_t)mo * d + i] - shift;
        m->means[(size_t)mo * d + i] = plus;
        m->means[(size_t)ms * d + i] = minus;
      }
      if (m->pooling == ORC_NO_POOLING) {
        m->var_wacc[vs] = m->var_wacc[vo];
        memcpy(m->var_acc  + (size_t)vs * d, m->var_acc  + (size_t)vo * d, d * sizeof(double));
        memcpy(m->vars     + (size_t)vs * d, m->vars     + (size_t)vo * d, d * sizeof(double));
        memcpy(m->vars_inv + (size_t)vs * d, m->vars_inv + (size_t)vo * d, d * sizeof(double));
        m->norm[vs] = m->norm[vo];
      }
      mix_push(m, s, ms, vs);
    }
  }
}

/* MixtureModel::eliminate, Mixtures.cpp:547-576: drop densities with count < min_obs (descending). */
void orc_eliminate(orc_model *m, double min_obs) {
  for (uint32_t s = 0; s < m->n_mix; s++) {
    for (int k = (int)m->mix_n[s] - 1; k >= 0; k--) {
      uint32_t mi = m->mix_mean[s][k], vi = m->mix_var[s][k];
      if (m->mean_wacc[mi] < min_obs) {
        for (uint32_t j = k; j + 1 < m->mix_n[s]; j++) {
          m->mix_mean[s][j] = m->mix_mean[s][j + 1];
          m->mix_var[s][j]  = m->mix_var[s][j + 1];
        }
        m->mix_n[s]--;
        m->mean_refs[mi] = 0;
        if (m->pooling == ORC_NO_POOLING) m->var_refs[vi] = 0;
      }
    }
  }
}

/* MixtureModel::num_densities, Mixtures.cpp:580-582 */
uint64_t orc_num_densities(const orc_model *m) {
  uint64_t n = 0;
  for (uint32_t i = 0; i < m->n_means; i++) n += m->mean_refs[i] != 0;
  return n;
}

/* ------------------------------------------------------------------------------------------------
 * alignment
 * ---------------------------------------------------------------------------------------------- */

/* TdpModel::score, TdpModel.cpp:19-29: the silence state always pays tdp_forward. */
typedef struct { double loop, forward, skip; uint32_t silence_state; } orc_tdp;

static double tdp_score(const orc_tdp *t, uint32_t state, uint32_t jump) {
  if (state == t->silence_state) return t->forward;
  switch (jump) {
    case 0: return t->loop;
    case 1: return t->forward;
    case 2: return t->skip;
  }
  return INFINITY;
}

double orc_tdp_score(double loop, double forward, double skip, uint32_t silence, uint32_t state, uint32_t jump) {
  orc_tdp t = {loop, forward, skip, silence};
  return tdp_score(&t, state, jump);
}

/* Aligner::align_sequence_full, Alignment.cpp:50-144.  Literal restatement including the 0-1-2 band
 * (:73-74,82,87), the tie order loop > forward > skip with strict '<' (:96-113), the TDP looked up with
 * the SOURCE position's state (:95,100,108), and the backtrace from (T-1, S-1) (:129-138).
 * out_states[t] = reference[s_t]; out_pos (optional) = automaton position s_t.  Returns Q(T-1,S-1).
 * `emis` (optional, [T*S]) is filled with the emission score of every in-band cell (NaN elsewhere)
 * and `margin` (optional, [T]) with the best-vs-second-best predecessor gap at the backtraced cell,
 * so that tests can classify a mismatch as a near-tie. */
double orc_align_full(const orc_model *m, double loop, double forward, double skip, uint32_t silence,
                      int max_approx, const float *x, uint64_t T, const uint32_t *ref, uint64_t S,
                      uint32_t *out_states, uint32_t *out_pos, double *margin) {
  orc_tdp tdp = {loop, forward, skip, silence};
  const size_t d = m->dim;
  int32_t *bp = malloc(S * T * sizeof(int32_t));
  double *gap = margin ? malloc(S * T * sizeof(double)) : NULL;
  for (size_t i = 0; i < S * T; i++) bp[i] = -1;
  size_t ncost = T > S ? T : S;   /* the reference sizes these by T (:62-63) and needs S <= T */
  double *prev = malloc(ncost * sizeof(double)), *cur = malloc(ncost * sizeof(double));
  for (size_t i = 0; i < ncost; i++) { prev[i] = INFINITY; cur[i] = INFINITY; }

  long max_state = 2;
  long min_state = (long)S - 1 - ((long)T - 2) * 2;
  prev[0] = orc_score(m, x, ref[0], max_approx);
  for (uint64_t t = 1; t < T; t++, min_state += 2, max_state += 2) {
    long lo = min_state > 0 ? min_state : 0;
    long hi = (long)S - 1 < max_state ? (long)S - 1 : max_state;
    for (long s = lo; s <= hi; s++) {
      double local = orc_score(m, x + t * d, ref[s], max_approx);
      double best = prev[s] + tdp_score(&tdp, ref[s], 0);
      double second = INFINITY;
      int taken = 0;
      if (s > 0) {
        double fwd = prev[s - 1] + tdp_score(&tdp, ref[s - 1], 1);
        if (fwd < best) { second = best; best = fwd; taken = 1; } else if (fwd < second) second = fwd;
      }
      if (s > 1) {
        double skp = prev[s - 2] + tdp_score(&tdp, ref[s - 2], 2);
        if (skp < best) { second = best; best = skp; taken = 2; } else if (skp < second) second = skp;
      }
      cur[s] = local + best;
      bp[s * T + t] = (int32_t)(s - taken);
      if (gap) gap[s * T + t] = second - best;
    }
    memcpy(prev, cur, ncost * sizeof(double));
  }
  double cost = cur[S - 1];
  long s = (long)S - 1;
  for (long t = (long)T - 1; t >= 0; t--) {
    if (out_states) out_states[t] = ref[s];
    if (out_pos) out_pos[t] = (uint32_t)s;
    if (margin) margin[t] = t > 0 ? gap[s * T + t] : INFINITY;
    if (t > 0) s = bp[s * T + t];
    if (s < 0) s = 0;   /* unreachable end state: the reference would read out of bounds here */
  }
  free(bp); free(gap); free(prev); free(cur);
  return cost;
}

/* Aligner::align_sequence_pruned, Alignment.cpp:149-288, as a DP over arrays instead of std::map
 * beams of heap nodes.  Differences from the full aligner that are reproduced: TDP looked up with the
 * DESTINATION state (:191); candidates reach a target in the order skip, forward, loop and replace only
 * on strictly smaller cost (:199-204); after each frame hypotheses above best + threshold are dropped
 * (:227-239); the backtrace starts at the HIGHEST position alive in the last frame (:245-256). */
double orc_align_pruned(const orc_model *m, double loop, double forward, double skip, uint32_t silence,
                        int max_approx, const float *x, uint64_t T, const uint32_t *ref, uint64_t S,
                        double threshold, uint32_t *out_states, uint32_t *out_pos) {
  orc_tdp tdp = {loop, forward, skip, silence};
  const size_t d = m->dim;
  int32_t *bp = malloc(S * T * sizeof(int32_t));
  double *prev = malloc(S * sizeof(double)), *cur = malloc(S * sizeof(double));
  char *palive = calloc(S, 1), *calive = calloc(S, 1);
  prev[0] = orc_score(m, x, ref[0], max_approx);
  palive[0] = 1;
  for (uint64_t t = 1; t < T; t++) {
    memset(calive, 0, S);
    for (uint64_t p = 0; p < S; p++) {
      if (!palive[p]) continue;
      for (uint32_t jump = 0; jump <= 2; jump++) {
        uint64_t s = p + jump;
        if (s >= S) break;
        double c = prev[p];
        c += tdp_score(&tdp, ref[s], jump);
        c += orc_score(m, x + t * d, ref[s], max_approx);
        if (!calive[s]) { calive[s] = 1; cur[s] = c; bp[s * T + t] = (int32_t)p; }
        else if (cur[s] > c) { cur[s] = c; bp[s * T + t] = (int32_t)p; }
      }
    }
    double best = INFINITY;
    for (uint64_t s = 0; s < S; s++) if (calive[s] && cur[s] < best) best = cur[s];
    double upper = best + threshold;
    for (uint64_t s = 0; s < S; s++) if (calive[s] && cur[s] > upper) calive[s] = 0;
    double *td = prev; prev = cur; cur = td;
    char *ta = palive; palive = calive; calive = ta;
  }
  long s = 0;
  for (uint64_t i = 0; i < S; i++) if (palive[i]) s = (long)i;
  double cost = prev[s];
  for (long t = (long)T - 1; t >= 1; t--) {
    if (out_states) out_states[t] = ref[s];
    if (out_pos) out_pos[t] = (uint32_t)s;
    s = bp[s * T + t];
  }
  if (out_states) out_states[0] = ref[0];
  if (out_pos) out_pos[0] = 0;
  free(bp); free(prev); free(cur); free(palive); free(calive);
  return cost;
}

/* The loop of Trainer::train, Training.cpp:168-183 with alignment-pruning false, over a batch.
 * frame_off/aut_off are offsets (in frames / positions) of each utterance.  OpenMP over utterances is
 * used only for CPU-baseline timing with `threads` > 1 (the reference itself is single-threaded here). */
double orc_align_batch(const orc_model *m, double loop, double forward, double skip, uint32_t silence,
                       int max_approx, const float *x, uint64_t U, const uint64_t *frame_off,
                       const uint32_t *automata, const uint32_t *aut_off, uint32_t *out_states,
                       double *costs, int threads) {
  double total = 0.0;
  if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 4) num_threads(threads) reduction(+ : total)
  for (int64_t u = 0; u < (int64_t)U; u++) {
    uint64_t f0 = frame_off[u], f1 = frame_off[u + 1];
    double c = orc_align_full(m, loop, forward, skip, silence, max_approx, x + f0 * m->dim, f1 - f0,
                              automata + aut_off[u], aut_off[u + 1] - aut_off[u], out_states + f0, NULL, NULL);
    if (costs) costs[u] = c;
    total += c;
  }
  return total;
}

/* ------------------------------------------------------------------------------------------------
 * recognition (BASELINE config 4)
 * ---------------------------------------------------------------------------------------------- */

/* Recognizer::recognizeSequence_pruned, Recognizer.cpp:103-232, restated over a lexicon given as arrays
 * (word w = positions word_states[word_off[w] .. word_off[w+1]) ).  Kept literally:
 *  - hypothesis table indexed max_len * word + position (:117-118), hypotheses visited in that order (:128);
 *  - the start hypothesis is slot 0 with score 0 (:121);
 *  - a hypothesis on the LAST position of its word only leaves the word: it enters position 0 and 1 of every word with
 *    word_penalty (0 for silence) + tdp(first_state, init + 1) and — for BOTH entry positions — the emission score of
 *    the word's FIRST state (:131-157);
 *  - inside a word: jumps 0/1/2 with the DESTINATION state's tdp (:162-187);
 *  - a candidate is dropped when its score before the emission already exceeds the target's (:142-145,172-175),
 *    otherwise it replaces the target on strict improvement (:153,183);
 *  - pruning against best + am_threshold, best word end per frame with strict '>' (:190-207);
 *  - traceback over word ends, silence not emitted (:222-231); Book::bkp is 16 bits (Recognizer.hpp:79).
 * Returns the number of words written to out_words (at most max_words are stored). */
uint64_t orc_recognize(const orc_model *m, double loop, double forward, double skip, uint32_t silence_state,
                       int max_approx, const float *x, uint64_t T, uint32_t n_words, const uint32_t *word_off,
                       const uint32_t *word_states, uint32_t silence_word, double word_penalty,
                       double am_threshold, uint32_t *out_words, uint64_t max_words, double *out_score) {
  orc_tdp tdp = {loop, forward, skip, silence_state};
  uint32_t max_len = 0, n_states = 0;
  for (uint32_t w = 0; w < n_words; w++) {
    uint32_t L = word_off[w + 1] - word_off[w];
    if (L > max_len) max_len = L;
    for (uint32_t p = word_off[w]; p < word_off[w + 1]; p++)
      if (word_states[p] + 1 > n_states) n_states = word_states[p] + 1;
  }
  typedef struct { double score; uint16_t word, bkp; int state_idx; } book;
  const size_t n_slots = (size_t)n_words * max_len;
  book *cur = (book *)malloc(n_slots * sizeof(book)), *nxt = (book *)malloc(n_slots * sizeof(book));
  book *tb = (book *)malloc((T + 1) * sizeof(book));
  double *am = (double *)malloc(n_states * sizeof(double));
  for (size_t i = 0; i < n_slots; i++) { book b = {INFINITY, 0, 0, 0}; cur[i] = b; nxt[i] = b; }
  for (uint64_t t = 0; t <= T; t++) { book b = {0.0, 0, 0, 0}; tb[t] = b; }
  for (uint32_t s = 0; s < n_states; s++) am[s] = INFINITY;
  cur[0].score = 0.0;
  for (uint64_t t = 1; t <= T; t++) {
    const float *xt = x + (t - 1) * m->dim;
    double best = INFINITY;
    for (size_t i = 0; i < n_slots; i++) {
      const book h = cur[i];
      if (h.score == INFINITY) continue;
      const uint32_t Lw = word_off[h.word + 1] - word_off[h.word];
      if ((uint32_t)h.state_idx == Lw - 1) {
        for (uint32_t w = 0; w < n_words; w++) {
          const double wp = (w != silence_word) ? word_penalty : 0.0;
          const uint32_t first = word_states[word_off[w]];
          for (uint32_t init = 0; init <= 1; init++) {
            book *tg = &nxt[(size_t)max_len * w + init];
            double ns = h.score + wp + tdp_score(&tdp, first, init + 1);
            if (ns > tg->score) continue;
            if (am[first] == INFINITY) am[first] = orc_score(m, xt, first, max_approx);
            ns += am[first];
            if (tg->score > ns) {
              tg->score = ns; tg->state_idx = (int)init; tg->bkp = (uint16_t)(t - 1); tg->word = (uint16_t)w;
              if (ns < best) best = ns;
            }
          }
        }
      } else {
        for (uint32_t jump = 0; jump <= 2; jump++) {
          const uint32_t np = (uint32_t)h.state_idx + jump;
          if (np >= Lw) break;
          const uint32_t st = word_states[word_off[h.word] + np];
          book *tg = &nxt[(size_t)max_len * h.word + np];
          double ns = h.score + tdp_score(&tdp, st, jump);
          if (ns > tg->score) continue;
          if (am[st] == INFINITY) am[st] = orc_score(m, xt, st, max_approx);
          ns += am[st];
          if (tg->score > ns) {
            tg->score = ns; tg->state_idx = (int)np; tg->bkp = h.bkp; tg->word = h.word;
            if (ns < best) best = ns;
          }
        }
      }
    }
    tb[t].score = INFINITY;
    for (size_t i = 0; i < n_slots; i++) {
      book *h = &nxt[i];
      if (h->score > best + am_threshold) { h->score = INFINITY; continue; }
      if ((uint32_t)h->state_idx == word_off[h->word + 1] - word_off[h->word] - 1 && tb[t].score > h->score) {
        tb[t].score = h->score; tb[t].word = h->word; tb[t].bkp = h->bkp;
      }
    }
    book *sw = cur; cur = nxt; nxt = sw;
    for (size_t i = 0; i < n_slots; i++) { book b = {INFINITY, 0, 0, 0}; nxt[i] = b; }
    for (uint32_t s = 0; s < n_states; s++) am[s] = INFINITY;
  }
  if (out_score) *out_score = tb[T].score;
  uint64_t n = 0;
  uint32_t *rev = (uint32_t *)malloc((T + 1) * sizeof(uint32_t));
  for (uint64_t t = T; t > 0; t = tb[t].bkp)
    if (tb[t].word != silence_word) rev[n++] = tb[t].word;
  for (uint64_t i = 0; i < n && i < max_words; i++) out_words[i] = rev[n - 1 - i];
  free(rev); free(cur); free(nxt); free(tb); free(am);
  return n;
}

/* ---- load-time feature processing ------------------------------------------------------------- */
/* SignalAnalysis::process_features, SignalAnalysis.cpp:379-399, on ONE utterance: raw [T][nf] -> out [T][nt],
 * nt = nf + n1 + n2.  add_deltas :320-337, mean / variance normalisation :389-394 (mean == NULL: not applied),
 * energy_max_normalization :341-350.  Needs T > step and n2 <= n1 <= nf, as the reference does. */
void orc_process_features(const float *raw, uint64_t T, uint32_t nf, uint32_t n1, uint32_t n2, uint32_t step,
                          const double *mean, const double *stddev, int energy_max_norm, float *out) {
  const uint32_t nt = nf + n1 + n2;
  for (uint64_t t = 0; t < T; t++)                                             /* :383-387 */
    for (uint32_t i = 0; i < nf; i++) out[t * nt + i] = raw[t * nf + i];
  for (uint64_t t = 0; t < T; t++) {                                           /* :322-328 */
    const uint64_t a = t > step ? t : step;
    for (uint32_t i = 0; i < n1; i++) out[t * nt + nf + i] = out[a * nt + i] - out[(a - step) * nt + i];
  }
  for (uint64_t t = 0; t < T; t++) {                                           /* :329-336 */
    const uint64_t b = (t < T - 1 - step ? t : T - 1 - step) + step;
    for (uint32_t i = 0; i < n2; i++) out[t * nt + nf + n1 + i] = out[b * nt + nf + i] - out[t * nt + nf + i];
  }
  if (mean) {                                                                  /* :389-394: two passes, f32 stores */
    for (uint64_t t = 0; t < T; t++) {
      for (uint32_t c = 0; c < nt; c++) out[t * nt + c] = (float)((double)out[t * nt + c] - mean[c]);
      for (uint32_t c = 0; c < nt; c++) out[t * nt + c] = (float)((double)out[t * nt + c] / stddev[c]);
    }
  }
  if (energy_max_norm) {                                                       /* :341-350 */
    float mx = -INFINITY;
    for (uint64_t t = 0; t < T; t++) mx = out[t * nt] > mx ? out[t * nt] : mx;
    for (uint64_t t = 0; t < T; t++) out[t * nt] -= mx;
  }
}
