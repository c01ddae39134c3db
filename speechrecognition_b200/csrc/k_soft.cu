// the NON-max-approx branch of the path.
//   * MixtureModel::sum_score (reference src/sietill/Mixtures.cpp:719-728): score = -log sum_k exp(-density_score_k),
//     densities in mixture order, plain sum starting from 0.0 (no log-sum-exp stabilisation: the reference has none,
//     so a frame far from every density scores +inf exactly as there);
//   * the posterior-weighted branch of MixtureModel::accumulate (Mixtures.cpp:300-357): p_k = exp(-score_k), normalised
//     by their in-order sum, entries below 1e-8 skipped, weight/sum_x/sum_xx += p, p*x, (p*x)*x.
// Density scores are the reference's f64 operation order (sgmm_exact_score), so they are bit-identical; exp / log are
// CUDA's f64 functions (<= 1 ulp from glibc's), hence sum scores and soft statistics agree to ~1e-15 relative, not
// bit for bit (tests demand 1e-12).  FP64-pipe bound; this branch has no f32 filter (every density contributes).
//
// Soft accumulation is deterministic: frames are stably sorted by aligned state (the radix sort of k_accum.cu), a
// state's frames are cut into items of kSoftItem consecutive frames, one CTA walks an item in frame order keeping
// the state's accumulators [2 dim + 1][densities] in shared memory, then one thread per accumulator adds the items'
// partials in item order.  The summation tree depends only on the data.
#include "sgmm_internal.cuh"

namespace {

ExactTables exact_tables(const sgmm_ctx* c) {
  ExactTables m{};
  m.meanT = c->d_meanT.p; m.invT = c->d_invT.p; m.mean_tail = c->d_mean_tail.p; m.inv_tail = c->d_inv_tail.p;
  m.norm = c->d_norm.p; m.logw = c->d_logw.p; m.dim = c->dim; m.npair = c->npair; m.Dtot = c->Dtot;
  m.meanR = c->d_meanR.p; m.invU = c->d_invU.p; m.var_of = c->d_var_of.p; m.dimP = c->dimP;
  return m;
}

// ---- sum scores of (frame, state) cells ---------------------------------------------------------------------
// One warp per 32 consecutive frames of one utterance (gathered) or of the dense range; lane = frame.  Table reads
// are warp-uniform (one density at a time, all lanes), features come from shared memory as f64.
struct SumParams {
  ExactTables m;
  const uint32_t* state_off;
  const float* x;
  uint64_t n_tasks;
  const uint4* tasks;
  const uint64_t* frame_off;
  const uint32_t* uniq;
  const uint32_t* uniq_off;
  const uint32_t* nuniq_pad;
  const uint64_t* E_off;
  double* E;
  uint16_t* Amin;
  int dense;
  uint64_t f0, nframes;
  uint32_t n_states;
  double* scores;
  uint32_t* argmin;
};

constexpr int kSumWarps = 4;

__device__ __forceinline__ double sum_score_lane(const ExactTables& m, const float* __restrict__ xs, uint32_t k0,
                                                 uint32_t k1) {
  double acc = 0.0;                                                        // Mixtures.cpp:721
  for (uint32_t k = k0; k < k1; k++) acc = __dadd_rn(acc, exp(-sgmm_exact_score(m, xs, k)));   // :723-725
  return -log(acc);                                                        // :727
}

__global__ void __launch_bounds__(kSumWarps * 32) k_emission_sum(SumParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = (uint64_t)blockIdx.x * kSumWarps + warp;
  if (task >= p.n_tasks) return;
  uint64_t fbase;
  uint32_t nvalid, nu, out_stride;
  const uint32_t* ulist = nullptr;
  double* out;
  uint32_t* amin = nullptr;
  uint16_t* amin16 = nullptr;
  if (p.dense) {
    const uint64_t r0 = task * 32;
    fbase = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)32, p.nframes - r0);
    nu = p.n_states;
    out_stride = p.n_states;
    out = p.scores ? p.scores + r0 * p.n_states : nullptr;
    amin = p.argmin ? p.argmin + r0 * p.n_states : nullptr;
  } else {
    const uint4 t = p.tasks[task];
    const uint64_t uf0 = p.frame_off[t.x];
    const uint32_t T = (uint32_t)(p.frame_off[t.x + 1] - uf0);
    fbase = uf0 + t.y;
    nvalid = min(32u, T - t.y);
    ulist = p.uniq + p.uniq_off[t.x];
    nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
    out_stride = p.nuniq_pad[t.x];
    out = p.E + p.E_off[t.x] + (uint64_t)t.y * out_stride;
    amin16 = p.Amin + p.E_off[t.x] + (uint64_t)t.y * out_stride;
  }
  if (lane >= (int)nvalid) return;
  const float* xf = p.x + (fbase + lane) * p.m.dim;
  for (uint32_t j = 0; j < nu; j++) {
    const uint32_t s = ulist ? ulist[j] : j;
    const double sc = sum_score_lane(p.m, xf, p.state_off[s], p.state_off[s + 1]);
    if (out) out[(uint64_t)lane * out_stride + j] = sc;
    if (amin) amin[(uint64_t)lane * out_stride + j] = 0;       // no density is singled out by the sum
    if (amin16) amin16[(uint64_t)lane * out_stride + j] = 0;
  }
}

// sum score of every frame against its ALIGNED state (Trainer::calc_am_score with max_approx_ == false)
struct SumAlignedParams {
  ExactTables m;
  const uint32_t* state_off;
  const float* x;
  const sgmm_alignment_item* ali;
  uint64_t n_frames;
  uint32_t n_states;
  double* fscore;
  uint32_t* fdens;
};

__global__ void __launch_bounds__(128) k_sum_aligned(SumAlignedParams p) {
  const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= p.n_frames) return;
  uint32_t s = p.ali[f].state;
  if (s >= p.n_states) s = 0;
  p.fscore[f] = sum_score_lane(p.m, p.x + f * p.m.dim, p.state_off[s], p.state_off[s + 1]);
  p.fdens[f] = p.m.Dtot;   // dummy slot: the per-frame density of the max-approx E-step does not exist here
}

}  // namespace
