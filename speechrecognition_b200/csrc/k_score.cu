// K1 (exact form) and K3: diagonal-covariance GMM scoring with the maximum approximation, in f64 and in
// the reference's exact operation order, so that scores and argmins are bit-identical to
// MixtureModel::density_score_sse / min_score (reference src/sietill/Mixtures.cpp:645-713).
//
// Layout: model tables are dim-pair-major double2 (tabT[ip * Dtot + j]) — the SSE2 reference keeps two
// f64 partial sums (even dims / odd dims), which is exactly one double2 lane pair here.  A warp scores
// 32 consecutive frames (one per lane) against a list of states; table reads are warp-uniform
// (one L1 wavefront each), features sit in shared memory as f64.  Bound by the FP64 pipe:
// 4 DP ops per (frame, density, dim) — sub, mul, mul, add; no FMA contraction, the reference has none.
#include "sgmm_internal.cuh"

namespace {

struct ModelView {
  const double2* meanT;
  const double2* invT;
  const double* mean_tail;
  const double* inv_tail;
  const double* norm;
  const double* logw;
  const uint32_t* state_off;
  uint32_t dim, npair, Dtot;
};

struct EmisParams {
  ModelView m;
  const float* x;
  uint64_t n_tasks;
  // gathered (per-utterance) mode
  const uint4* tasks;
  const uint64_t* frame_off;
  const uint32_t* uniq;
  const uint32_t* uniq_off;
  const uint32_t* nuniq_pad;
  const uint64_t* E_off;
  double* E;
  uint16_t* Amin;
  // dense mode
  int dense;
  uint64_t f0, nframes;
  uint32_t n_states;
  double* scores;
  uint32_t* argmin;
};

// Scores NK consecutive densities k .. k+NK-1 for the calling lane's frame (NK independent f64 chains
// for ILP; each chain keeps the reference's order).  xs: this warp's features, xs[ip * stride + lane].
template <int NK>
__device__ __forceinline__ void score_densities(const ModelView& m, uint32_t k, const double2* __restrict__ xs,
                                                int xstride, double xtail, double (&sc)[NK]) {
  double l0[NK], l1[NK];
#pragma unroll
  for (int q = 0; q < NK; q++) { l0[q] = 0.0; l1[q] = 0.0; }
  const double2* mp = m.meanT + k;
  const double2* vp = m.invT + k;
#pragma unroll 2
  for (uint32_t ip = 0; ip < m.npair; ip++) {
    const double2 xx = xs[ip * xstride];
#pragma unroll
    for (int q = 0; q < NK; q++) {
      const double2 mu = __ldg(mp + q);
      const double2 iv = __ldg(vp + q);
      double a = __dsub_rn(xx.x, mu.x);
      double b = __dsub_rn(xx.y, mu.y);
      a = __dmul_rn(a, a);
      b = __dmul_rn(b, b);
      a = __dmul_rn(a, iv.x);
      b = __dmul_rn(b, iv.y);
      l0[q] = __dadd_rn(l0[q], a);
      l1[q] = __dadd_rn(l1[q], b);
    }
    mp += m.Dtot;
    vp += m.Dtot;
  }
#pragma unroll
  for (int q = 0; q < NK; q++) {
    double dist = __dadd_rn(l0[q], l1[q]);                       // add_double_sse_register, Mixtures.cpp:633-639
    if (m.dim & 1) {                                             // odd tail, Mixtures.cpp:672-675
      const double c = __dsub_rn(xtail, __ldg(m.mean_tail + k + q));
      dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), __ldg(m.inv_tail + k + q)));
    }
    double s = __dadd_rn(__ldg(m.norm + k + q), __dmul_rn(dist, 0.5));  // norm + dist/2, Mixtures.cpp:679
    sc[q] = __dsub_rn(s, __ldg(m.logw + k + q));                        // -= log weight, Mixtures.cpp:680
  }
}

// min_score (Mixtures.cpp:696-713): start (1e10, 0), strict '<' in density order.
__device__ __forceinline__ void min_over_state(const ModelView& m, uint32_t k0, uint32_t k1,
                                               const double2* __restrict__ xs, int xstride, double xtail,
                                               double& best, uint32_t& best_k) {
  best = 1e10;
  best_k = 0;
  uint32_t k = k0;
  for (; k + 2 <= k1; k += 2) {
    double sc[2];
    score_densities<2>(m, k, xs, xstride, xtail, sc);
    if (sc[0] < best) { best = sc[0]; best_k = k - k0; }
    if (sc[1] < best) { best = sc[1]; best_k = k + 1 - k0; }
  }
  if (k < k1) {
    double sc[1];
    score_densities<1>(m, k, xs, xstride, xtail, sc);
    if (sc[0] < best) { best = sc[0]; best_k = k - k0; }
  }
}

// sum_score (Mixtures.cpp:719-728): acc += exp(-score_k) in density order from 0.0, result -log(acc).  Same score
// chains as min_over_state (two densities in flight); the additions stay in mixture order.
__device__ __forceinline__ double sum_over_state(const ModelView& m, uint32_t k0, uint32_t k1,
                                                 const double2* __restrict__ xs, int xstride, double xtail) {
  double acc = 0.0;
  uint32_t k = k0;
  for (; k + 2 <= k1; k += 2) {
    double sc[2];
    score_densities<2>(m, k, xs, xstride, xtail, sc);
    const double e0 = exp(-sc[0]), e1 = exp(-sc[1]);
    acc = __dadd_rn(__dadd_rn(acc, e0), e1);
  }
  if (k < k1) {
    double sc[1];
    score_densities<1>(m, k, xs, xstride, xtail, sc);
    acc = __dadd_rn(acc, exp(-sc[0]));
  }
  return -log(acc);
}

// Cooperative, coalesced load of `nvalid` consecutive frames into this warp's shared f64 tile.
__device__ __forceinline__ void load_frames(const float* __restrict__ x, uint64_t fbase, uint32_t nvalid,
                                            uint32_t dim, uint32_t npair, double2* xs, double* xt, int lane) {
  const float* src = x + fbase * dim;
  const uint32_t total = nvalid * dim;
  double* xs_d = reinterpret_cast<double*>(xs);
  for (uint32_t idx = lane; idx < total; idx += 32) {
    const uint32_t r = idx / dim, i = idx - r * dim;
    const double v = (double)__ldg(src + idx);
    if ((i >> 1) < npair) xs_d[((i >> 1) * 32 + r) * 2 + (i & 1)] = v;
    else xt[r] = v;
  }
  __syncwarp();
}

constexpr int kEmisWarps = 4;
__host__ __device__ inline size_t emis_smem_per_warp(uint32_t npair) {
  return (size_t)npair * 32 * sizeof(double2) + 32 * sizeof(double) + 32 * 33 * sizeof(double);
}

template <bool SUM>
__global__ void __launch_bounds__(kEmisWarps * 32) k_emission_f64(EmisParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = (uint64_t)blockIdx.x * kEmisWarps + warp;
  if (task >= p.n_tasks) return;
  unsigned char* base = smem + warp * emis_smem_per_warp(p.m.npair);
  double2* xs = reinterpret_cast<double2*>(base);
  double* xt = reinterpret_cast<double*>(base + (size_t)p.m.npair * 32 * sizeof(double2));
  double* stage = xt + 32;

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
    out = p.E + p.E_off[t.x] + (uint64_t)t.y * out_stride;       // t.y is a multiple of 16: first cell of its tile
    amin16 = p.Amin + p.E_off[t.x] + (uint64_t)t.y * out_stride;
  }
  load_frames(p.x, fbase, nvalid, p.m.dim, p.m.npair, xs, xt, lane);
  const double xtail = xt[lane];
  const double2* my_xs = xs + lane;

  for (uint32_t j0 = 0; j0 < nu; j0 += 32) {
    const uint32_t jn = min(32u, nu - j0);
    for (uint32_t jj = 0; jj < jn; jj++) {
      const uint32_t s = ulist ? ulist[j0 + jj] : j0 + jj;
      double best;
      uint32_t bk = 0;
      if (SUM) best = sum_over_state(p.m, p.m.state_off[s], p.m.state_off[s + 1], my_xs, 32, xtail);
      else min_over_state(p.m, p.m.state_off[s], p.m.state_off[s + 1], my_xs, 32, xtail, best, bk);
      if (amin16) {
        // gathered: the tiled layout (sgmm_cell_rel) runs along the frames, lane = frame stores straight away
        if (lane < (int)nvalid) {
          const uint64_t cell = sgmm_cell_rel(out_stride, lane, j0 + jj);
          out[cell] = best;
          amin16[cell] = (uint16_t)bk;
        }
        continue;
      }
      stage[lane * 33 + jj] = best;
      if (amin && lane < (int)nvalid) amin[(uint64_t)lane * out_stride + j0 + jj] = bk;
    }
    if (amin16) continue;
    __syncwarp();
    if (out && lane < (int)jn)
      for (uint32_t r = 0; r < nvalid; r++) out[(uint64_t)r * out_stride + j0 + lane] = stage[r * 33 + lane];
    __syncwarp();
  }
}

// K3: every frame against its ALIGNED state only (MixtureModel::accumulate's call to min_score,
// Mixtures.cpp:296-298, and Trainer::calc_am_score, Training.cpp:605).  One lane per frame; lanes of a
// warp usually share a state (alignments are piecewise constant), so table reads mostly coalesce.
struct RescoreParams {
  ModelView m;
  const float* x;
  const sgmm_alignment_item* ali;
  uint64_t n_frames;
  uint32_t n_states;
  int first_pass;
  uint32_t* fdens;     // global density id per frame
  double* fscore;      // min score per frame
  uint32_t* cnt;       // per-density frame count (integer atomics: order-independent)
};

constexpr int kRescoreWarps = 4;
__global__ void __launch_bounds__(kRescoreWarps * 32) k_rescore_f64(RescoreParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t r0 = ((uint64_t)blockIdx.x * kRescoreWarps + warp) * 32;
  if (r0 >= p.n_frames) return;
  const uint32_t nvalid = (uint32_t)min((uint64_t)32, p.n_frames - r0);
  const uint64_t f = r0 + lane;
  const bool valid = lane < (int)nvalid;
  uint32_t s = valid ? p.ali[f].state : 0;
  if (s >= p.n_states) s = 0;  // guarded on the host; never index out of the table
  const uint32_t k0 = p.m.state_off[s], k1 = p.m.state_off[s + 1];
  double best = 0.0;
  uint32_t bk = 0;
  if (!p.first_pass) {
    unsigned char* base = smem + warp * ((size_t)p.m.npair * 32 * sizeof(double2) + 32 * sizeof(double));
    double2* xs = reinterpret_cast<double2*>(base);
    double* xt = reinterpret_cast<double*>(base + (size_t)p.m.npair * 32 * sizeof(double2));
    load_frames(p.x, r0, nvalid, p.m.dim, p.m.npair, xs, xt, lane);
    if (valid) min_over_state(p.m, k0, k1, xs + lane, 32, xt[lane], best, bk);
  }
  if (valid) {
    // A state whose mixture is empty (every density eliminated) contributes nothing: the reference's loop
    // over mixtures_[m] (Mixtures.cpp:333) runs zero times.  Such frames go to the dummy slot Dtot.
    const uint32_t gid = (k1 > k0) ? k0 + bk : p.m.Dtot;
    if (k1 == k0 && !p.first_pass) best = 1e10;   // min_score of an empty mixture, Mixtures.cpp:699
    p.fdens[f] = gid;
    p.fscore[f] = best;
    atomicAdd(p.cnt + gid, 1u);
  }
}

// K3 when the model is the one the resident alignment was computed with: the alignment pass has already
// scored every (frame, state-of-its-utterance) cell, so the frame's min score and argmin density are a lookup.
struct CachedParams {
  const sgmm_alignment_item* ali;
  const uint16_t* fcol;
  const double* E;
  const uint16_t* Amin;
  const uint64_t* frame_off;
  const uint64_t* E_off;
  const uint32_t* nuniq_pad;
  const uint32_t* state_off;
  uint32_t n_utts, Dtot;
  uint64_t n_frames;
  uint32_t* fdens;
  double* fscore;
  uint32_t* cnt;
};

__global__ void __launch_bounds__(256) k_rescore_cached(CachedParams p) {
  const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= p.n_frames) return;
  uint32_t lo = 0, hi = p.n_utts;            // utterance of frame f: largest u with frame_off[u] <= f
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (p.frame_off[mid] <= f) lo = mid; else hi = mid;
  }
  const uint64_t e = p.E_off[lo] + sgmm_cell_rel(p.nuniq_pad[lo], (uint32_t)(f - p.frame_off[lo]), p.fcol[f]);
  const uint32_t s = p.ali[f].state;
  const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
  const uint32_t gid = (k1 > k0) ? k0 + p.Amin[e] : p.Dtot;
  p.fdens[f] = gid;
  p.fscore[f] = p.E[e];
  atomicAdd(p.cnt + gid, 1u);
}

// K3 of the tolerance mode (option "tolerance"): the alignment came from f32 scores, but the tensor filter's VERDICTS of
// that pass are still in place — for every (frame, state) cell the certified f64 argmin, or "uncertain".  One thread
// per frame looks its aligned cell up: a certified winner is evaluated in f64 (the reference's operation order, one
// density); an uncertain cell (0.25 %) is scanned by the whole warp, lanes = densities, first minimum wins.  The E-step
// therefore sees the exact min_score / argmin of Mixtures.cpp:296-298 for the alignment it is given.
struct VerdictParams {
  CachedParams c;
  const uint32_t* cand;
  const float* x;
  ExactTables m;
};

__global__ void __launch_bounds__(256) k_rescore_verdict(VerdictParams p) {
  const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31;
  const bool valid = f < p.c.n_frames;
  uint32_t k0 = 0, k1 = 0, v = 0;
  if (valid) {
    uint32_t lo = 0, hi = p.c.n_utts;            // utterance of frame f: largest u with frame_off[u] <= f
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (p.c.frame_off[mid] <= f) lo = mid; else hi = mid;
    }
    const uint64_t e = p.c.E_off[lo] + sgmm_cell_rel(p.c.nuniq_pad[lo], (uint32_t)(f - p.c.frame_off[lo]), p.c.fcol[f]);
    const uint32_t s = p.c.ali[f].state;
    k0 = p.c.state_off[s]; k1 = p.c.state_off[s + 1];
    v = p.cand[e];
  }
  // a pair verdict (sgmm_tc::kPair | a | b << 15): the argmin is one of two columns
  const bool pair = valid && k1 > k0 && (v & 0xc0000000u) == 0x40000000u && (v & 0x7fffu) < k1 - k0 && ((v >> 15) & 0x7fffu) < k1 - k0;
  const bool unc = valid && k1 > k0 && !(v < k1 - k0) && !pair;
  double best = 1e10;                            // min_score's start value, Mixtures.cpp:699
  uint32_t bk = 0;
  if (valid && k1 > k0 && !unc) {
    const uint32_t ka = pair ? (v & 0x7fffu) : v;
    const double sc = sgmm_exact_score_rm(p.m, p.x + f * p.m.dim, k0 + ka);
    if (sc < best) { best = sc; bk = ka; }
    if (pair) {
      const uint32_t kb = (v >> 15) & 0x7fffu;     // kb > ka: on equal scores the lower column stays
      const double sb = sgmm_exact_score_rm(p.m, p.x + f * p.m.dim, k0 + kb);
      if (sb < best) { best = sb; bk = kb; }
    }
  }
  uint32_t todo = __ballot_sync(0xffffffffu, unc);
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const uint64_t fs = __shfl_sync(0xffffffffu, f, src);
    const uint32_t a0 = __shfl_sync(0xffffffffu, k0, src), a1 = __shfl_sync(0xffffffffu, k1, src);
    double rb = 1e10;
    uint32_t rk = 0xffffffffu;
    for (uint32_t k = a0 + lane; k < a1; k += 32) {
      const double sc = sgmm_exact_score_rm(p.m, p.x + fs * p.m.dim, k);
      if (sc < rb) { rb = sc; rk = k - a0; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, rb, o);
      const uint32_t ok = __shfl_xor_sync(0xffffffffu, rk, o);
      if (ob < rb || (ob == rb && ok < rk)) { rb = ob; rk = ok; }
    }
    if ((int)lane == src) { best = rb; bk = rk == 0xffffffffu ? 0u : rk; }
  }
  if (valid) {
    const uint32_t gid = (k1 > k0) ? k0 + bk : p.c.Dtot;
    p.c.fdens[f] = gid;
    p.c.fscore[f] = best;
    atomicAdd(p.c.cnt + gid, 1u);
  }
}

ModelView model_view(const sgmm_ctx* c) {
  ModelView m;
  m.meanT = c->d_meanT.p; m.invT = c->d_invT.p; m.mean_tail = c->d_mean_tail.p; m.inv_tail = c->d_inv_tail.p;
  m.norm = c->d_norm.p; m.logw = c->d_logw.p; m.state_off = c->d_state_off.p;
  m.dim = c->dim; m.npair = c->npair; m.Dtot = c->Dtot;
  return m;
}

}  // namespace

static int launch_emission_f64(sgmm_ctx* c, bool sum, const uint4* d_tasks, uint64_t n_tasks, bool dense,
                               uint64_t dense_f0, uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin) {
  if (n_tasks == 0) return 0;
  EmisParams p{};
  p.m = model_view(c);
  p.x = c->d_x.p;
  p.n_tasks = n_tasks;
  p.tasks = d_tasks;
  p.frame_off = c->d_frame_off.p; p.uniq = c->d_uniq.p; p.uniq_off = c->d_uniq_off.p;
  p.nuniq_pad = c->d_nuniq_pad.p; p.E_off = c->d_E_off.p; p.E = c->d_E.p; p.Amin = c->d_Amin.p;
  p.dense = dense ? 1 : 0; p.f0 = dense_f0; p.nframes = dense_nframes; p.n_states = c->n_states;
  p.scores = d_scores; p.argmin = d_argmin;
  const size_t smem = kEmisWarps * emis_smem_per_warp(c->npair);
  SGMM_CHECK(c, cudaFuncSetAttribute(k_emission_f64<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SGMM_CHECK(c, cudaFuncSetAttribute(k_emission_f64<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const uint64_t grid = (n_tasks + kEmisWarps - 1) / kEmisWarps;
  if (sum) k_emission_f64<true><<<(unsigned)grid, kEmisWarps * 32, smem, c->stream>>>(p);
  else k_emission_f64<false><<<(unsigned)grid, kEmisWarps * 32, smem, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

int sgmm_launch_emission(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                         uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin) {
  return launch_emission_f64(c, false, d_tasks, n_tasks, dense, dense_f0, dense_nframes, d_scores, d_argmin);
}

// score() = sum_score (option max_approx = 0): the same kernel with the sum reduction over a state's densities
int sgmm_launch_emission_sum(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                             uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin) {
  return launch_emission_f64(c, true, d_tasks, n_tasks, dense, dense_f0, dense_nframes, d_scores, d_argmin);
}

int sgmm_launch_rescore(sgmm_ctx* c, bool first_pass) {
  if (c->n_frames == 0) return 0;
  RescoreParams p{};
  p.m = model_view(c);
  p.x = c->d_x.p; p.ali = c->d_ali.p; p.n_frames = c->n_frames; p.n_states = c->n_states;
  p.first_pass = first_pass ? 1 : 0;
  p.fdens = c->d_fdens.p; p.fscore = c->d_fscore.p; p.cnt = c->d_cnt_u32.p;
  const size_t smem = kRescoreWarps * ((size_t)c->npair * 32 * sizeof(double2) + 32 * sizeof(double));
  SGMM_CHECK(c, cudaFuncSetAttribute(k_rescore_f64, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const uint64_t warps = (c->n_frames + 31) / 32;
  const uint64_t grid = (warps + kRescoreWarps - 1) / kRescoreWarps;
  k_rescore_f64<<<(unsigned)grid, kRescoreWarps * 32, smem, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

int sgmm_launch_rescore_verdict(sgmm_ctx* c) {
  if (c->n_frames == 0) return 0;
  VerdictParams p{};
  p.c.ali = c->d_ali.p; p.c.fcol = c->d_fcol.p;
  p.c.frame_off = c->d_frame_off.p; p.c.E_off = c->d_E_off.p; p.c.nuniq_pad = c->d_nuniq_pad.p;
  p.c.state_off = c->d_state_off.p; p.c.n_utts = c->n_utts; p.c.Dtot = c->Dtot; p.c.n_frames = c->n_frames;
  p.c.fdens = c->d_fdens.p; p.c.fscore = c->d_fscore.p; p.c.cnt = c->d_cnt_u32.p;
  p.cand = c->d_ci1.p; p.x = c->d_x.p;
  p.m.meanT = c->d_meanT.p; p.m.invT = c->d_invT.p; p.m.mean_tail = c->d_mean_tail.p; p.m.inv_tail = c->d_inv_tail.p;
  p.m.norm = c->d_norm.p; p.m.logw = c->d_logw.p; p.m.dim = c->dim; p.m.npair = c->npair; p.m.Dtot = c->Dtot;
  p.m.meanR = c->d_meanR.p; p.m.invU = c->d_invU.p; p.m.var_of = c->d_var_of.p; p.m.dimP = c->dimP;
  k_rescore_verdict<<<(unsigned)((c->n_frames + 255) / 256), 256, 0, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

int sgmm_launch_rescore_cached(sgmm_ctx* c) {
  if (c->n_frames == 0) return 0;
  CachedParams p{};
  p.ali = c->d_ali.p; p.fcol = c->d_fcol.p; p.E = c->d_E.p; p.Amin = c->d_Amin.p;
  p.frame_off = c->d_frame_off.p; p.E_off = c->d_E_off.p; p.nuniq_pad = c->d_nuniq_pad.p;
  p.state_off = c->d_state_off.p; p.n_utts = c->n_utts; p.Dtot = c->Dtot; p.n_frames = c->n_frames;
  p.fdens = c->d_fdens.p; p.fscore = c->d_fscore.p; p.cnt = c->d_cnt_u32.p;
  k_rescore_cached<<<(unsigned)((c->n_frames + 255) / 256), 256, 0, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
