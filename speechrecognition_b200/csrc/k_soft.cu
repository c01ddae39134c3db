// the NON-max-approx branch of the path.
//   * MixtureModel::sum_score (reference src/sietill/Mixtures.cpp:719-728): score = -log sum_k exp(-density_score_k),
//     densities in mixture order, plain sum starting from 0.0 (no log-sum-exp stabilisation: the reference has none,
//     so a frame far from every density scores +inf exactly as there);
//   * the posterior-weighted branch of MixtureModel::accumulate (Mixtures.cpp:300-357): p_k = exp(-score_k), normalised
//     by their in-order sum, entries below 1e-8 skipped, weight/sum_x/sum_xx += p, p*x, (p*x)*x.
// Density scores are the reference's f64 operation order (sgmm_exact_score), so they are bit-identical; exp / log are
// CUDA's f64 functions (<= 1 ulp from glibc's), hence sum scores and soft statistics agree to ~1e-15 relative, not
// bit for bit (tests demand 1e-12).  FP64-pipe bound; this branch has no f32 filter (every density contributes).
// The (frame, state) emission pass with sum scores is k_emission_f64<true> in k_score.cu.
//
// Soft accumulation is deterministic: frames are stably sorted by aligned state (the radix sort of k_accum.cu), a
// state's frames are cut into items of kSoftItem consecutive frames, one CTA walks an item in frame order keeping
// the state's accumulators [2 dim + 1][densities] in shared memory, then one thread per accumulator adds the items'
// partials in item order.  The summation tree depends only on the data.
#include <algorithm>
#include <initializer_list>

#include "sgmm_internal.cuh"

namespace {

ExactTables exact_tables(const sgmm_ctx* c) {
  ExactTables m{};
  m.meanT = c->d_meanT.p; m.invT = c->d_invT.p; m.mean_tail = c->d_mean_tail.p; m.inv_tail = c->d_inv_tail.p;
  m.norm = c->d_norm.p; m.logw = c->d_logw.p; m.dim = c->dim; m.npair = c->npair; m.Dtot = c->Dtot;
  m.meanR = c->d_meanR.p; m.invU = c->d_invU.p; m.var_of = c->d_var_of.p; m.dimP = c->dimP;
  return m;
}

__device__ __forceinline__ double sum_score_lane(const ExactTables& m, const float* __restrict__ xs, uint32_t k0,
                                                 uint32_t k1) {
  double acc = 0.0;                                                        // Mixtures.cpp:721
  for (uint32_t k = k0; k < k1; k++) acc = __dadd_rn(acc, exp(-sgmm_exact_score(m, xs, k)));   // :723-725
  return -log(acc);                                                        // :727
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


// ---- posterior-weighted accumulation ----------------------------------------------------------------------------
constexpr uint32_t kSoftItem = 2048;     // frames per item (one CTA), walked in sub-blocks of SB frames
constexpr int kSoftThreads = 512;

// keys for the stable sort by aligned state + per-state frame counts (integer atomics: order-independent)
__global__ void __launch_bounds__(256) k_soft_keys(const sgmm_alignment_item* __restrict__ ali, uint64_t n,
                                                   uint32_t n_states, uint32_t Dtot, uint32_t* __restrict__ keys,
                                                   uint32_t* __restrict__ cnt, uint32_t* __restrict__ fdens) {
  const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n) return;
  uint32_t s = ali[f].state;
  if (s >= n_states) s = 0;
  keys[f] = s;
  fdens[f] = Dtot;
  atomicAdd(cnt + s, 1u);
}

// seg_off[s] = first sorted position of state s, item_off[s] = first item of state s (n_states is small)
__global__ void k_soft_offsets(const uint32_t* __restrict__ cnt, uint32_t n_states, uint32_t* __restrict__ seg_off,
                               uint32_t* __restrict__ item_off) {
  if (threadIdx.x || blockIdx.x) return;
  uint32_t a = 0, b = 0;
  for (uint32_t s = 0; s < n_states; s++) {
    seg_off[s] = a; item_off[s] = b;
    a += cnt[s];
    b += (cnt[s] + kSoftItem - 1) / kSoftItem;
  }
  seg_off[n_states] = a; item_off[n_states] = b;
}

struct SoftParams {
  ExactTables m;
  const uint32_t* state_off;
  const float* x;
  const uint32_t* sorted;      // frame ids, stably sorted by aligned state
  const uint32_t* seg_off;
  const uint32_t* item_off;
  uint32_t n_states;
  uint32_t SB;                 // frames per sub-block
  uint32_t maxE;               // partial stride: max densities per state * (2 dim + 1)
  double* partial;             // [item][maxE]
  double* fscore;              // per frame: sum score (-log of the normaliser)
};

__global__ void __launch_bounds__(kSoftThreads, 1) k_soft_partial(SoftParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t item = blockIdx.x;
  if (item >= p.item_off[p.n_states]) return;
  uint32_t lo = 0, hi = p.n_states;          // state owning this item: largest s with item_off[s] <= item
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (p.item_off[mid] <= item) lo = mid; else hi = mid;
  }
  const uint32_t s = lo;
  const uint32_t beg = p.seg_off[s] + (item - p.item_off[s]) * kSoftItem;
  const uint32_t end = min(p.seg_off[s + 1], beg + kSoftItem);
  const uint32_t k0 = p.state_off[s], nd = p.state_off[s + 1] - k0, d = p.m.dim;
  const uint32_t W = 2 * d + 1, Wp = W | 1u, nE = nd * W, ndp = nd | 1u, SB = p.SB;
  const uint32_t maxnd = p.maxE / W, nwords = (maxnd + 31) / 32;
  double* acc = reinterpret_cast<double*>(smem);            // [nd][Wp] (odd row stride: conflict-free column walks)
  double* P = acc + (size_t)maxnd * Wp;                     // [SB][ndp]
  double* X = P + (size_t)SB * (maxnd | 1u);                // [SB][d]
  double* sums = X + (size_t)SB * d;                        // [SB]
  uint32_t* fr = reinterpret_cast<uint32_t*>(sums + SB);    // [SB]
  uint32_t* masks = fr + SB;                                // [SB][nwords]: densities whose posterior is not skipped
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (uint32_t e = tid; e < nd * Wp; e += kSoftThreads) acc[e] = 0.0;

  for (uint32_t i0 = beg; i0 < end; i0 += SB) {
    const uint32_t nf = min(SB, end - i0);
    if (tid < nf) fr[tid] = p.sorted[i0 + tid];
    __syncthreads();
    for (uint32_t q = tid; q < nf * d; q += kSoftThreads) {
      const uint32_t j = q / d, i = q - j * d;
      X[j * d + i] = (double)__ldg(p.x + (uint64_t)fr[j] * d + i);
    }
    // membership_probabilities[k] = exp(-1 * density_score), Mixtures.cpp:310
    for (uint32_t q = tid; q < nf * nd; q += kSoftThreads) {
      const uint32_t j = q / nd, k = q - j * nd;
      P[j * ndp + k] = exp(-sgmm_exact_score(p.m, p.x + (uint64_t)fr[j] * d, k0 + k));
    }
    __syncthreads();
    if (tid < nf) {                       // std::accumulate(..., 0.0) in density order, :319-320
      double a = 0.0;
      for (uint32_t k = 0; k < nd; k++) a = __dadd_rn(a, P[tid * ndp + k]);
      sums[tid] = a;
      p.fscore[fr[tid]] = -log(a);        // == sum_score(x, s), :719-728 (same terms, same order)
    }
    __syncthreads();
    // divides<double> (:322-323) and the skip test (:336-338: "< 1e-8" — a NaN weight is NOT skipped, as in the
    // reference): warp = frame, one ballot per 32 densities
    for (uint32_t j = warp; j < nf; j += kSoftThreads / 32) {
      for (uint32_t r = 0; r < nwords; r++) {
        const uint32_t k = 32 * r + lane;
        bool keep = false;
        if (k < nd) {
          const double w = __ddiv_rn(P[j * ndp + k], sums[j]);
          P[j * ndp + k] = w;
          keep = !(w < 1e-8);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) masks[j * nwords + r] = m;
      }
    }
    __syncthreads();
    // one task per (accumulator column c, word of 32 densities): frames in order, only the kept densities — a frame's
    // posterior mass sits on one or two of them
    for (uint32_t idx = tid; idx < W * nwords; idx += kSoftThreads) {
      const uint32_t wi = idx / W, c = idx - wi * W;
      for (uint32_t j = 0; j < nf; j++) {
        uint32_t m = masks[j * nwords + wi];
        while (m) {
          const uint32_t k = 32 * wi + (uint32_t)__ffs(m) - 1u;
          m &= m - 1;
          const double w = P[j * ndp + k];
          double v = acc[k * Wp + c];
          if (c < d) v = __dadd_rn(v, __dmul_rn(w, X[j * d + c]));                       // x + scale * y, :46-56
          else if (c < 2 * d) { const double y = X[j * d + (c - d)]; v = __dadd_rn(v, __dmul_rn(__dmul_rn(w, y), y)); }   // :58-68
          else v = __dadd_rn(v, w);                                                       // weights, :343-344
          acc[k * Wp + c] = v;
        }
      }
    }
    __syncthreads();
  }
  double* out = p.partial + (size_t)item * p.maxE;
  for (uint32_t e = tid; e < nE; e += kSoftThreads) {       // partial layout [column][density]
    const uint32_t c = e / nd, k = e - c * nd;
    out[e] = acc[k * Wp + c];
  }
}

// one thread per (density, accumulator column): add the items' partials in item order
__global__ void __launch_bounds__(256) k_soft_final(const double* __restrict__ partial, uint32_t maxE,
                                                    const uint32_t* __restrict__ item_off,
                                                    const uint32_t* __restrict__ state_off, uint32_t n_states,
                                                    uint32_t Dtot, uint32_t dim, double* __restrict__ accum) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t W = 2 * dim + 1;
  if (t >= (uint64_t)Dtot * W) return;
  const uint32_t j = (uint32_t)(t / W), c = (uint32_t)(t - (uint64_t)j * W);
  uint32_t lo = 0, hi = n_states;            // state of density j: largest s with state_off[s] <= j
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (state_off[mid] <= j) lo = mid; else hi = mid;
  }
  const uint32_t s = lo, nd = state_off[s + 1] - state_off[s], k = j - state_off[s];
  double a = 0.0;
  for (uint32_t it = item_off[s]; it < item_off[s + 1]; it++)
    a = __dadd_rn(a, partial[(size_t)it * maxE + (size_t)c * nd + k]);
  double* count = accum;
  double* sum_x = accum + Dtot;
  double* sum_xx = accum + Dtot + (size_t)Dtot * dim;
  if (c < dim) sum_x[(size_t)j * dim + c] = a;
  else if (c < 2 * dim) sum_xx[(size_t)j * dim + (c - dim)] = a;
  else count[j] = a;
}

}  // namespace

int sgmm_launch_sum_aligned(sgmm_ctx* c) {
  if (c->n_frames == 0) return 0;
  SumAlignedParams p{};
  p.m = exact_tables(c);
  p.state_off = c->d_state_off.p;
  p.x = c->d_x.p; p.ali = c->d_ali.p; p.n_frames = c->n_frames; p.n_states = c->n_states;
  p.fscore = c->d_fscore.p; p.fdens = c->d_fdens.p;
  k_sum_aligned<<<(unsigned)((c->n_frames + 127) / 128), 128, 0, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}

// d_acc[count | sum_x | sum_xx] = posterior-weighted statistics of the resident alignment; d_fscore = sum scores
int sgmm_launch_accumulate_soft(sgmm_ctx* c) {
  const uint64_t n = c->n_frames;
  const uint32_t N = c->n_states, D = c->Dtot, dim = c->dim, W = 2 * dim + 1;
  const uint32_t maxE = c->max_dens * W;
  uint32_t SB = 0;
  size_t smem = 0;
  for (uint32_t sb : {32u, 16u, 8u}) {
    const size_t need = 8 * ((size_t)c->max_dens * (W | 1u) + (size_t)sb * (c->max_dens | 1u) + (size_t)sb * dim + sb) +
                        4 * (size_t)sb * (1 + (c->max_dens + 31) / 32);
    if (need <= 227 * 1024) { SB = sb; smem = need; break; }
  }
  if (!SB) SGMM_FAIL(c, "accumulate_soft: a state's accumulators (densities x (2 dim + 1) doubles) exceed shared memory");
  SGMM_CHECK(c, c->d_cnt_u32.reserve(std::max<size_t>(D + 2, N + 1)));
  SGMM_CHECK(c, c->d_seg_off.reserve(std::max<size_t>(D + 2, N + 1)));
  SGMM_CHECK(c, c->d_item_off.reserve(std::max<size_t>(D + 1, N + 1)));
  const size_t max_items = n / kSoftItem + N + 1;
  SGMM_CHECK(c, c->d_partial.reserve(max_items * (size_t)maxE));
  SGMM_CHECK(c, cudaMemsetAsync(c->d_cnt_u32.p, 0, (N + 1) * sizeof(uint32_t), c->stream));
  SGMM_CHECK(c, c->d_soft_keys.reserve(n));   // the sort ping-pongs through d_keys[]: its input lives elsewhere
  uint32_t* keys = c->d_soft_keys.p;
  if (n) {
    k_soft_keys<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d_ali.p, n, N, D, keys, c->d_cnt_u32.p, c->d_fdens.p);
    SGMM_LAUNCH_CHECK(c);
  }
  const uint32_t* sorted = nullptr;
  if (sgmm_sort_frames(c, keys, N, &sorted)) return 1;
  StageTimer st(c, SGMM_T_REDUCE);
  k_soft_offsets<<<1, 32, 0, c->stream>>>(c->d_cnt_u32.p, N, c->d_seg_off.p, c->d_item_off.p);
  SGMM_LAUNCH_CHECK(c);
  if (n) {
    SoftParams p{};
    p.m = exact_tables(c);
    p.state_off = c->d_state_off.p; p.x = c->d_x.p; p.sorted = sorted;
    p.seg_off = c->d_seg_off.p; p.item_off = c->d_item_off.p; p.n_states = N; p.SB = SB; p.maxE = maxE;
    p.partial = c->d_partial.p; p.fscore = c->d_fscore.p;
    SGMM_CHECK(c, cudaFuncSetAttribute(k_soft_partial, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_soft_partial<<<(unsigned)max_items, kSoftThreads, smem, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
  }
  const uint64_t tot = (uint64_t)D * W;
  if (tot) {
    k_soft_final<<<(unsigned)((tot + 255) / 256), 256, 0, c->stream>>>(c->d_partial.p, maxE, c->d_item_off.p,
                                                                      c->d_state_off.p, N, D, dim, c->d_acc.p);
    SGMM_LAUNCH_CHECK(c);
  }
  return 0;
}
