// K1, FP32-FFMA form with certified refinement ("filter in f32, decide in f64").
//
// For every (frame, state) the reference needs  min_k / argmin_k  of the f64 density scores
// (Mixtures.cpp:645-713).  This kernel finds the two smallest f32 scores with FFMA-pipe arithmetic
// (direct form  c_k + sum_i h_ki (x_i - mu_ki)^2,  h = 1/(2 sigma^2); every term is non-negative, so there is
// no cancellation), then
//   * if  best2 - best1  exceeds a rigorous bound on the f32 evaluation error of both candidates, the f32
//     argmin IS the f64 argmin: its score is re-evaluated in f64 in the reference's exact operation order
//     (one density instead of all of them), giving the bit-identical reference score;
//   * otherwise (a near-tie, probability ~1e-4 per cell) the lane rescans the state's densities in f64,
//     exactly as the reference does.
// Either way the outputs are bit-identical to min_score(); the f32 pass only decides where f64 work is spent.
//
// Error bound (standard forward analysis, u = 2^-24, all terms >= 0): for density k with
// dist = sum_i h_i t_i^2, M_k = sum_i h_i mu_i^2:
//   |score32 - score64| <= u * ((d + 8) * dist + 2 * sqrt(dist * M_k) + |c_k|)   (+ higher order)
// (mu and h are rounded to f32: |delta t| <= u (|mu| + |t|); Cauchy-Schwarz on sum h |t| |mu|).  The kernel
// uses eps = 2^-22 (4x slack) and the per-state maxima of M_k and |c_k|.
//
// Work split: one warp scores 64 frames (2 per lane, features in registers) against the density tables of
// its task's state list; tables are staged through warp-private shared memory in chunks (broadcast LDS).
#include "sgmm_internal.cuh"

namespace sgmm_ffma {

constexpr int kWarps = 4;
constexpr int kChunk = 8;   // densities per staged chunk
constexpr int R = 2;        // frames per lane

struct Params {
  // f32 tables, [D][4*DV4] (zero padded), and per-density constant
  const float4* mu32;
  const float4* h32;
  const float* c32;
  // per-state guard constants {cmin, cabs_max, M_max, unused}
  const float4* sguard;
  // exact f64 tables (same as k_score.cu)
  const double2* meanT;
  const double2* invT;
  const double* mean_tail;
  const double* inv_tail;
  const double* norm;
  const double* logw;
  const uint32_t* state_off;
  uint32_t dim, npair, Dtot, n_states;
  const float* x;
  uint64_t n_tasks;
  // gathered mode
  const uint4* tasks;          // {utt, t0, _, _}, t0 multiple of 64
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
  double* scores;
  uint32_t* argmin;
  unsigned long long* fallback_count;   // diagnostics: cells that needed the f64 rescan
};

// Exact f64 density score, reference order (same arithmetic as k_score.cu::score_densities<1>), with the
// frame's features read from global memory (f32 -> f64 is exact).
__device__ __forceinline__ double exact_score(const Params& p, const float* __restrict__ xf, uint32_t k) {
  double l0 = 0.0, l1 = 0.0;
  const double2* mp = p.meanT + k;
  const double2* vp = p.invT + k;
  for (uint32_t ip = 0; ip < p.npair; ip++) {
    const double2 mu = __ldg(mp), iv = __ldg(vp);
    double a = __dsub_rn((double)__ldg(xf + 2 * ip), mu.x);
    double b = __dsub_rn((double)__ldg(xf + 2 * ip + 1), mu.y);
    a = __dmul_rn(a, a); b = __dmul_rn(b, b);
    a = __dmul_rn(a, iv.x); b = __dmul_rn(b, iv.y);
    l0 = __dadd_rn(l0, a); l1 = __dadd_rn(l1, b);
    mp += p.Dtot; vp += p.Dtot;
  }
  double dist = __dadd_rn(l0, l1);
  if (p.dim & 1) {
    const double c = __dsub_rn((double)__ldg(xf + p.dim - 1), __ldg(p.mean_tail + k));
    dist = __dadd_rn(dist, __dmul_rn(__dmul_rn(c, c), __ldg(p.inv_tail + k)));
  }
  const double s = __dadd_rn(__ldg(p.norm + k), __dmul_rn(dist, 0.5));
  return __dsub_rn(s, __ldg(p.logw + k));
}

template <int DV4>
__global__ void __launch_bounds__(kWarps * 32) k_emission_ffma(Params p) {
  __shared__ float4 s_mu[kWarps][kChunk][DV4];
  __shared__ float4 s_h[kWarps][kChunk][DV4];
  __shared__ float s_c[kWarps][kChunk];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t task = (uint64_t)blockIdx.x * kWarps + warp;
  if (task >= p.n_tasks) return;

  uint64_t fbase;
  uint32_t nvalid, nu, out_stride;
  const uint32_t* ulist = nullptr;
  double* out;
  uint32_t* amin = nullptr;
  uint16_t* amin16 = nullptr;
  if (p.dense) {
    const uint64_t r0 = task * (32 * R);
    fbase = p.f0 + r0;
    nvalid = (uint32_t)min((uint64_t)(32 * R), p.nframes - r0);
    nu = p.n_states;
    out_stride = p.n_states;
    out = p.scores ? p.scores + r0 * p.n_states : nullptr;
    amin = p.argmin ? p.argmin + r0 * p.n_states : nullptr;
  } else {
    const uint4 t = p.tasks[task];
    const uint64_t uf0 = p.frame_off[t.x];
    const uint32_t T = (uint32_t)(p.frame_off[t.x + 1] - uf0);
    fbase = uf0 + t.y;
    nvalid = min((uint32_t)(32 * R), T - t.y);
    ulist = p.uniq + p.uniq_off[t.x];
    nu = p.uniq_off[t.x + 1] - p.uniq_off[t.x];
    out_stride = p.nuniq_pad[t.x];
    out = p.E + p.E_off[t.x] + (uint64_t)t.y * out_stride;
    amin16 = p.Amin + p.E_off[t.x] + (uint64_t)t.y * out_stride;
  }

  // features of this lane's R frames in registers (zero padded to 4*DV4)
  float xr[R][4 * DV4];
  bool valid[R];
#pragma unroll
  for (int r = 0; r < R; r++) {
    const uint32_t row = r * 32 + lane;
    valid[r] = row < nvalid;
    const float* xf = p.x + (fbase + (valid[r] ? row : 0)) * p.dim;
#pragma unroll
    for (int i = 0; i < 4 * DV4; i++) xr[r][i] = (i < (int)p.dim) ? __ldg(xf + i) : 0.f;
  }

  const float eps = 2.384185791015625e-07f;   // 2^-22
  const float dterm = (float)(p.dim + 8);
  unsigned long long n_fallback = 0;

  for (uint32_t j = 0; j < nu; j++) {
    const uint32_t s = ulist ? ulist[j] : j;
    const uint32_t k0 = p.state_off[s], k1 = p.state_off[s + 1];
    float b1[R], b2[R];
    uint32_t i1[R];
#pragma unroll
    for (int r = 0; r < R; r++) { b1[r] = INFINITY; b2[r] = INFINITY; i1[r] = 0; }
    for (uint32_t kc = k0; kc < k1; kc += kChunk) {
      const uint32_t nk = min((uint32_t)kChunk, k1 - kc);
      __syncwarp();
      for (uint32_t q = lane; q < nk * DV4; q += 32) {
        const uint32_t dk = q / DV4, v = q - dk * DV4;
        s_mu[warp][dk][v] = __ldg(p.mu32 + (size_t)(kc + dk) * DV4 + v);
        s_h[warp][dk][v] = __ldg(p.h32 + (size_t)(kc + dk) * DV4 + v);
      }
      if (lane < (int)nk) s_c[warp][lane] = __ldg(p.c32 + kc + lane);
      __syncwarp();
      for (uint32_t dk = 0; dk < nk; dk++) {
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
#pragma unroll
        for (int v = 0; v < DV4; v++) {
          const float4 m = s_mu[warp][dk][v];
          const float4 h = s_h[warp][dk][v];
#pragma unroll
          for (int r = 0; r < R; r++) {
            float t;
            t = xr[r][4 * v + 0] - m.x; acc[r] = fmaf(t * h.x, t, acc[r]);
            t = xr[r][4 * v + 1] - m.y; acc[r] = fmaf(t * h.y, t, acc[r]);
            t = xr[r][4 * v + 2] - m.z; acc[r] = fmaf(t * h.z, t, acc[r]);
            t = xr[r][4 * v + 3] - m.w; acc[r] = fmaf(t * h.w, t, acc[r]);
          }
        }
        const float c = s_c[warp][dk];
        const uint32_t kk = kc + dk - k0;
#pragma unroll
        for (int r = 0; r < R; r++) {
          const float sc = acc[r] + c;
          b2[r] = fminf(b2[r], fmaxf(b1[r], sc));
          if (sc < b1[r]) { b1[r] = sc; i1[r] = kk; }
        }
      }
    }
    // certify or rescan, then the exact f64 score of the winner
    const float4 g = __ldg(p.sguard + s);   // {cmin, cabs, Mmax, -}
#pragma unroll
    for (int r = 0; r < R; r++) {
      if (!valid[r]) continue;
      const uint32_t row = r * 32 + lane;
      const float* xf = p.x + (fbase + row) * p.dim;
      double best = 1e10;
      uint32_t bk = 0;
      if (k1 > k0) {
        const float dist2 = fmaxf(b2[r] - g.x, 0.f);
        const float guard = 2.f * eps * (dterm * dist2 + 2.f * sqrtf(dist2 * g.z) + g.y);
        const bool certain = (b2[r] - b1[r]) > guard;   // false for NaN; true when b2 = +inf (single density)
        if (certain) {
          const double sc = exact_score(p, xf, k0 + i1[r]);
          if (sc < best) { best = sc; bk = i1[r]; }     // min_score's own 1e10 start, Mixtures.cpp:699
        } else {
          n_fallback++;
          for (uint32_t k = k0; k < k1; k++) {
            const double sc = exact_score(p, xf, k);
            if (sc < best) { best = sc; bk = k - k0; }
          }
        }
      }
      // gathered (amin16): the tiled layout of sgmm_cell_rel; the task starts on a tile boundary
      const uint64_t cell = amin16 ? sgmm_cell_rel(out_stride, row, j) : (uint64_t)row * out_stride + j;
      if (out) out[cell] = best;
      if (amin) amin[cell] = bk;
      if (amin16) amin16[cell] = (uint16_t)bk;
    }
  }
  if (p.fallback_count && n_fallback) atomicAdd(p.fallback_count, n_fallback);
}

}  // namespace sgmm_ffma

int sgmm_launch_emission_ffma(sgmm_ctx* c, const uint4* d_tasks, uint64_t n_tasks, bool dense, uint64_t dense_f0,
                              uint64_t dense_nframes, double* d_scores, uint32_t* d_argmin) {
  using namespace sgmm_ffma;
  if (n_tasks == 0) return 0;
  Params p{};
  p.mu32 = c->d_mu32.p; p.h32 = c->d_h32.p; p.c32 = c->d_c32.p; p.sguard = c->d_sguard.p;
  p.meanT = c->d_meanT.p; p.invT = c->d_invT.p; p.mean_tail = c->d_mean_tail.p; p.inv_tail = c->d_inv_tail.p;
  p.norm = c->d_norm.p; p.logw = c->d_logw.p; p.state_off = c->d_state_off.p;
  p.dim = c->dim; p.npair = c->npair; p.Dtot = c->Dtot; p.n_states = c->n_states;
  p.x = c->d_x.p; p.n_tasks = n_tasks; p.tasks = d_tasks;
  p.frame_off = c->d_frame_off.p; p.uniq = c->d_uniq.p; p.uniq_off = c->d_uniq_off.p;
  p.nuniq_pad = c->d_nuniq_pad.p; p.E_off = c->d_E_off.p; p.E = c->d_E.p; p.Amin = c->d_Amin.p;
  p.dense = dense ? 1 : 0; p.f0 = dense_f0; p.nframes = dense_nframes;
  p.scores = d_scores; p.argmin = d_argmin;
  p.fallback_count = c->d_fallbacks.p;
  const unsigned grid = (unsigned)((n_tasks + kWarps - 1) / kWarps);
  switch (c->dv4) {
    case 7: k_emission_ffma<7><<<grid, kWarps * 32, 0, c->stream>>>(p); break;
    case 9: k_emission_ffma<9><<<grid, kWarps * 32, 0, c->stream>>>(p); break;
    case 12: k_emission_ffma<12><<<grid, kWarps * 32, 0, c->stream>>>(p); break;
    case 16: k_emission_ffma<16><<<grid, kWarps * 32, 0, c->stream>>>(p); break;
    default: SGMM_FAIL(c, "ffma scorer: unsupported feature dimension");
  }
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
