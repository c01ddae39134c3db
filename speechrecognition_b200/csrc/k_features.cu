// SURVEY.md §8 f3 (second half): load-time feature processing — SignalAnalysis::process_features (reference
// src/sietill/SignalAnalysis.cpp:379-399) with add_deltas (:320-337), the mean / variance normalisation (:389-394)
// and energy_max_normalization (:341-350) — on the device, so that raw .mm2 frames can be uploaded as they lie in
// the file and the processed [frames][total] f32 matrix never exists on the host.  Every operation is the
// reference's, with its roundings: deltas are f32 differences, the normalisation converts to f64, subtracts / divides
// and rounds back to f32 after EACH of the two steps (std::transform into a float range), the energy maximum is taken
// over the normalised column 0.  Results are bit-identical.  HBM-bound: 4 * n_file bytes in, 4 * total out per frame.
#include "sgmm_internal.cuh"

namespace {

struct FeatParams {
  const float* raw;            // [F][nf]
  float* out;                  // [F][nt]
  const uint64_t* frame_off;   // [U + 1]
  uint32_t n_utts;
  uint64_t n_frames;
  uint32_t nf, n1, n2, nt, step;
  const double* mean;          // [nt] or NULL
  const double* stddev;        // [nt] or NULL
};

// first derivative of raw column i at frame t of an utterance starting at row f0 (:322-328): the difference is
// taken between frames max(t, step) and max(t, step) - step
__device__ __forceinline__ float delta1(const FeatParams& p, uint64_t f0, uint32_t t, uint32_t i) {
  const uint32_t a = max(t, p.step);
  return __fsub_rn(__ldg(p.raw + (f0 + a) * p.nf + i), __ldg(p.raw + (f0 + a - p.step) * p.nf + i));
}

constexpr int kFeatFrames = 32;   // frames per CTA: their utterances are looked up once, then elements are coalesced

__global__ void __launch_bounds__(256) k_feat_process(FeatParams p) {
  __shared__ uint64_t s_f0[kFeatFrames];
  __shared__ uint32_t s_T[kFeatFrames];
  const uint64_t fb = (uint64_t)blockIdx.x * kFeatFrames;
  if (threadIdx.x < kFeatFrames && fb + threadIdx.x < p.n_frames) {
    const uint64_t f = fb + threadIdx.x;
    uint32_t lo = 0, hi = p.n_utts;            // utterance of frame f: largest u with frame_off[u] <= f
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (p.frame_off[mid] <= f) lo = mid; else hi = mid;
    }
    s_f0[threadIdx.x] = p.frame_off[lo];
    s_T[threadIdx.x] = (uint32_t)(p.frame_off[lo + 1] - p.frame_off[lo]);
  }
  __syncthreads();
  const uint32_t nfr = (uint32_t)min((uint64_t)kFeatFrames, p.n_frames - fb);
  for (uint32_t q = threadIdx.x; q < nfr * p.nt; q += 256) {
    const uint32_t j = q / p.nt, c = q - j * p.nt;
    const uint64_t f = fb + j, f0 = s_f0[j];
    const uint32_t T = s_T[j], t = (uint32_t)(f - f0);
    float v;
    if (c < p.nf) {
      v = __ldg(p.raw + f * p.nf + c);                                     // :383-387
    } else if (c < p.nf + p.n1) {
      v = delta1(p, f0, t, c - p.nf);                                      // :322-328
    } else {                                                               // :329-336: forward difference of the
      const uint32_t i = c - p.nf - p.n1;                                  // first derivatives (column n_file + i)
      const uint32_t b = min(t, T - 1 - p.step) + p.step;
      v = __fsub_rn(delta1(p, f0, b, i), delta1(p, f0, t, i));
    }
    if (p.mean) {                                                          // :389-394, two f32 roundings
      v = (float)__dsub_rn((double)v, __ldg(p.mean + c));
      v = (float)__ddiv_rn((double)v, __ldg(p.stddev + c));
    }
    p.out[fb * p.nt + q] = v;
  }
}

// :341-350 — one warp per utterance: maximum of column 0, then subtract it from every frame's column 0
__global__ void __launch_bounds__(128) k_feat_energy_max(float* __restrict__ out, const uint64_t* __restrict__ frame_off,
                                                         uint32_t n_utts, uint32_t nt) {
  const uint32_t u = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (u >= n_utts) return;
  const uint64_t f0 = frame_off[u], f1 = frame_off[u + 1];
  float m = -INFINITY;
  for (uint64_t f = f0 + lane; f < f1; f += 32) m = fmaxf(m, out[f * nt]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  for (uint64_t f = f0 + lane; f < f1; f += 32) out[f * nt] = __fsub_rn(out[f * nt], m);
}

}  // namespace

extern "C" int sgmm_process_features(sgmm_ctx* c, const float* raw, uint64_t n_frames, uint32_t n_file,
                                     uint32_t n_first, uint32_t n_second, uint32_t deriv_step, uint32_t n_utts,
                                     const uint64_t* frame_off, const double* mean, const double* stddev,
                                     int energy_max_norm, float* out) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (!raw || !frame_off || n_utts == 0) SGMM_FAIL(c, "process_features: null argument");
  if ((mean == nullptr) != (stddev == nullptr)) SGMM_FAIL(c, "process_features: mean and stddev go together");
  if (n_file == 0 || n_first > n_file || n_second > n_first)
    SGMM_FAIL(c, "process_features: need n_second <= n_first <= n_file (add_deltas reads those columns)");
  if (frame_off[0] != 0 || frame_off[n_utts] != n_frames) SGMM_FAIL(c, "process_features: frame offsets do not cover the frames");
  for (uint32_t u = 0; u < n_utts; u++)   // the reference indexes frame T - 1 - deriv_step (:330): UB for shorter segments
    if (frame_off[u + 1] < frame_off[u] + deriv_step + 1)
      SGMM_FAIL(c, "process_features: every utterance needs more than deriv_step frames");
  const uint32_t nt = n_file + n_first + n_second;
  if (c->feat_pending) {
    SGMM_CHECK(c, cudaStreamSynchronize(c->copy_stream));
    c->feat_pending = 0;
  }
  SGMM_CHECK(c, c->d_raw.reserve((size_t)n_frames * n_file));
  SGMM_CHECK(c, c->d_x.reserve((size_t)n_frames * nt + 8));
  SGMM_CHECK(c, c->d_feat_off.reserve(n_utts + 1));
  SGMM_CHECK(c, c->d_feat_norm.reserve(2 * (size_t)nt));
  {
    StageTimer st(c, SGMM_T_H2D);
    SGMM_CHECK(c, cudaMemcpyAsync(c->d_raw.p, raw, (size_t)n_frames * n_file * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    SGMM_CHECK(c, cudaMemcpyAsync(c->d_feat_off.p, frame_off, (n_utts + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    if (mean) {
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_feat_norm.p, mean, nt * sizeof(double), cudaMemcpyHostToDevice, c->stream));
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_feat_norm.p + nt, stddev, nt * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    }
  }
  FeatParams p{};
  p.raw = c->d_raw.p; p.out = c->d_x.p; p.frame_off = c->d_feat_off.p; p.n_utts = n_utts; p.n_frames = n_frames;
  p.nf = n_file; p.n1 = n_first; p.n2 = n_second; p.nt = nt; p.step = deriv_step;
  p.mean = mean ? c->d_feat_norm.p : nullptr; p.stddev = mean ? c->d_feat_norm.p + nt : nullptr;
  const uint64_t total = n_frames * nt;
  if (total) {
    k_feat_process<<<(unsigned)((n_frames + kFeatFrames - 1) / kFeatFrames), 256, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
    if (energy_max_norm) {
      k_feat_energy_max<<<(n_utts + 3) / 4, 128, 0, c->stream>>>(c->d_x.p, c->d_feat_off.p, n_utts, nt);
      SGMM_LAUNCH_CHECK(c);
    }
  }
  if (out) {
    StageTimer st(c, SGMM_T_D2H);
    SGMM_CHECK(c, cudaMemcpyAsync(out, c->d_x.p, total * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  }
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  // the processed matrix is now the context's resident feature matrix (as after sgmm_set_features)
  c->n_frames = n_frames; c->feat_dim = nt;
  c->have_alignment = false;
  c->ali_has_cols = false;
  c->tc_A_valid = false;
  return 0;
}
