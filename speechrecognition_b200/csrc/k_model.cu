// K5 — the model tables, and the device-resident M-step.
//
//  * sgmm_set_model: parameters arrive from the host (state-major f64 rows), ONE upload of the four arrays; every table
//    the scorers read — dim-pair-major f64 transposes, padded row-major twins, distinct variance rows, the FFMA scorer's
//    f32 tables with their per-state guard constants, the tf32 and fp16 hi/lo splits in the swizzled shared-memory image
//    the GEMM kernels copy verbatim — is derived ON THE DEVICE by the kernels below.  Every one of them is an exact
//    conversion, a fixed-order sum inside one thread, or an order-independent max, written with the same operation order
//    as the host builder it replaces (kept as the checker behind option "host_tables" = 1), so the tables are
//    bit-identical (tests/test_gpu_model.py).
//  * sgmm_mstep: MixtureModel::finalize + calculate_variance (reference Mixtures.cpp:374-461, :251-275) from the
//    statistics sgmm_accumulate* left on the device (already all-reduced), in the reference's operation order; the new
//    parameters never leave the GPU, the tables above are rebuilt from them.
//  * Both build into a spare table set and swap it in only after every step succeeded: a failed call leaves the
//    context's model exactly as it was.
#include <cuda_fp16.h>
#include <math_constants.h>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "sgmm_internal.cuh"

namespace {

template <typename T>
int upload(sgmm_ctx* c, DevBuf<T>& buf, const T* src, size_t n) {
  SGMM_CHECK(c, buf.reserve(n ? n : 1));
  if (n) SGMM_CHECK(c, cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// layout: everything that depends on the topology (state_off), the dimension and the `pooled` flag only
// ---------------------------------------------------------------------------------------------------------------
uint32_t dv4_of(uint32_t dim) { return dim <= 28 ? 7 : dim <= 36 ? 9 : dim <= 48 ? 12 : dim <= 64 ? 16 : 0; }

int model_layout(sgmm_ctx* c, SgmmModel& m, uint32_t dim, uint32_t n_states, const uint32_t* state_off, bool pooled,
                 const uint32_t* var_of, uint32_t n_var_rows) {
  const uint32_t D = state_off[n_states];
  uint32_t max_d = 0;
  for (uint32_t s = 0; s < n_states; s++) max_d = std::max(max_d, state_off[s + 1] - state_off[s]);
  m.dim = dim; m.npair = dim / 2; m.n_states = n_states; m.Dtot = D; m.max_dens = max_d;
  m.h_state_off.assign(state_off, state_off + n_states + 1);
  m.dimP = (dim + 3) & ~3u;                  // rows 32-byte aligned for 256-bit loads
  m.n_var_rows = n_var_rows;
  m.h_var_of.assign(var_of, var_of + D);
  m.dv4 = dv4_of(dim);
  m.pooled_states = pooled;
  const size_t npair = m.npair;
  SGMM_CHECK(c, m.d_mean_src.reserve((size_t)D * dim)); SGMM_CHECK(c, m.d_inv_src.reserve((size_t)D * dim));
  SGMM_CHECK(c, m.d_norm.reserve(D)); SGMM_CHECK(c, m.d_logw.reserve(D));
  SGMM_CHECK(c, m.d_meanT.reserve(std::max<size_t>(1, npair * D))); SGMM_CHECK(c, m.d_invT.reserve(std::max<size_t>(1, npair * D)));
  SGMM_CHECK(c, m.d_mean_tail.reserve(D)); SGMM_CHECK(c, m.d_inv_tail.reserve(D));
  SGMM_CHECK(c, m.d_meanR.reserve((size_t)D * m.dimP)); SGMM_CHECK(c, m.d_invU.reserve((size_t)n_var_rows * m.dimP));
  if (upload(c, m.d_state_off, state_off, (size_t)n_states + 1) || upload(c, m.d_var_of, var_of, (size_t)D)) return 1;
  if (m.dv4) {
    SGMM_CHECK(c, m.d_mu32.reserve((size_t)D * m.dv4)); SGMM_CHECK(c, m.d_h32.reserve((size_t)D * m.dv4));
    SGMM_CHECK(c, m.d_mu32T.reserve((size_t)D * m.dv4)); SGMM_CHECK(c, m.d_c32.reserve(D));
    SGMM_CHECK(c, m.d_sguard.reserve(n_states));
  }
  // Tensor-core operand layout (k_score_tc.cu).  Columns = densities, state-major, every state padded to a multiple of
  // 16 columns with never-winning dummies.  B row of density j:
  //   full  : [ h_0..h_{d-1} | g_0..g_{d-1} | c | 0.. ]   h = inv_var/2, g = -inv_var*mu, c = norm - logw + sum h mu^2
  //   pooled: [ g_0..g_{d-1} | c | 0.. ]  when all densities of every state share one variance vector (the x^2 block is
  //           then the same for all candidates of a state and drops out of the argmin).
  const uint32_t Kused = pooled ? dim + 1 : 2 * dim + 1;
  const uint32_t KP = (Kused + 7) / 8 * 8;     // whole k-steps of 8 (32-byte swizzle atoms: no further padding)
  m.tc_ok = KP / 8 <= 13;                      // A (hi, lo) = 16 * ksteps TMEM columns per 128 frames, <= 256
  m.tc_f16_ok = false;
  m.tc_rem16 = 0;
  m.h_col_off.assign(n_states + 1, 0);
  for (uint32_t s = 0; s < n_states; s++)
    m.h_col_off[s + 1] = m.h_col_off[s] + (state_off[s + 1] - state_off[s] + 15) / 16 * 16;
  if (!m.tc_ok) return 0;
  const uint32_t ks = KP / 8;
  // tile width: 256 densities (N = 256 MMAs run at the full tensor rate, N = 128 ones do not — DESIGN.md §4) when some
  // state is wider than 128 columns and three half-stage slots + one half's A parts fit shared memory
  uint32_t widest = 0;
  for (uint32_t s = 0; s < n_states; s++) widest = std::max(widest, m.h_col_off[s + 1] - m.h_col_off[s]);
  const bool fits256 = 1024 + 2 * (size_t)ks * 4096 + 3 * (size_t)ks * 8192 + 256 + 5120 <= 227 * 1024;
  const uint32_t TR = (c->tc_wide != 0 && fits256 && (widest > 128 || c->tc_wide == 1)) ? 256u : 128u;
  m.tc_tile_rows = TR;
  m.tc_KP = KP; m.tc_ksteps = (Kused + 7) / 8; m.tc_Dpad = m.h_col_off[n_states]; m.tc_full = pooled ? 0 : 1;
  // error model of one 3xTF32 dot product against |a|.|b|: hi/lo split leaves 2^-22 per operand and the dropped lo.lo
  // term another 2^-22; fp32 accumulation of 3 * 8 * ksteps terms, assumed to truncate (2^-23 each).
  m.tc_eps = 3.f * 2.384185791015625e-07f + 3.f * 8.f * (float)m.tc_ksteps * 1.1920928955078125e-07f;
  // every tile occupies a full ring stage: [hi, lo][k-step][TR rows][8 floats] in the 32-byte-swizzled shared-memory
  // image the MMA descriptors name (rows >= N stay zero and are never multiplied: the MMA's N says so)
  m.h_tile_first.assign(n_states + 1, 0);
  std::vector<uint2> tile_info, tile_info16;
  const bool want16 = pooled && TR == 256 && c->tc_f16 != 0;
  const uint32_t K16 = dim + 1, ks16 = (K16 + 15) / 16;
  size_t img = 0, img16 = 0;
  for (uint32_t s = 0; s < n_states; s++) {
    const uint32_t nc = m.h_col_off[s + 1] - m.h_col_off[s];
    for (uint32_t n0 = 0; n0 < nc; n0 += TR) {
      const uint32_t N = std::min(TR, nc - n0);
      tile_info.push_back(make_uint2((uint32_t)(img / 4), N));            // 16-byte units
      img += (size_t)2 * ks * TR * 8;
      if (want16) {
        tile_info16.push_back(make_uint2((uint32_t)(img16 / 8), N));
        img16 += (size_t)2 * ks16 * TR * 16;
      }
    }
    m.h_tile_first[s + 1] = (uint32_t)tile_info.size();
  }
  m.img_elems = img; m.img16_elems = want16 ? img16 : 0;
  if (img / 4 > 0xffffffffull) SGMM_FAIL(c, "set_model: tensor-core operand image exceeds 64 GB");
  SGMM_CHECK(c, m.d_Bimg.reserve(img)); SGMM_CHECK(c, m.d_sguard_tc.reserve(n_states));
  if (upload(c, m.d_tile_first, m.h_tile_first.data(), m.h_tile_first.size()) ||
      upload(c, m.d_tile_info, tile_info.data(), tile_info.size()) ||
      upload(c, m.d_col_off, m.h_col_off.data(), m.h_col_off.size()))
    return 1;
  if (want16) {
    // fp16 operand form (pooled models, 256-column tiles; k_tc_gemm256<KS, true>): K = d + 1 takes ceil((d + 1) / 16)
    // k-steps of 16 halves (3 at d = 33, against 5 tf32 k-steps of 8) at the same cycles per MMA
    SGMM_CHECK(c, m.d_Bimg16.reserve(img16)); SGMM_CHECK(c, m.d_sguard16.reserve(n_states));
    if (upload(c, m.d_tile_info16, tile_info16.data(), tile_info16.size())) return 1;
    m.tc_ks16 = ks16;
    const uint32_t rem = K16 % 16;
    m.tc_rem16 = (c->tc_rem != 0 && K16 >= 16 && rem > 0 && 3 * rem <= 16) ? rem : 0;
    // error model of one split dot product against |a|.|b| (cf. tc_eps): 2^-22 per operand split, 2^-22 for the
    // dropped lo.lo term; fp32 accumulation of 3 * 16 * ksteps products, assumed to truncate (2^-23 each)
    m.tc_eps16 = 3.f * 2.384185791015625e-07f + 3.f * 16.f * (float)ks16 * 1.1920928955078125e-07f;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// device table builder
// ---------------------------------------------------------------------------------------------------------------
struct BuildParams {
  const double* mean;      // [D][dim]
  const double* inv;       // [D][dim]
  const double* norm;      // [D]
  const double* logw;      // [D]
  const uint32_t* state_off;
  const uint32_t* var_of;
  uint32_t dim, npair, dimP, D, n_states, dv4;
  double2* meanT; double2* invT; double* mean_tail; double* inv_tail; double* meanR; double* invU;
  float* mu32; float* h32; float* mu32T; float* c32; float4* sguard;
  // tensor operands
  const uint32_t* col_off; const uint32_t* tile_first; const uint2* tile_info; const uint2* tile_info16;
  float* Bimg; __half* Bimg16; float4* sguard_tc; float4* sguard16;
  uint32_t KP, ks, TR, nsq, ks16, rem16;
  double inv2eps16;
  int* flags;
};

// exact f64 tables: dim-pair-major transposes, odd tails, padded row-major twins, distinct variance rows
__global__ void k_model_exact(BuildParams p) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t total = (size_t)p.D * p.dimP;
  if (idx >= total) return;
  const uint32_t j = (uint32_t)(idx / p.dimP), i = (uint32_t)(idx % p.dimP);
  const double mu = i < p.dim ? p.mean[(size_t)j * p.dim + i] : 0.0;
  const double iv = i < p.dim ? p.inv[(size_t)j * p.dim + i] : 0.0;
  p.meanR[idx] = mu;
  const uint32_t v = p.var_of[j];
  if (j == 0 || p.var_of[j - 1] != v) p.invU[(size_t)v * p.dimP + i] = iv;   // first density of its variance row
  if (i < 2 * p.npair) {
    double* mt = reinterpret_cast<double*>(p.meanT + (size_t)(i / 2) * p.D + j);
    double* vt = reinterpret_cast<double*>(p.invT + (size_t)(i / 2) * p.D + j);
    mt[i & 1] = mu; vt[i & 1] = iv;
  } else if (i == p.dim - 1) {    // odd dimension: the tail
    p.mean_tail[j] = mu; p.inv_tail[j] = iv;
  }
}
__global__ void k_model_zero_tails(BuildParams p) {   // even dimensions: the tails are defined (0) but unused
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < p.D) { p.mean_tail[j] = 0.0; p.inv_tail[j] = 0.0; }
}

__device__ __forceinline__ float nextafter_up(float v) {     // nextafterf(v, +inf) for finite v >= 0
  if (v != v || v == CUDART_INF_F) return v;
  if (v == 0.f) return __uint_as_float(1u);
  return v > 0.f ? __uint_as_float(__float_as_uint(v) + 1u) : __uint_as_float(__float_as_uint(v) - 1u);
}
__device__ __forceinline__ float nextafter_down(float v) {   // nextafterf(v, -inf)
  if (v != v || v == -CUDART_INF_F) return v;
  if (v == 0.f) return __uint_as_float(0x80000001u);
  return v > 0.f ? __uint_as_float(__float_as_uint(v) - 1u) : __uint_as_float(__float_as_uint(v) + 1u);
}

// std::min / std::max semantics of the host builder: (b < a) ? b : a and (a < b) ? b : a — a NaN candidate is ignored
__device__ __forceinline__ double hmin(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double hmax(double a, double b) { return (a < b) ? b : a; }

template <typename Op>
__device__ __forceinline__ double block_reduce(double v, double* sh, Op op) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double r = sh[0];
  for (int i = 1; i < (int)(blockDim.x >> 5); i++) r = op(r, sh[i]);
  return r;
}

// f32 tables of the FFMA scorer / two-stage rescan / pruned sum: one block per state
__global__ void __launch_bounds__(128) k_model_f32(BuildParams p) {
  __shared__ double sh[4];
  const uint32_t s = blockIdx.x;
  const uint32_t j0 = p.state_off[s], j1 = p.state_off[s + 1];
  const uint32_t row = 4 * p.dv4;
  double cmin = CUDART_INF, cabs = 0.0, Mmax = 0.0;
  for (uint32_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
    double M = 0.0;
    for (uint32_t i = 0; i < row; i++) {
      float mu32 = 0.f, h32 = 0.f;
      if (i < p.dim) {
        const double mu = p.mean[(size_t)j * p.dim + i], h = __dmul_rn(0.5, p.inv[(size_t)j * p.dim + i]);
        mu32 = __double2float_rn(mu); h32 = __double2float_rn(h);
        M = __dadd_rn(M, __dmul_rn(__dmul_rn(h, mu), mu));
      }
      p.mu32[(size_t)j * row + i] = mu32; p.h32[(size_t)j * row + i] = h32;
      p.mu32T[((size_t)(i / 4) * p.D + j) * 4 + (i & 3)] = mu32;
    }
    const double cc = __dsub_rn(p.norm[j], p.logw[j]);
    p.c32[j] = __double2float_rn(cc);
    cmin = hmin(cmin, cc); cabs = hmax(cabs, fabs(cc)); Mmax = hmax(Mmax, M);
  }
  cmin = block_reduce(cmin, sh, [](double a, double b) { return hmin(a, b); });
  cabs = block_reduce(cabs, sh, [](double a, double b) { return hmax(a, b); });
  Mmax = block_reduce(Mmax, sh, [](double a, double b) { return hmax(a, b); });
  if (threadIdx.x == 0)   // guard constants rounded conservatively (cmin down, the others up)
    p.sguard[s] = make_float4(nextafter_down(__double2float_rn(cmin)), nextafter_up(__double2float_rn(cabs)),
                              nextafter_up(__double2float_rn(Mmax)), 0.f);
}

__device__ __forceinline__ float tf32_rna(float v) {   // round to nearest, ties away (cvt.rna.tf32.f32), as the host builder
  uint32_t u = __float_as_uint(v);
  if ((u & 0x7f800000u) == 0x7f800000u) return v;
  u += 0x1000u; u &= 0xffffe000u;
  return __uint_as_float(u);
}

// tf32 hi/lo image + per-state guards: one block per state, one thread per column (density or dummy)
__global__ void __launch_bounds__(128) k_model_tf32(BuildParams p) {
  __shared__ double sh[4];
  const uint32_t s = blockIdx.x;
  const uint32_t j0 = p.state_off[s], nreal = p.state_off[s + 1] - j0;
  const uint32_t ncol = p.col_off[s + 1] - p.col_off[s];
  const uint32_t t0 = p.tile_first[s];
  double Gmax = 0.0, Hmax = 0.0, cabs = 0.0;
  for (uint32_t q = threadIdx.x; q < ncol; q += blockDim.x) {
    const uint2 ti = p.tile_info[t0 + q / p.TR];
    float* base = p.Bimg + (size_t)ti.x * 4;
    const uint32_t r = q % p.TR;
    auto put = [&](uint32_t k, double v) {
      const float hi = tf32_rna(__double2float_rn(v));
      const float lo = tf32_rna(__double2float_rn(__dsub_rn(v, (double)hi)));
      const uint32_t lin = (k / 8) * p.TR * 32 + r * 32 + (k % 8) * 4;
      const uint32_t phys = lin ^ (((lin >> 7) & 1u) << 4);
      base[phys / 4] = hi;
      base[(size_t)p.ks * p.TR * 8 + phys / 4] = lo;
    };
    if (q >= nreal) { put(p.nsq + p.dim, 1e30); continue; }
    const uint32_t j = j0 + q;
    double cc = __dsub_rn(p.norm[j], p.logw[j]), G = 0.0, H = 0.0;
    for (uint32_t i = 0; i < p.dim; i++) {
      const double mu = p.mean[(size_t)j * p.dim + i], iv = p.inv[(size_t)j * p.dim + i];
      const double h = __dmul_rn(0.5, iv), g = __dmul_rn(-iv, mu);
      if (p.nsq) { put(i, h); H = __dadd_rn(H, __dmul_rn(h, h)); }
      put(p.nsq + i, g); G = __dadd_rn(G, __dmul_rn(g, g));
      cc = __dadd_rn(cc, __dmul_rn(__dmul_rn(h, mu), mu));
    }
    put(p.nsq + p.dim, cc);
    Gmax = hmax(Gmax, __dsqrt_rn(G)); Hmax = hmax(Hmax, __dsqrt_rn(H)); cabs = hmax(cabs, fabs(cc));
  }
  Gmax = block_reduce(Gmax, sh, [](double a, double b) { return hmax(a, b); });
  Hmax = block_reduce(Hmax, sh, [](double a, double b) { return hmax(a, b); });
  cabs = block_reduce(cabs, sh, [](double a, double b) { return hmax(a, b); });
  if (threadIdx.x == 0)
    p.sguard_tc[s] = make_float4(__double2float_rn(__dmul_rn(Gmax, 1.0001)), __double2float_rn(__dmul_rn(Hmax, 1.0001)),
                                 __double2float_rn(__dmul_rn(cabs, 1.0001)), 0.f);
}

// fp16 hi/lo image.  B row of density j of state s: t_s * [ g_0..g_{d-1} | c | 0.. ], t_s = the power of two that brings
// the state's largest |entry| into [2^13, 2^14), every entry split into fp16 hi + fp16 lo.  Per-state guard constants, in
// the slots the epilogue multiplies with k_tc_prep16's {||s x||, s}: {max ||t g|| + flush terms, max |t c| + flush, 0}.
__global__ void __launch_bounds__(128) k_model_f16(BuildParams p) {
  __shared__ double sh[4];
  const uint32_t s = blockIdx.x;
  const uint32_t j0 = p.state_off[s], nreal = p.state_off[s + 1] - j0;
  const uint32_t ncol = p.col_off[s + 1] - p.col_off[s];
  const uint32_t t0 = p.tile_first[s];
  double bmax = 0.0;
  for (uint32_t q = threadIdx.x; q < nreal; q += blockDim.x) {
    const uint32_t j = j0 + q;
    double cc = __dsub_rn(p.norm[j], p.logw[j]);
    for (uint32_t i = 0; i < p.dim; i++) {
      const double mu = p.mean[(size_t)j * p.dim + i], iv = p.inv[(size_t)j * p.dim + i];
      bmax = hmax(bmax, fabs(__dmul_rn(iv, mu)));
      cc = __dadd_rn(cc, __dmul_rn(__dmul_rn(__dmul_rn(0.5, iv), mu), mu));
    }
    bmax = hmax(bmax, fabs(cc));
  }
  bmax = block_reduce(bmax, sh, [](double a, double b) { return hmax(a, b); });
  if (!(bmax < 1e300)) {                       // non-finite model entries: the context keeps the tf32 form
    if (threadIdx.x == 0) atomicOr(p.flags, 1);
    return;
  }
  int e = 0;
  frexp(hmax(bmax, 1e-300), &e);               // bmax = m 2^e, m in [0.5, 1)
  const double ts = ldexp(1.0, 14 - e);        // bmax * ts in [2^13, 2^14)
  double Gmax = 0.0, cabs = 0.0, flg = 0.0, flc = 0.0;
  const uint32_t K = p.dim + 1;
  for (uint32_t q = threadIdx.x; q < ncol; q += blockDim.x) {
    const uint2 ti = p.tile_info16[t0 + q / p.TR];
    __half* base = p.Bimg16 + (size_t)ti.x * 8;
    const uint32_t r = q % p.TR;
    const bool real = q < nreal;
    const uint32_t j = j0 + q;
    double cc = real ? __dsub_rn(p.norm[j], p.logw[j]) : 0.0, G = 0.0, flrow = 0.0;
    for (uint32_t k = 0; k < K; k++) {
      double v;
      if (!real) v = k == p.dim ? 60000.0 : 0.0;   // never-winning dummy (a certified dummy is rescanned)
      else if (k < p.dim) {
        const double mu = p.mean[(size_t)j * p.dim + k], iv = p.inv[(size_t)j * p.dim + k];
        v = __dmul_rn(__dmul_rn(-iv, mu), ts);
        cc = __dadd_rn(cc, __dmul_rn(__dmul_rn(__dmul_rn(0.5, iv), mu), mu));
      } else v = __dmul_rn(cc, ts);
      const __half hi = __float2half_rn(__double2float_rn(v));
      const double rem = __dsub_rn(v, (double)__half2float(hi));
      const __half lo = __float2half_rn(__double2float_rn(rem));
      auto put = [&](uint32_t img, uint32_t kpos, __half val) {   // img 0: the hi slot's image, 1: the lo slot's
        const uint32_t lin = (kpos / 16) * p.TR * 32 + r * 32 + (kpos % 16) * 2;
        const uint32_t phys = lin ^ (((lin >> 7) & 1u) << 4);
        base[(size_t)img * p.ks16 * p.TR * 16 + phys / 2] = val;
      };
      const uint32_t kfull = K - p.rem16;                         // terms in whole k-steps (all of them when not packed)
      if (p.rem16 && k >= kfull) {   // packed remainder: [hi | hi | lo] in the hi slot's last k-step (sgmm_internal.cuh)
        put(0, k, hi); put(0, k + p.rem16, hi); put(0, k + 2 * p.rem16, lo);
      } else { put(0, k, hi); put(1, k, lo); }
      double fl = 0.0;                              // subnormal parts: see k_tc_prep16
      if (rem != 0.0 && fabs(rem) < 6.103515625e-05) fl = __dadd_rn(fl, hmax(fabs(rem), 2.98023223876953125e-08));
      if (v != 0.0 && fabs(v) < 6.103515625e-05) fl = __dadd_rn(fl, fabs(v));
      if (real) {
        if (k < p.dim) { G = __dadd_rn(G, __dmul_rn(v, v)); flrow = hmax(flrow, fl); }
        else { cabs = hmax(cabs, fabs(v)); flc = hmax(flc, fl); }
      }
    }
    if (real) { Gmax = hmax(Gmax, __dsqrt_rn(G)); flg = hmax(flg, flrow); }
  }
  Gmax = block_reduce(Gmax, sh, [](double a, double b) { return hmax(a, b); });
  cabs = block_reduce(cabs, sh, [](double a, double b) { return hmax(a, b); });
  flg = block_reduce(flg, sh, [](double a, double b) { return hmax(a, b); });
  flc = block_reduce(flc, sh, [](double a, double b) { return hmax(a, b); });
  if (threadIdx.x == 0) {
    // a flushed / rounded B entry b_i errs by fl_i |a_i| <= fl_i ||a||: fold max_i fl_i * sqrt(d) / (2 eps) into the norm
    const double gn = __dadd_rn(Gmax, __dmul_rn(__dmul_rn(flg, __dsqrt_rn((double)p.dim)), p.inv2eps16));
    const double cn = __dadd_rn(cabs, __dmul_rn(flc, p.inv2eps16));
    // .w: 1 / t_s (a power of two) — un-scales an accumulator value into a score in the tolerance mode's epilogue
    p.sguard16[s] = make_float4(__double2float_rn(__dmul_rn(gn, 1.0001)), __double2float_rn(__dmul_rn(cn, 1.0001)), 0.f,
                                __double2float_rn(__ddiv_rn(1.0, ts)));
  }
}

int model_build_device(sgmm_ctx* c, SgmmModel& m) {
  BuildParams p{};
  p.mean = m.d_mean_src.p; p.inv = m.d_inv_src.p; p.norm = m.d_norm.p; p.logw = m.d_logw.p;
  p.state_off = m.d_state_off.p; p.var_of = m.d_var_of.p;
  p.dim = m.dim; p.npair = m.npair; p.dimP = m.dimP; p.D = m.Dtot; p.n_states = m.n_states; p.dv4 = m.dv4;
  p.meanT = m.d_meanT.p; p.invT = m.d_invT.p; p.mean_tail = m.d_mean_tail.p; p.inv_tail = m.d_inv_tail.p;
  p.meanR = m.d_meanR.p; p.invU = m.d_invU.p;
  p.mu32 = reinterpret_cast<float*>(m.d_mu32.p); p.h32 = reinterpret_cast<float*>(m.d_h32.p);
  p.mu32T = reinterpret_cast<float*>(m.d_mu32T.p); p.c32 = m.d_c32.p; p.sguard = m.d_sguard.p;
  p.col_off = m.d_col_off.p; p.tile_first = m.d_tile_first.p; p.tile_info = m.d_tile_info.p; p.tile_info16 = m.d_tile_info16.p;
  p.Bimg = m.d_Bimg.p; p.Bimg16 = m.d_Bimg16.p; p.sguard_tc = m.d_sguard_tc.p; p.sguard16 = m.d_sguard16.p;
  p.KP = m.tc_KP; p.ks = m.tc_KP / 8; p.TR = m.tc_tile_rows; p.nsq = m.tc_full ? m.dim : 0; p.ks16 = m.tc_ks16; p.rem16 = m.tc_rem16;
  p.inv2eps16 = m.tc_eps16 > 0.f ? 1.0 / (2.0 * (double)m.tc_eps16) : 0.0;
  SGMM_CHECK(c, c->d_model_flags.reserve(4));
  SGMM_CHECK(c, cudaMemsetAsync(c->d_model_flags.p, 0, 4 * sizeof(int), c->stream));
  p.flags = c->d_model_flags.p;
  const size_t total = (size_t)m.Dtot * m.dimP;
  k_model_exact<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(p);
  SGMM_LAUNCH_CHECK(c);
  if (!(m.dim & 1)) {
    k_model_zero_tails<<<(m.Dtot + 255) / 256, 256, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
  }
  if (m.dv4) {
    k_model_f32<<<m.n_states, 128, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
  }
  if (m.tc_ok) {
    SGMM_CHECK(c, cudaMemsetAsync(m.d_Bimg.p, 0, m.img_elems * sizeof(float), c->stream));
    k_model_tf32<<<m.n_states, 128, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
    if (m.img16_elems) {
      SGMM_CHECK(c, cudaMemsetAsync(m.d_Bimg16.p, 0, m.img16_elems * sizeof(__half), c->stream));
      k_model_f16<<<m.n_states, 128, 0, c->stream>>>(p);
      SGMM_LAUNCH_CHECK(c);
    }
  }
  int flags[4] = {0, 0, 0, 0};
  SGMM_CHECK(c, cudaMemcpyAsync(flags, c->d_model_flags.p, sizeof(flags), cudaMemcpyDeviceToHost, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  m.tc_f16_ok = m.tc_ok && m.img16_elems > 0 && !(flags[0] & 1);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// host table builder — the checker of the device builder (option "host_tables" = 1).  Same tables, same bits.
// ---------------------------------------------------------------------------------------------------------------
int model_build_host(sgmm_ctx* c, SgmmModel& m, const double* mean, const double* inv_var, const double* norm,
                     const double* log_weight) {
  const uint32_t dim = m.dim, D = m.Dtot, n_states = m.n_states, npair = m.npair, dimP = m.dimP;
  const uint32_t* state_off = m.h_state_off.data();
  std::vector<double2> mT((size_t)npair * D), vT((size_t)npair * D);
  std::vector<double> mt(D, 0.0), vt(D, 0.0);
  std::vector<double> mR((size_t)D * dimP, 0.0), vU((size_t)m.n_var_rows * dimP, 0.0);
  for (uint32_t j = 0; j < D; j++) {
    const double* mu = mean + (size_t)j * dim;
    const double* iv = inv_var + (size_t)j * dim;
    for (uint32_t ip = 0; ip < npair; ip++) {
      mT[(size_t)ip * D + j] = make_double2(mu[2 * ip], mu[2 * ip + 1]);
      vT[(size_t)ip * D + j] = make_double2(iv[2 * ip], iv[2 * ip + 1]);
    }
    if (dim & 1) { mt[j] = mu[dim - 1]; vt[j] = iv[dim - 1]; }
    memcpy(&mR[(size_t)j * dimP], mu, dim * sizeof(double));
    memcpy(&vU[(size_t)m.h_var_of[j] * dimP], iv, dim * sizeof(double));
  }
  if (upload(c, m.d_meanT, mT.data(), mT.size()) || upload(c, m.d_invT, vT.data(), vT.size()) ||
      upload(c, m.d_mean_tail, mt.data(), mt.size()) || upload(c, m.d_inv_tail, vt.data(), vt.size()) ||
      upload(c, m.d_meanR, mR.data(), mR.size()) || upload(c, m.d_invU, vU.data(), vU.size()))
    return 1;
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  const uint32_t dv4 = m.dv4;
  if (dv4) {
    const size_t row = 4 * (size_t)dv4;
    std::vector<float> mu32((size_t)D * row, 0.f), h32((size_t)D * row, 0.f), c32(D, 0.f), mu32T((size_t)D * row);
    std::vector<float4> sg(n_states, make_float4(0.f, 0.f, 0.f, 0.f));
    for (uint32_t s = 0; s < n_states; s++) {
      double cmin = INFINITY, cabs = 0.0, Mmax = 0.0;
      for (uint32_t j = state_off[s]; j < state_off[s + 1]; j++) {
        double M = 0.0;
        for (uint32_t i = 0; i < dim; i++) {
          const double mu = mean[(size_t)j * dim + i], h = 0.5 * inv_var[(size_t)j * dim + i];
          mu32[j * row + i] = (float)mu; h32[j * row + i] = (float)h;
          M += h * mu * mu;
        }
        const double cc = norm[j] - log_weight[j];
        c32[j] = (float)cc;
        cmin = std::min(cmin, cc); cabs = std::max(cabs, std::fabs(cc)); Mmax = std::max(Mmax, M);
      }
      sg[s] = make_float4(nextafterf((float)cmin, -INFINITY), nextafterf((float)cabs, INFINITY),
                          nextafterf((float)Mmax, INFINITY), 0.f);
    }
    for (uint32_t j = 0; j < D; j++)
      for (uint32_t v = 0; v < dv4; v++) memcpy(&mu32T[((size_t)v * D + j) * 4], &mu32[j * row + 4 * v], 4 * sizeof(float));
    if (upload(c, m.d_mu32, reinterpret_cast<const float4*>(mu32.data()), (size_t)D * dv4) ||
        upload(c, m.d_h32, reinterpret_cast<const float4*>(h32.data()), (size_t)D * dv4) ||
        upload(c, m.d_c32, c32.data(), (size_t)D) || upload(c, m.d_sguard, sg.data(), sg.size()) ||
        upload(c, m.d_mu32T, reinterpret_cast<const float4*>(mu32T.data()), (size_t)D * dv4))
      return 1;
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  }
  if (!m.tc_ok) return 0;
  const bool pooled = !m.tc_full;
  const uint32_t KP = m.tc_KP, ks = KP / 8, TR = m.tc_tile_rows, nsq = pooled ? 0 : dim;
  const std::vector<uint32_t>& col_off = m.h_col_off;
  {
    std::vector<float> img(m.img_elems, 0.f);
    std::vector<float4> sgt(n_states, make_float4(0.f, 0.f, 0.f, 0.f));
    auto tf32 = [](float v) {   // round to nearest, ties away (cvt.rna.tf32.f32)
      uint32_t u; memcpy(&u, &v, 4);
      if ((u & 0x7f800000u) == 0x7f800000u) return v;
      u += 0x1000u; u &= 0xffffe000u;
      float r; memcpy(&r, &u, 4); return r;
    };
    size_t tile = 0;
    for (uint32_t s = 0; s < n_states; s++) {
      double Gmax = 0.0, Hmax = 0.0, cabs = 0.0;
      const uint32_t nreal = state_off[s + 1] - state_off[s], nc = col_off[s + 1] - col_off[s];
      for (uint32_t q = 0; q < nc; q++) {
        float* base = img.data() + (tile + q / TR) * ((size_t)2 * ks * TR * 8);
        const uint32_t r = q % TR;
        auto put = [&](uint32_t k, double v) {
          const float hi = tf32((float)v);
          const uint32_t lin = (k / 8) * TR * 32 + r * 32 + (k % 8) * 4;
          const uint32_t phys = lin ^ (((lin >> 7) & 1u) << 4);
          base[phys / 4] = hi;
          base[(size_t)ks * TR * 8 + phys / 4] = tf32((float)(v - (double)hi));
        };
        if (q >= nreal) { put(nsq + dim, 1e30); continue; }
        const uint32_t j = state_off[s] + q;
        double cc = norm[j] - log_weight[j], G = 0.0, H = 0.0;
        for (uint32_t i = 0; i < dim; i++) {
          const double mu = mean[(size_t)j * dim + i], iv = inv_var[(size_t)j * dim + i];
          const double h = 0.5 * iv, g = -iv * mu;
          if (!pooled) { put(i, h); H += h * h; }
          put(nsq + i, g); G += g * g;
          cc += h * mu * mu;
        }
        put(nsq + dim, cc);
        Gmax = std::max(Gmax, std::sqrt(G)); Hmax = std::max(Hmax, std::sqrt(H)); cabs = std::max(cabs, std::fabs(cc));
      }
      sgt[s] = make_float4((float)(Gmax * 1.0001), (float)(Hmax * 1.0001), (float)(cabs * 1.0001), 0.f);
      tile += (nc + TR - 1) / TR;
    }
    if (upload(c, m.d_Bimg, img.data(), img.size()) || upload(c, m.d_sguard_tc, sgt.data(), sgt.size())) return 1;
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  }
  if (!m.img16_elems) return 0;
  {
    const uint32_t K = dim + 1, ks16 = m.tc_ks16;
    const double inv2eps = 1.0 / (2.0 * (double)m.tc_eps16);
    std::vector<__half> img(m.img16_elems, __float2half(0.f));
    std::vector<float4> sg(n_states, make_float4(0.f, 0.f, 0.f, 0.f));
    std::vector<double> row(K);
    size_t tile = 0;
    bool finite = true;
    for (uint32_t s = 0; s < n_states && finite; s++) {
      const uint32_t nreal = state_off[s + 1] - state_off[s], nc = col_off[s + 1] - col_off[s];
      double bmax = 0.0;
      for (uint32_t j = state_off[s]; j < state_off[s + 1]; j++) {
        double cc = norm[j] - log_weight[j];
        for (uint32_t i = 0; i < dim; i++) {
          const double mu = mean[(size_t)j * dim + i], iv = inv_var[(size_t)j * dim + i];
          bmax = std::max(bmax, std::fabs(iv * mu));
          cc += 0.5 * iv * mu * mu;
        }
        bmax = std::max(bmax, std::fabs(cc));
      }
      if (!(bmax < 1e300)) { finite = false; break; }      // non-finite model entries: keep the tf32 form
      int e = 0;
      std::frexp(std::max(bmax, 1e-300), &e);              // bmax = m 2^e, m in [0.5, 1)
      const double ts = std::ldexp(1.0, 14 - e);           // bmax * ts in [2^13, 2^14)
      double Gmax = 0.0, cabs = 0.0, flg = 0.0, flc = 0.0;
      for (uint32_t q = 0; q < nc; q++) {
        __half* base = img.data() + (tile + q / TR) * ((size_t)2 * ks16 * TR * 16);
        const uint32_t r = q % TR;
        std::fill(row.begin(), row.end(), 0.0);
        if (q >= nreal) {
          row[dim] = 60000.0;                              // never-winning dummy (a certified dummy is rescanned)
        } else {
          const uint32_t j = state_off[s] + q;
          double cc = norm[j] - log_weight[j];
          for (uint32_t i = 0; i < dim; i++) {
            const double mu = mean[(size_t)j * dim + i], iv = inv_var[(size_t)j * dim + i];
            row[i] = -iv * mu * ts;
            cc += 0.5 * iv * mu * mu;
          }
          row[dim] = cc * ts;
        }
        double G = 0.0, flrow = 0.0;
        for (uint32_t k = 0; k < K; k++) {
          const __half hi = __float2half_rn((float)row[k]);
          const double rem = row[k] - (double)__half2float(hi);
          const __half lo = __float2half_rn((float)rem);
          auto put = [&](uint32_t img, uint32_t kpos, __half val) {
            const uint32_t lin = (kpos / 16) * TR * 32 + r * 32 + (kpos % 16) * 2;
            const uint32_t phys = lin ^ (((lin >> 7) & 1u) << 4);
            base[(size_t)img * ks16 * TR * 16 + phys / 2] = val;
          };
          const uint32_t rem16 = m.tc_rem16, kfull = K - rem16;
          if (rem16 && k >= kfull) { put(0, k, hi); put(0, k + rem16, hi); put(0, k + 2 * rem16, lo); }
          else { put(0, k, hi); put(1, k, lo); }
          double fl = 0.0;                                  // subnormal parts: see k_tc_prep16
          if (rem != 0.0 && std::fabs(rem) < 6.103515625e-05) fl += std::max(std::fabs(rem), 2.98023223876953125e-08);
          if (row[k] != 0.0 && std::fabs(row[k]) < 6.103515625e-05) fl += std::fabs(row[k]);
          if (q < nreal) {
            if (k < dim) { G += row[k] * row[k]; flrow = std::max(flrow, fl); }
            else { cabs = std::max(cabs, std::fabs(row[k])); flc = std::max(flc, fl); }
          }
        }
        if (q < nreal) { Gmax = std::max(Gmax, std::sqrt(G)); flg = std::max(flg, flrow); }
      }
      // a flushed / rounded B entry b_i errs by fl_i |a_i| <= fl_i ||a||: fold max_i fl_i * sqrt(d) / (2 eps) into the norm
      sg[s] = make_float4((float)((Gmax + flg * std::sqrt((double)dim) * inv2eps) * 1.0001),
                          (float)((cabs + flc * inv2eps) * 1.0001), 0.f, (float)(1.0 / ts));
      tile += (nc + TR - 1) / TR;
    }
    if (finite) {
      if (upload(c, m.d_Bimg16, img.data(), img.size()) || upload(c, m.d_sguard16, sg.data(), sg.size())) return 1;
      SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
    }
    m.tc_f16_ok = finite;
  }
  return 0;
}

// the new model becomes the context's; what it held goes to the spare set (its buffers are reused by the next build)
void model_commit(sgmm_ctx* c) {
  std::swap(static_cast<SgmmModel&>(*c), c->spare);
  c->model_version++;     // (the tensor scorer's A rows stay valid: they depend on the features and the operand form only)
}

// ---------------------------------------------------------------------------------------------------------------
// M-step kernels (Mixtures.cpp:374-461, :251-275)
// ---------------------------------------------------------------------------------------------------------------
struct MstepParams {
  const double* acc;        // [count D | sum_x D*d | sum_xx D*d | ...]
  const uint32_t* state_off;
  uint32_t D, dim, n_states;
  int pooling;              // 0 none, 1 mixture, 2 global
  double seed;              // minimal_variance_value_ (Mixtures.cpp:243): every variance accumulator starts from it
  double dlog2pi;           // dimension * log(2 * M_PI), the host's value (:268-270)
  double* mean; double* inv; double* var; double* norm; double* logw; double* weight;
  double* state_total;      // [n_states] total_mixture_observations
  double* gvar;             // global pooling: [dim] variance row, [dim] inverse, [1] norm
};

// one block per state: means, weights, and the variances of NO / MIXTURE pooling
__global__ void __launch_bounds__(512) k_mstep_state(MstepParams p) {
  extern __shared__ double shm[];     // [dim] pooled mean / variance scratch, [1] total
  const uint32_t s = blockIdx.x, d = p.dim;
  const uint32_t j0 = p.state_off[s], j1 = p.state_off[s + 1];
  const double* cnt = p.acc;
  const double* sx = p.acc + p.D;
  const double* sxx = p.acc + p.D + (size_t)p.D * d;
  double* total = shm + d;
  if (threadIdx.x == 0) {             // total_mixture_observations: in-order sum from 0.0 (:381-386)
    double t = 0.0;
    uint32_t j = j0;
    for (; j + 8 <= j1; j += 8) {     // (eight loads in flight, added in order: the chain is the additions, not the L2)
      double v[8];
#pragma unroll
      for (int q = 0; q < 8; q++) v[q] = cnt[j + q];
#pragma unroll
      for (int q = 0; q < 8; q++) t = __dadd_rn(t, v[q]);
    }
    for (; j < j1; j++) t = __dadd_rn(t, cnt[j]);
    *total = t;
    p.state_total[s] = t;
  }
  __syncthreads();
  const double tot = *total;
  // means (:389-392), weights and their logs (:402-406)
  for (uint32_t idx = threadIdx.x; idx < (j1 - j0) * d; idx += blockDim.x) {
    const uint32_t j = j0 + idx / d, i = idx % d;
    p.mean[(size_t)j * d + i] = __ddiv_rn(sx[(size_t)j * d + i], cnt[j]);
  }
  for (uint32_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
    const double w = __ddiv_rn(cnt[j], tot);
    p.weight[j] = w;
    p.logw[j] = log(w);
  }
  if (p.pooling == 0) {
    // calculate_variance per density around its own mean (:394-397, :251-275): acc / weight, minus mean^2, inverse,
    // norm = (d log 2pi + sum_i log var_i) / 2 with the logs added in dimension order
    __syncthreads();
    for (uint32_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
      const double w = __dadd_rn(0.0, cnt[j]);
      double n = p.dlog2pi;
      for (uint32_t i = 0; i < d; i++) {
        const double mu = p.mean[(size_t)j * d + i];
        double v = __ddiv_rn(__dadd_rn(p.seed, sxx[(size_t)j * d + i]), w);
        v = __dsub_rn(v, __dmul_rn(mu, mu));
        p.var[(size_t)j * d + i] = v;
        p.inv[(size_t)j * d + i] = __ddiv_rn(1.0, v);
        n = __dadd_rn(n, log(v));
      }
      p.norm[j] = __ddiv_rn(n, 2.0);
    }
  } else if (p.pooling == 1) {
    // variance of the mixture around its pooled mean (:408-427): sums over the densities in order
    for (uint32_t i = threadIdx.x; i < d; i += blockDim.x) {
      double va = p.seed, pm = 0.0;
      uint32_t j = j0;
      for (; j + 8 <= j1; j += 8) {   // same in-order sums, eight densities' loads in flight
        double a[8], b[8];
#pragma unroll
        for (int q = 0; q < 8; q++) { a[q] = sxx[(size_t)(j + q) * d + i]; b[q] = sx[(size_t)(j + q) * d + i]; }
#pragma unroll
        for (int q = 0; q < 8; q++) { va = __dadd_rn(va, a[q]); pm = __dadd_rn(pm, b[q]); }
      }
      for (; j < j1; j++) {
        va = __dadd_rn(va, sxx[(size_t)j * d + i]);
        pm = __dadd_rn(pm, sx[(size_t)j * d + i]);
      }
      pm = __ddiv_rn(pm, tot);
      double v = __ddiv_rn(va, tot);            // var_weight_accumulators_ received the same counts in the same order
      v = __dsub_rn(v, __dmul_rn(pm, pm));
      shm[i] = v;
    }
    __syncthreads();
    double n = p.dlog2pi;
    if (threadIdx.x == 0) {
      for (uint32_t i = 0; i < d; i++) n = __dadd_rn(n, log(shm[i]));
      n = __ddiv_rn(n, 2.0);
      *total = n;
    }
    __syncthreads();
    n = *total;
    for (uint32_t idx = threadIdx.x; idx < (j1 - j0) * d; idx += blockDim.x) {
      const uint32_t j = j0 + idx / d, i = idx % d;
      p.var[(size_t)j * d + i] = shm[i];
      p.inv[(size_t)j * d + i] = __ddiv_rn(1.0, shm[i]);
    }
    for (uint32_t j = j0 + threadIdx.x; j < j1; j += blockDim.x) p.norm[j] = n;
  }
}

// global pooling (:432-450): one variance row around the global mean; thread i walks every density in order
__global__ void __launch_bounds__(128) k_mstep_global(MstepParams p) {
  __shared__ double total_obs;
  const uint32_t d = p.dim;
  const double* cnt = p.acc;
  const double* sx = p.acc + p.D;
  const double* sxx = p.acc + p.D + (size_t)p.D * d;
  if (threadIdx.x == 0) {
    double t = 0.0;                              // total_observations: sum of the per-mixture totals (:429)
    for (uint32_t s = 0; s < p.n_states; s++) t = __dadd_rn(t, p.state_total[s]);
    total_obs = t;
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < d; i += blockDim.x) {
    double va = p.seed, gm = 0.0;
    for (uint32_t j = 0; j < p.D; j++) {
      va = __dadd_rn(va, sxx[(size_t)j * d + i]);
      gm = __dadd_rn(gm, sx[(size_t)j * d + i]);
    }
    gm = __ddiv_rn(gm, total_obs);
    double w = 0.0;                              // var_weight_accumulators_[0]: every density's count, in order
    for (uint32_t j = 0; j < p.D; j++) w = __dadd_rn(w, cnt[j]);
    double v = __ddiv_rn(va, w);
    v = __dsub_rn(v, __dmul_rn(gm, gm));
    p.gvar[i] = v;
    p.gvar[d + i] = __ddiv_rn(1.0, v);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double n = p.dlog2pi;
    for (uint32_t i = 0; i < d; i++) n = __dadd_rn(n, log(p.gvar[i]));
    p.gvar[2 * d] = __ddiv_rn(n, 2.0);
  }
}
__global__ void k_mstep_global_spread(MstepParams p) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)p.D * p.dim) return;
  const uint32_t i = (uint32_t)(idx % p.dim);
  p.var[idx] = p.gvar[i];
  p.inv[idx] = p.gvar[p.dim + i];
  if (i == 0) p.norm[idx / p.dim] = p.gvar[2 * p.dim];
}

}  // namespace

extern "C" {

int sgmm_set_model(sgmm_ctx* c, uint32_t dim, uint32_t n_states, const uint32_t* state_off,
                   const double* mean, const double* inv_var, const double* norm, const double* log_weight) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (!state_off || !mean || !inv_var || !norm || !log_weight) SGMM_FAIL(c, "set_model: null argument");
  if (dim < 2 || dim > 128) SGMM_FAIL(c, "set_model: dim must be in [2, 128]");
  if (n_states == 0 || n_states > 65535) SGMM_FAIL(c, "set_model: n_states must be in [1, 65535] (StateIdx is 16 bit)");
  const uint32_t D = state_off[n_states];
  if (state_off[0] != 0 || D == 0) SGMM_FAIL(c, "set_model: bad state_off");
  uint32_t max_d = 0;
  for (uint32_t s = 0; s < n_states; s++) {
    if (state_off[s + 1] < state_off[s]) SGMM_FAIL(c, "set_model: state_off not monotone");
    max_d = std::max(max_d, state_off[s + 1] - state_off[s]);
  }
  // the per-cell argmin is stored in 16 bits (d_Amin; DensityIdx is uint16_t in the reference too, Types.hpp:17)
  if (max_d > 65535) SGMM_FAIL(c, "set_model: a state has more than 65535 densities (DensityIdx is 16 bit)");
  // variance rows: consecutive densities with bit-identical rows share one (pooled states: one row per state, global
  // pooling: one row); `pooled` = every state's densities share their row
  std::vector<uint32_t> var_of(D);
  uint32_t n_rows = 0;
  bool pooled = true;
  {
    std::vector<uint8_t> first_of_state(D, 0);
    for (uint32_t s = 0; s < n_states; s++) if (state_off[s] < D && state_off[s + 1] > state_off[s]) first_of_state[state_off[s]] = 1;
    for (uint32_t j = 0; j < D; j++) {
      const bool same = j > 0 && memcmp(inv_var + (size_t)j * dim, inv_var + (size_t)(j - 1) * dim, dim * sizeof(double)) == 0;
      if (!same) n_rows++;
      var_of[j] = n_rows - 1;
      if (!same && !first_of_state[j]) pooled = false;
    }
  }
  SgmmModel& m = c->spare;
  sgmm_stage_reset(c);
  sgmm_stage_begin(c, SGMM_T_MSTEP);
  if (model_layout(c, m, dim, n_states, state_off, pooled, var_of.data(), n_rows)) return 1;
  if (upload(c, m.d_mean_src, mean, (size_t)D * dim) || upload(c, m.d_inv_src, inv_var, (size_t)D * dim) ||
      upload(c, m.d_norm, norm, (size_t)D) || upload(c, m.d_logw, log_weight, (size_t)D))
    return 1;
  if (c->host_tables ? model_build_host(c, m, mean, inv_var, norm, log_weight) : model_build_device(c, m)) return 1;
  if (!c->d_fallbacks.p) {
    SGMM_CHECK(c, c->d_fallbacks.reserve(16));   // [0] rescanned cells, [1..15] k_tc_gemm cycle profile (tc_debug & 32)
    SGMM_CHECK(c, cudaMemsetAsync(c->d_fallbacks.p, 0, 16 * sizeof(unsigned long long), c->stream));
  }
  sgmm_stage_end(c, SGMM_T_MSTEP);
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));  // the caller's arrays and the staging vectors may die now
  model_commit(c);
  c->var_valid = false;
  return 0;
}

int sgmm_mstep(sgmm_ctx* c, int pooling, double var_seed) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (c->Dtot == 0) SGMM_FAIL(c, "mstep: no model");
  if (pooling < 0 || pooling > 2) SGMM_FAIL(c, "mstep: pooling must be 0 (none), 1 (mixture) or 2 (global)");
  if (!c->stats_valid || c->stats_model_D != c->Dtot) SGMM_FAIL(c, "mstep: no statistics resident: call sgmm_accumulate first");
  sgmm_stage_reset(c);
  sgmm_stage_begin(c, SGMM_T_MSTEP);
  const uint32_t D = c->Dtot, dim = c->dim, N = c->n_states;
  // variance rows by construction (Mixtures.hpp:20-24): none = one per density, mixture = one per state, global = one
  std::vector<uint32_t> var_of(D);
  uint32_t n_rows = 0;
  if (pooling == 0) { for (uint32_t j = 0; j < D; j++) var_of[j] = j; n_rows = D; }
  else if (pooling == 1) {
    for (uint32_t s = 0; s < N; s++) {
      if (c->h_state_off[s + 1] > c->h_state_off[s]) n_rows++;
      for (uint32_t j = c->h_state_off[s]; j < c->h_state_off[s + 1]; j++) var_of[j] = n_rows - 1;
    }
  } else { n_rows = 1; }
  const bool pooled = pooling != 0 || c->max_dens <= 1;
  SgmmModel& m = c->spare;
  const std::vector<uint32_t> state_off = c->h_state_off;   // (model_layout overwrites the spare's copy, not ours)
  if (model_layout(c, m, dim, N, state_off.data(), pooled, var_of.data(), n_rows)) return 1;
  SGMM_CHECK(c, c->d_mstep_tmp.reserve((size_t)D * dim + D + N + 2 * dim + 8));
  MstepParams p{};
  p.acc = c->d_acc.p; p.state_off = m.d_state_off.p; p.D = D; p.dim = dim; p.n_states = N;
  p.pooling = pooling; p.seed = var_seed; p.dlog2pi = (double)dim * std::log(2 * M_PI);
  p.mean = m.d_mean_src.p; p.inv = m.d_inv_src.p; p.norm = m.d_norm.p; p.logw = m.d_logw.p;
  p.var = c->d_mstep_tmp.p; p.weight = p.var + (size_t)D * dim; p.state_total = p.weight + D; p.gvar = p.state_total + N;
  k_mstep_state<<<N, 512, (dim + 2) * sizeof(double), c->stream>>>(p);   // one block per state: 106 blocks at the bench shape, so wide ones
  SGMM_LAUNCH_CHECK(c);
  if (pooling == 2) {
    k_mstep_global<<<1, 128, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
    k_mstep_global_spread<<<(unsigned)(((size_t)D * dim + 255) / 256), 256, 0, c->stream>>>(p);
    SGMM_LAUNCH_CHECK(c);
  }
  if (model_build_device(c, m)) return 1;
  model_commit(c);
  c->var_valid = true;
  sgmm_stage_end(c, SGMM_T_MSTEP);
  return 0;
}

int sgmm_get_model(sgmm_ctx* c, double* mean, double* inv_var, double* norm, double* log_weight, double* var, double* weight) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (c->Dtot == 0) SGMM_FAIL(c, "get_model: no model");
  const size_t D = c->Dtot, d = c->dim;
  if ((var || weight) && !c->var_valid) SGMM_FAIL(c, "get_model: variances / weights exist only after sgmm_mstep");
  if (mean) SGMM_CHECK(c, cudaMemcpyAsync(mean, c->d_mean_src.p, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (inv_var) SGMM_CHECK(c, cudaMemcpyAsync(inv_var, c->d_inv_src.p, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (norm) SGMM_CHECK(c, cudaMemcpyAsync(norm, c->d_norm.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (log_weight) SGMM_CHECK(c, cudaMemcpyAsync(log_weight, c->d_logw.p, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (var) SGMM_CHECK(c, cudaMemcpyAsync(var, c->d_mstep_tmp.p, D * d * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (weight) SGMM_CHECK(c, cudaMemcpyAsync(weight, c->d_mstep_tmp.p + D * d, D * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// test hook: raw bytes of one derived table of the current model (tests compare the device builder with the host one)
int sgmm_debug_table(sgmm_ctx* c, const char* name, void* out, size_t cap, size_t* bytes) {
  if (!c || !name || !bytes) return 1;
  cudaSetDevice(c->device);
  const void* src = nullptr;
  size_t n = 0;
  const size_t D = c->Dtot, N = c->n_states;
#define SGMM_TAB(key, buf, count) if (!strcmp(name, key)) { src = (buf).p; n = (size_t)(count) * sizeof(*(buf).p); }
  SGMM_TAB("meanT", c->d_meanT, (size_t)c->npair * D) SGMM_TAB("invT", c->d_invT, (size_t)c->npair * D)
  SGMM_TAB("mean_tail", c->d_mean_tail, D) SGMM_TAB("inv_tail", c->d_inv_tail, D)
  SGMM_TAB("meanR", c->d_meanR, D * c->dimP) SGMM_TAB("invU", c->d_invU, (size_t)c->n_var_rows * c->dimP)
  SGMM_TAB("var_of", c->d_var_of, D) SGMM_TAB("norm", c->d_norm, D) SGMM_TAB("logw", c->d_logw, D)
  SGMM_TAB("mu32", c->d_mu32, c->dv4 ? D * c->dv4 : 0) SGMM_TAB("h32", c->d_h32, c->dv4 ? D * c->dv4 : 0)
  SGMM_TAB("mu32T", c->d_mu32T, c->dv4 ? D * c->dv4 : 0) SGMM_TAB("c32", c->d_c32, c->dv4 ? D : 0)
  SGMM_TAB("sguard", c->d_sguard, c->dv4 ? N : 0)
  SGMM_TAB("Bimg", c->d_Bimg, c->tc_ok ? c->img_elems : 0) SGMM_TAB("sguard_tc", c->d_sguard_tc, c->tc_ok ? N : 0)
  SGMM_TAB("Bimg16", c->d_Bimg16, c->tc_f16_ok ? c->img16_elems : 0) SGMM_TAB("sguard16", c->d_sguard16, c->tc_f16_ok ? N : 0)
  SGMM_TAB("tile_info", c->d_tile_info, c->tc_ok ? c->h_tile_first.back() : 0)
  SGMM_TAB("col_off", c->d_col_off, c->tc_ok ? N + 1 : 0)
#undef SGMM_TAB
  if (!src && n == 0 && strcmp(name, "mu32") && strcmp(name, "h32") && strcmp(name, "mu32T") && strcmp(name, "c32") &&
      strcmp(name, "sguard") && strcmp(name, "Bimg") && strcmp(name, "sguard_tc") && strcmp(name, "Bimg16") &&
      strcmp(name, "sguard16") && strcmp(name, "tile_info") && strcmp(name, "col_off"))
    SGMM_FAIL(c, std::string("debug_table: unknown table ") + name);
  *bytes = n;
  if (!out || n == 0) return 0;
  if (cap < n) SGMM_FAIL(c, "debug_table: buffer too small");
  SGMM_CHECK(c, cudaMemcpyAsync(out, src, n, cudaMemcpyDeviceToHost, c->stream));
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
