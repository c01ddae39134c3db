// K4: deterministic EM sufficient statistics — the scatter loop of MixtureModel::accumulate
// (reference src/sietill/Mixtures.cpp:333-357) as  stable sort by density  ->  fixed-shape segmented sum.
//
//  1. stable LSD radix sort (8-bit digits) of frame ids by the density each frame was assigned to; one
//     warp owns one contiguous chunk, ranks are computed with warp match/ballot and prefix sums — no
//     atomics decide a position, so the permutation is the unique stable one.
//  2. every density's (frame-ordered) run is cut into blocks of SGMM_RED_BLOCK frames; one warp sums a
//     block sequentially in frame order (lane = feature dimension, f64), then one warp per density adds
//     the block partials in block order.  The summation tree depends only on the data, never on
//     scheduling: bit-reproducible run to run and independent of the GPU count per shard.
//  Counts are integers (atomicAdd on u32 in K3), exact by construction.
#include "sgmm_internal.cuh"

namespace {

constexpr int kSortWarps = 8;

__global__ void __launch_bounds__(kSortWarps * 32) k_sort_hist(const uint32_t* __restrict__ keys, uint64_t n,
                                                              int shift, uint32_t n_chunks,
                                                              uint32_t* __restrict__ hist) {
  __shared__ uint32_t h[kSortWarps][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * kSortWarps + warp;
  for (int b = lane; b < 256; b += 32) h[warp][b] = 0;
  __syncwarp();
  if (chunk < n_chunks) {
    const uint64_t beg = (uint64_t)chunk * SGMM_SORT_CHUNK;
    const uint64_t end = min(n, beg + SGMM_SORT_CHUNK);
    for (uint64_t i = beg; i < end; i += 32) {
      const bool valid = i + lane < end;
      const uint32_t d = valid ? ((keys[i + lane] >> shift) & 255u) : (256u + lane);
      const uint32_t peers = __match_any_sync(0xffffffffu, d);
      if (valid && lane == __ffs(peers) - 1) h[warp][d] += __popc(peers);
      __syncwarp();
    }
    for (int b = lane; b < 256; b += 32) hist[(size_t)b * n_chunks + chunk] = h[warp][b];
  }
}

// Exclusive prefix sum of n u32 values by ONE block (n is at most a few hundred thousand here): tiles of 4096
// values, four consecutive ones per thread (coalesced 16-byte accesses), warp-shuffle scans, a running carry.
// mode 0: v = in[i];  mode 1: v = ceil(in[i] / SGMM_RED_BLOCK).  out has n + 1 entries (out[n] = total).
__global__ void __launch_bounds__(1024) k_excl_scan(const uint32_t* in, uint32_t* out,  // may alias
                                                    uint32_t n, int mode) {
  __shared__ uint32_t wexcl[32];     // exclusive scan of the warp totals of the current tile
  __shared__ uint32_t tile_total;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t carry = 0;
  for (uint32_t base = 0; base < n; base += 4096) {
    const uint32_t i = base + tid * 4;
    uint32_t v[4] = {0, 0, 0, 0};
    if (i + 3 < n) {
      const uint4 q = *reinterpret_cast<const uint4*>(in + i);
      v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    } else {
      for (uint32_t j = 0; j < 4; j++) if (i + j < n) v[j] = in[i + j];
    }
    if (mode)
      for (uint32_t j = 0; j < 4; j++) v[j] = (v[j] + SGMM_RED_BLOCK - 1) / SGMM_RED_BLOCK;
    const uint32_t t = v[0] + v[1] + v[2] + v[3];
    uint32_t inc = t;                                         // inclusive scan of the thread totals inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
      if ((int)lane >= o) inc += u;
    }
    if (lane == 31) wexcl[warp] = inc;                        // warp total
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = wexcl[lane];
      uint32_t winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, winc, o);
        if ((int)lane >= o) winc += u;
      }
      wexcl[lane] = winc - w;
      if (lane == 31) tile_total = winc;
    }
    __syncthreads();
    uint32_t run = carry + wexcl[warp] + (inc - t);
    if (i + 3 < n) {
      uint4 q;
      q.x = run; run += v[0]; q.y = run; run += v[1]; q.z = run; run += v[2]; q.w = run;
      *reinterpret_cast<uint4*>(out + i) = q;
    } else {
      for (uint32_t j = 0; j < 4; j++) if (i + j < n) { out[i + j] = run; run += v[j]; }
    }
    carry += tile_total;
    __syncthreads();                                          // wexcl / tile_total are rewritten by the next tile
  }
  if (tid == 0) out[n] = carry;
}

__global__ void __launch_bounds__(kSortWarps * 32) k_sort_scatter(const uint32_t* __restrict__ keys_in,
                                                                 const uint32_t* __restrict__ vals_in,  // NULL: iota
                                                                 uint32_t* __restrict__ keys_out,
                                                                 uint32_t* __restrict__ vals_out, uint64_t n,
                                                                 int shift, uint32_t n_chunks,
                                                                 const uint32_t* __restrict__ base) {
  __shared__ uint32_t off[kSortWarps][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * kSortWarps + warp;
  if (chunk >= n_chunks) return;
  for (int b = lane; b < 256; b += 32) off[warp][b] = base[(size_t)b * n_chunks + chunk];
  __syncwarp();
  const uint64_t beg = (uint64_t)chunk * SGMM_SORT_CHUNK;
  const uint64_t end = min(n, beg + SGMM_SORT_CHUNK);
  for (uint64_t i = beg; i < end; i += 32) {
    const bool valid = i + lane < end;
    const uint32_t key = valid ? keys_in[i + lane] : 0;
    const uint32_t val = valid ? (vals_in ? vals_in[i + lane] : (uint32_t)(i + lane)) : 0;
    const uint32_t d = valid ? ((key >> shift) & 255u) : (256u + lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    const uint32_t b0 = valid ? off[warp][d] : 0;
    __syncwarp();
    if (valid && lane == __ffs(peers) - 1) off[warp][d] = b0 + __popc(peers);
    __syncwarp();
    if (valid) {
      keys_out[b0 + rank] = key;
      vals_out[b0 + rank] = val;
    }
  }
}

// One warp per (density, block) item: sequential f64 sums of x and x*x over the block's frames.
template <int ND>
__global__ void __launch_bounds__(128) k_seg_partial(const float* __restrict__ x, uint32_t dim,
                                                     const uint32_t* __restrict__ sorted_frames,
                                                     const uint32_t* __restrict__ seg_off,
                                                     const uint32_t* __restrict__ item_off, uint32_t Dtot,
                                                     double* __restrict__ partial) {
  const int lane = threadIdx.x & 31;
  const uint32_t item = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (item >= item_off[Dtot]) return;
  // density owning this item: largest k with item_off[k] <= item
  uint32_t lo = 0, hi = Dtot;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (item_off[mid] <= item) lo = mid; else hi = mid;
  }
  const uint32_t k = lo;
  const uint32_t b = item - item_off[k];
  const uint32_t beg = seg_off[k] + b * SGMM_RED_BLOCK;
  const uint32_t end = min(seg_off[k + 1], beg + SGMM_RED_BLOCK);
  double sx[ND], sxx[ND];
#pragma unroll
  for (int q = 0; q < ND; q++) { sx[q] = 0.0; sxx[q] = 0.0; }
  for (uint32_t i0 = beg; i0 < end; i0 += 32) {
    const uint32_t nf = min(32u, end - i0);
    const uint32_t myf = lane < (int)nf ? sorted_frames[i0 + lane] : 0;
    for (uint32_t j0 = 0; j0 < nf; j0 += 8) {
      float v[8][ND];
#pragma unroll
      for (int jj = 0; jj < 8; jj++) {
        const uint32_t f = __shfl_sync(0xffffffffu, myf, (j0 + jj) & 31);
#pragma unroll
        for (int q = 0; q < ND; q++) {
          const uint32_t i = lane + 32 * q;
          v[jj][q] = (j0 + jj < nf && i < dim) ? __ldg(x + (uint64_t)f * dim + i) : 0.f;
        }
      }
#pragma unroll
      for (int jj = 0; jj < 8; jj++) {
        if (j0 + jj < nf) {
#pragma unroll
          for (int q = 0; q < ND; q++) {
            const double y = (double)v[jj][q];
            sx[q] = __dadd_rn(sx[q], y);                 // scale_add with weight 1, Mixtures.cpp:46-56
            sxx[q] = __dadd_rn(sxx[q], __dmul_rn(y, y)); // scale_add_square, Mixtures.cpp:58-68
          }
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < ND; q++) {
    const uint32_t i = lane + 32 * q;
    if (i < dim) {
      partial[(size_t)item * 2 * dim + i] = sx[q];
      partial[(size_t)item * 2 * dim + dim + i] = sxx[q];
    }
  }
}

// One warp per density: add its block partials in block order; write count / sum_x / sum_xx.
__global__ void __launch_bounds__(128) k_seg_final(const double* __restrict__ partial, uint32_t dim,
                                                   const uint32_t* __restrict__ cnt,
                                                   const uint32_t* __restrict__ item_off, uint32_t Dtot,
                                                   double* __restrict__ acc) {
  const int lane = threadIdx.x & 31;
  const uint32_t k = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (k >= Dtot) return;
  const uint32_t i0 = item_off[k], i1 = item_off[k + 1];
  double* count = acc;
  double* sum_x = acc + Dtot;
  double* sum_xx = acc + Dtot + (size_t)Dtot * dim;
  if (lane == 0) count[k] = (double)cnt[k];
  for (uint32_t i = lane; i < dim; i += 32) {
    double a = 0.0, b = 0.0;
    for (uint32_t it = i0; it < i1; it++) {
      a = __dadd_rn(a, partial[(size_t)it * 2 * dim + i]);
      b = __dadd_rn(b, partial[(size_t)it * 2 * dim + dim + i]);
    }
    sum_x[(size_t)k * dim + i] = a;
    sum_xx[(size_t)k * dim + i] = b;
  }
}

// Deterministic sum of the per-frame scores: fixed chunking, fixed tree.
constexpr int kScoreBlocks = 296;
__global__ void __launch_bounds__(256) k_score_partial(const double* __restrict__ fscore, uint64_t n,
                                                       double* __restrict__ part) {
  __shared__ double sh[256];
  const uint64_t per = (n + gridDim.x - 1) / gridDim.x;
  const uint64_t beg = min(n, (uint64_t)blockIdx.x * per), end = min(n, beg + per);
  double a = 0.0;
  for (uint64_t i = beg + threadIdx.x; i < end; i += 256) a = __dadd_rn(a, fscore[i]);
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] = __dadd_rn(sh[threadIdx.x], sh[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}
// out[0] = the score sum, out[1] = the number of frames it covers (both travel in the all-reduce payload, so that
// calc_am_score divides the global sum by the global frame count)
__global__ void k_score_final(const double* __restrict__ part, int n, double n_frames, double* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double a = 0.0;
    for (int i = 0; i < n; i++) a = __dadd_rn(a, part[i]);
    out[0] = a;
    out[1] = n_frames;
  }
}

}  // namespace

// Stable sort of the frame ids 0 .. n_frames-1 by keys[f] in [0, n_keys): *sorted receives the permutation (device).
int sgmm_sort_frames(sgmm_ctx* c, const uint32_t* keys, uint32_t n_keys, const uint32_t** sorted) {
  const uint64_t n = c->n_frames;
  if (n >= (1ull << 32)) SGMM_FAIL(c, "accumulate: more than 2^32 frames per context");
  const uint32_t n_chunks = (uint32_t)((n + SGMM_SORT_CHUNK - 1) / SGMM_SORT_CHUNK);
  SGMM_CHECK(c, c->d_keys[0].reserve(n)); SGMM_CHECK(c, c->d_keys[1].reserve(n));
  SGMM_CHECK(c, c->d_vals[0].reserve(n)); SGMM_CHECK(c, c->d_vals[1].reserve(n));
  SGMM_CHECK(c, c->d_hist.reserve((size_t)256 * n_chunks + 1));
  int bits = 1;
  while ((1ull << bits) < (uint64_t)n_keys) bits++;
  const int passes = (bits + 7) / 8;
  const uint32_t sort_grid = (n_chunks + kSortWarps - 1) / kSortWarps;
  const uint32_t* keys_in = keys;
  const uint32_t* vals_in = nullptr;
  int cur = 0;
  StageTimer st(c, SGMM_T_SORT);
  for (int pass = 0; pass < passes && n > 0; pass++) {
    const int shift = 8 * pass;
    k_sort_hist<<<sort_grid, kSortWarps * 32, 0, c->stream>>>(keys_in, n, shift, n_chunks, c->d_hist.p);
    SGMM_LAUNCH_CHECK(c);
    k_excl_scan<<<1, 1024, 0, c->stream>>>(c->d_hist.p, c->d_hist.p, 256 * n_chunks, 0);
    SGMM_LAUNCH_CHECK(c);
    k_sort_scatter<<<sort_grid, kSortWarps * 32, 0, c->stream>>>(keys_in, vals_in, c->d_keys[cur].p,
                                                                c->d_vals[cur].p, n, shift, n_chunks,
                                                                c->d_hist.p);
    SGMM_LAUNCH_CHECK(c);
    keys_in = c->d_keys[cur].p;
    vals_in = c->d_vals[cur].p;
    cur ^= 1;
  }
  *sorted = vals_in;
  return 0;
}

// Sorts frames by c->d_fdens, reduces into c->d_acc = [count | sum_x | sum_xx | score_sum].
int sgmm_launch_accumulate(sgmm_ctx* c) {
  const uint64_t n = c->n_frames;
  const uint32_t D = c->Dtot, dim = c->dim;
  SGMM_CHECK(c, c->d_seg_off.reserve(D + 2)); SGMM_CHECK(c, c->d_item_off.reserve(D + 1));
  const size_t max_items = n / SGMM_RED_BLOCK + D + 1;
  SGMM_CHECK(c, c->d_partial.reserve(max_items * 2 * dim));
  const uint32_t* vals_in = nullptr;
  if (sgmm_sort_frames(c, c->d_fdens.p, D + 1, &vals_in)) return 1;   // keys are 0..D (D = dummy slot of empty mixtures)
  {
    StageTimer st(c, SGMM_T_REDUCE);
    k_excl_scan<<<1, 1024, 0, c->stream>>>(c->d_cnt_u32.p, c->d_seg_off.p, D + 1, 0);
    SGMM_LAUNCH_CHECK(c);
    k_excl_scan<<<1, 1024, 0, c->stream>>>(c->d_cnt_u32.p, c->d_item_off.p, D, 1);
    SGMM_LAUNCH_CHECK(c);
    if (n > 0) {
      const uint32_t grid = (uint32_t)((max_items + 3) / 4);
      const int nd = (dim + 31) / 32;
      if (nd == 1) k_seg_partial<1><<<grid, 128, 0, c->stream>>>(c->d_x.p, dim, vals_in, c->d_seg_off.p, c->d_item_off.p, D, c->d_partial.p);
      else if (nd == 2) k_seg_partial<2><<<grid, 128, 0, c->stream>>>(c->d_x.p, dim, vals_in, c->d_seg_off.p, c->d_item_off.p, D, c->d_partial.p);
      else if (nd <= 4) k_seg_partial<4><<<grid, 128, 0, c->stream>>>(c->d_x.p, dim, vals_in, c->d_seg_off.p, c->d_item_off.p, D, c->d_partial.p);
      else SGMM_FAIL(c, "accumulate: feature dimension > 128 not supported");
      SGMM_LAUNCH_CHECK(c);
    }
    k_seg_final<<<(D + 3) / 4, 128, 0, c->stream>>>(c->d_partial.p, dim, c->d_cnt_u32.p, c->d_item_off.p, D, c->d_acc.p);
    SGMM_LAUNCH_CHECK(c);
  }
  return 0;
}

// d_acc[D (1 + 2 dim)] = sum of c->d_fscore, the slot after it = the frame count
int sgmm_launch_score_sum(sgmm_ctx* c) {
  SGMM_CHECK(c, c->d_score_part.reserve(kScoreBlocks));
  double* out = c->d_acc.p + (size_t)c->Dtot * (1 + 2 * (size_t)c->dim);
  k_score_partial<<<kScoreBlocks, 256, 0, c->stream>>>(c->d_fscore.p, c->n_frames, c->d_score_part.p);
  SGMM_LAUNCH_CHECK(c);
  k_score_final<<<1, 32, 0, c->stream>>>(c->d_score_part.p, kScoreBlocks, (double)c->n_frames, out);
  SGMM_LAUNCH_CHECK(c);
  return 0;
}
