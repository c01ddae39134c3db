// sgmm_score_sum_static — an allocation-free variant of the deterministic score sum (k_score_partial / k_score_final in
// k_accum.cu, the numerator of Trainer::calc_am_score) for callers that cannot call cudaMalloc: code under CUDA graph
// capture, tight serving loops, harnesses that forbid allocation.  No counterpart in the reference.
// Nothing here touches an existing code path.
//
// The partial sums live in a FILE-SCOPE device array of fixed size instead of a cudaMalloc'ed buffer:
//   * kStaticChunk = 4096 values per partial, kStaticMaxChunks = 16384 partials  ->  at most 67 108 864 values per call,
//     128 KB of device memory that every GPU which loads this library holds until the process exits;
//   * every call shares the array and finds the previous call's partials in it.  Nothing has to start at zero: a call
//     with n values overwrites partials [0, ceil(n / 4096)) and reads back exactly those, so stale entries are never
//     read (the two-calls-in-a-row test covers a long call followed by a short one);
//   * calls that OVERLAP on the GPU (two streams, two host threads, two graphs replayed concurrently) race on the
//     array.  Serialise them.
// In host code the array's name is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current, and is what the kernels are given.
#include <mutex>

#include "sgmm_internal.cuh"

namespace {

constexpr uint32_t kStaticChunk = 4096;
constexpr uint32_t kStaticMaxChunks = 16384;
constexpr uint64_t kStaticMaxValues = (uint64_t)kStaticChunk * kStaticMaxChunks;

static __device__ __align__(256) double g_static_partials[kStaticMaxChunks];

constexpr int kMaxDevices = 64;
double* g_partials_addr[kMaxDevices] = {nullptr};   // device address of g_static_partials on every device seen so far
std::mutex g_partials_mutex;

// partial c = sum of values [4096 c, 4096 (c + 1)): 256 threads x 16 strided values, then a fixed shared-memory tree
__global__ void __launch_bounds__(256) k_static_partial(const double* __restrict__ v, uint64_t n, double* __restrict__ part) {
  __shared__ double sh[256];
  const uint64_t beg = (uint64_t)blockIdx.x * kStaticChunk;
  const uint64_t end = min(n, beg + kStaticChunk);
  double a = 0.0;
  for (uint64_t i = beg + threadIdx.x; i < end; i += 256) a = __dadd_rn(a, v[i]);
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] = __dadd_rn(sh[threadIdx.x], sh[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

// out[0] = sum of partials [0, n_part): thread t adds the contiguous run [16 t, 16 t + 16) in order, then the same tree
__global__ void __launch_bounds__(1024) k_static_final(const double* __restrict__ part, uint32_t n_part, double* __restrict__ out) {
  __shared__ double sh[1024];
  const uint32_t per = kStaticMaxChunks / 1024;
  double a = 0.0;
  for (uint32_t i = threadIdx.x * per; i < min(n_part, (threadIdx.x + 1) * per); i++) a = __dadd_rn(a, part[i]);
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] = __dadd_rn(sh[threadIdx.x], sh[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0];
}

}  // namespace

extern "C" int sgmm_score_sum_static(sgmm_ctx* c, const double* d_values, uint64_t n, double* d_sum, void* stream) {
  if (!c) return 1;
  if (!d_sum || (!d_values && n)) SGMM_FAIL(c, "score_sum_static: null argument");
  if (n > kStaticMaxValues)
    SGMM_FAIL(c, "score_sum_static: " + std::to_string(n) + " values exceed the static scratch (at most " +
                     std::to_string(kStaticMaxValues) + " per call); use sgmm_am_score_sum or split the call");
  if (c->device < 0 || c->device >= kMaxDevices) SGMM_FAIL(c, "score_sum_static: device index out of range");
  SGMM_CHECK(c, cudaSetDevice(c->device));
  double* part = nullptr;
  {
    // set-up, once per device, with that device current: the symbol's DEVICE address (the bare name is a host stand-in)
    std::lock_guard<std::mutex> lock(g_partials_mutex);
    if (!g_partials_addr[c->device]) {
      void* p = nullptr;
      SGMM_CHECK(c, cudaGetSymbolAddress(&p, g_static_partials));
      if (!p) SGMM_FAIL(c, "score_sum_static: cudaGetSymbolAddress returned a null address");
      g_partials_addr[c->device] = static_cast<double*>(p);
    }
    part = g_partials_addr[c->device];
  }
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : c->stream;
  const uint32_t n_part = (uint32_t)((n + kStaticChunk - 1) / kStaticChunk);
  if (n_part) {
    k_static_partial<<<n_part, 256, 0, st>>>(d_values, n, part);
    SGMM_LAUNCH_CHECK(c);
  }
  k_static_final<<<1, 1024, 0, st>>>(part, n_part, d_sum);
  SGMM_LAUNCH_CHECK(c);
  return 0;   // enqueued, not synchronised: the result is in *d_sum once `stream` has passed this point
}
