// ipoke_inn_flow_loss_static: the multi-CTA FlowLoss reduction (variant of inn_flow_loss_kernel, elementwise.cuh) with
// its scratch memory in a file-scope __device__ array, for callers that must not allocate device memory (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Declared in include/ipoke_inn.h.
//
// This file calls no cudaMalloc / cudaFree / allocating library routine.  The scratch array exists once per GPU that
// loads the library (6400 bytes, held until the process exits) and is SHARED by every call on that GPU: it keeps the last
// call's contents, the ticket word is cleared on the caller's stream before every launch, and two calls that overlap
// on the same GPU (different streams, or different host threads) race on it.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <mutex>

#include "ipoke_inn.h"

namespace {

constexpr int kStaticMaxBlocks = IPOKE_INN_FLOW_LOSS_STATIC_MAX_BLOCKS;   // 256 CTAs of 1024 threads
constexpr int kElemsPerBlock = 65536;                                      // 64 elements per thread
constexpr long long kStaticMaxElems = IPOKE_INN_FLOW_LOSS_STATIC_MAX_ELEMS;
static_assert(kStaticMaxElems == (long long)kStaticMaxBlocks * kElemsPerBlock, "limit = CTAs x elements per CTA");
constexpr int kTicketSlot = 3 * kStaticMaxBlocks;                          // one double-sized slot after the partials
constexpr int kScratchDoubles = 3 * kStaticMaxBlocks + 32;                 // 800 doubles = 6400 bytes

// In host code this name is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) double g_flow_loss_scratch[kScratchDoubles];

// Same arithmetic as inn_flow_loss_kernel: per-thread fp64 sums, fixed-order tree per CTA, fp64 partials parked in
// scratch[3 * cta ..], the CTA that draws the last ticket adds them in CTA order (deterministic for a given grid).
__global__ void __launch_bounds__(1024) inn_flow_loss_static_kernel(const float* __restrict__ z,
                                                                    const float* __restrict__ logdet, int B, int C,
                                                                    float nll_w, float ld_w, float* out4, float* dz,
                                                                    float* dlogdet, const float* __restrict__ ref,
                                                                    double* scratch) {
  __shared__ double r_a[1024];
  __shared__ double r_b[1024];
  __shared__ double r_c[1024];
  __shared__ unsigned int s_ticket;
  double a = 0.0, l = 0.0, r = 0.0;
  const long long total = static_cast<long long>(B) * C;
  const float inv_b = 1.0f / static_cast<float>(B);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const float v = z[i];
    a += 0.5 * static_cast<double>(v) * v;
    if (dz) dz[i] = nll_w * v * inv_b;
    if (ref) {
      const float w = ref[i];
      r += 0.5 * static_cast<double>(w) * w;
    }
  }
  for (long long b = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; b < B; b += stride) {
    l += logdet[b];
    if (dlogdet) dlogdet[b] = -ld_w * inv_b;
  }
  r_a[threadIdx.x] = a;
  r_b[threadIdx.x] = l;
  r_c[threadIdx.x] = r;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      r_a[threadIdx.x] += r_a[threadIdx.x + s];
      r_b[threadIdx.x] += r_b[threadIdx.x + s];
      r_c[threadIdx.x] += r_c[threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    scratch[3 * blockIdx.x] = r_a[0];
    scratch[3 * blockIdx.x + 1] = r_b[0];
    scratch[3 * blockIdx.x + 2] = r_c[0];
    __threadfence();
    s_ticket = atomicAdd(reinterpret_cast<unsigned int*>(scratch + kTicketSlot), 1u);
  }
  __syncthreads();
  if (s_ticket != gridDim.x - 1 || threadIdx.x != 0) return;
  __threadfence();
  double sa = 0.0, sl = 0.0, sr = 0.0;
  for (unsigned int k = 0; k < gridDim.x; ++k) {
    sa += scratch[3 * k];
    sl += scratch[3 * k + 1];
    sr += scratch[3 * k + 2];
  }
  const double nll = sa / B, nld = -sl / B;
  out4[0] = static_cast<float>(nll_w * nll + ld_w * nld);
  out4[1] = static_cast<float>(nll);
  out4[2] = static_cast<float>(nld);
  if (ref) out4[3] = static_cast<float>(sr / B);
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
double* g_scratch_ptr[kMaxDevices] = {};   // device address of g_flow_loss_scratch, per device ordinal

// Once per device, with that device current: resolve the symbol's device address.  Allocates nothing.
int scratch_for_current_device(double** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return IPOKE_INN_E_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    double* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_flow_loss_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return IPOKE_INN_E_INTERNAL;
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return IPOKE_INN_OK;
}

}  // namespace

extern "C" int ipoke_inn_flow_loss_static(const float* z, const float* logdet, const float* ref_sample, int B, int C,
                                          float nll_weight, float logdet_weight, float* out4, float* dz, float* dlogdet,
                                          void* stream) {
  if (!z || !logdet || !out4 || B <= 0 || C <= 0) return IPOKE_INN_E_INVALID;
  const long long total = static_cast<long long>(B) * C;
  if (total > kStaticMaxElems) return IPOKE_INN_E_TOO_LARGE;   // beyond what the fixed scratch array is sized for
  double* scratch = nullptr;
  if (int rc = scratch_for_current_device(&scratch)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the ticket must start at zero: the array keeps whatever the last call (possibly an aborted one) left behind
  cudaError_t e = cudaMemsetAsync(scratch + kTicketSlot, 0, sizeof(double), s);
  if (e != cudaSuccess) return (int)e;
  const int blocks = (int)std::max<long long>(1, std::min<long long>(kStaticMaxBlocks, (total + kElemsPerBlock - 1) / kElemsPerBlock));
  inn_flow_loss_static_kernel<<<blocks, 1024, 0, s>>>(z, logdet, B, C, nll_weight, logdet_weight, out4, dz, dlogdet,
                                                      ref_sample, scratch);
  e = cudaGetLastError();
  return e == cudaSuccess ? IPOKE_INN_OK : (int)e;
}
