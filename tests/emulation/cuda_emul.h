// TEST INFRASTRUCTURE: a small CPU emulation of the CUDA execution model for the MaCow kernels.
// Every thread of a CTA is a fiber (ucontext); __syncthreads() yields to a scheduler that resumes the fibers of the
// block one after the other, in ascending or descending thread order (emul::descending) -- so a missing barrier
// between a shared-memory write and a read by another thread shows up as a wrong result in one of the two orders.
// CTAs run one after the other.  Warp shuffles go through a block-wide exchange (valid when every thread of the block
// executes the shuffle, which is how the kernels use it).  Never linked into libipoke_inn.so.
#pragma once
#include <ucontext.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define MACOW_HOST_EMULATION 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __shared__
#define __align__(n)
#define __launch_bounds__(...)

struct float4 { float x, y, z, w; };
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
struct float2 { float x, y; };
struct dim3e { unsigned x, y, z; };
static dim3e threadIdx = {0, 0, 0}, blockDim = {1, 1, 1}, blockIdx = {0, 0, 0}, gridDim = {1, 1, 1};
template <typename T> static inline T __ldg(const T* p) { return *p; }
using std::max;
using std::min;

namespace emul {

constexpr size_t kStackBytes = 256 * 1024;
static bool descending = false;
static int grid_z = 0;                    // blockIdx.z of the CTAs a launch() call runs (set by the caller)

struct Runner {
  std::vector<ucontext_t> ctx;
  std::vector<char> stacks;
  std::vector<char> done;
  ucontext_t main;
  int cur = 0;
  const std::function<void()>* body = nullptr;
};
static Runner* g_runner = nullptr;

static void trampoline() {
  Runner* r = g_runner;
  (*r->body)();
  r->done[r->cur] = 1;
  swapcontext(&r->ctx[r->cur], &r->main);
}

static inline void barrier() {
  Runner* r = g_runner;
  swapcontext(&r->ctx[r->cur], &r->main);
}

// One CTA of `nthreads` threads running `body` (the kernel call) to completion.
static void run_block(int nthreads, const std::function<void()>& body) {
  Runner r;
  r.ctx.resize(nthreads);
  r.stacks.resize((size_t)nthreads * kStackBytes);
  r.done.assign(nthreads, 0);
  r.body = &body;
  g_runner = &r;
  blockDim = {(unsigned)nthreads, 1, 1};
  for (int t = 0; t < nthreads; ++t) {
    getcontext(&r.ctx[t]);
    r.ctx[t].uc_stack.ss_sp = r.stacks.data() + (size_t)t * kStackBytes;
    r.ctx[t].uc_stack.ss_size = kStackBytes;
    r.ctx[t].uc_link = &r.main;
    makecontext(&r.ctx[t], trampoline, 0);
  }
  bool alive = true;
  while (alive) {
    alive = false;
    for (int i = 0; i < nthreads; ++i) {
      const int t = descending ? nthreads - 1 - i : i;
      if (r.done[t]) continue;
      r.cur = t;
      threadIdx = {(unsigned)t, 0, 0};
      swapcontext(&r.main, &r.ctx[t]);
      if (!r.done[t]) alive = true;
    }
  }
  g_runner = nullptr;
}

// kernel<<<dim3(gx, gy), nthreads>>>: `call` invokes the kernel function with its parameters.
static void launch(int gx, int gy, int nthreads, const std::function<void()>& call) {
  gridDim = {(unsigned)gx, (unsigned)gy, 1};
  for (int by = 0; by < gy; ++by) {
    for (int bx = 0; bx < gx; ++bx) {
      blockIdx = {(unsigned)bx, (unsigned)by, (unsigned)grid_z};
      run_block(nthreads, call);
    }
  }
}

static float g_shfl[1024];

}  // namespace emul

static inline void __syncthreads() { emul::barrier(); }
static inline float __shfl_xor_sync(unsigned, float v, int o) {
  const unsigned t = threadIdx.x;
  emul::g_shfl[t] = v;
  emul::barrier();
  threadIdx.x = t;                                   // the scheduler restored it already; kept for clarity
  const unsigned partner = t ^ (unsigned)o;
  const float r = partner < blockDim.x ? emul::g_shfl[partner] : v;
  emul::barrier();
  return r;
}
