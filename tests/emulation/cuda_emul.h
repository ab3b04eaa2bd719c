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

// Context switch between fibers.  glibc's swapcontext saves / restores the signal mask with two system calls per
// switch, which dominated the tests' run time; on x86-64 a fiber switch only needs the callee-saved registers and the
// stack pointer (System V ABI), so a dozen instructions replace it there.  Other hosts keep ucontext.
#if defined(__x86_64__) && !defined(EMUL_FORCE_UCONTEXT)
#define EMUL_FAST_SWITCH 1
extern "C" void emul_switch_stack(void** save_sp, void* load_sp);
asm(R"(
    .text
    .p2align 4
    .globl emul_switch_stack
    .type emul_switch_stack, @function
emul_switch_stack:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
    .size emul_switch_stack, .-emul_switch_stack
)");
#else
#define EMUL_FAST_SWITCH 0
#endif

struct Runner {
#if EMUL_FAST_SWITCH
  std::vector<void*> sp;          // saved stack pointer of every fiber
  void* main_sp = nullptr;
#else
  std::vector<ucontext_t> ctx;
  ucontext_t main;
#endif
  char* stacks = nullptr;          // nthreads * kStackBytes, from a pool that lives as long as the library
  std::vector<char> done;
  int cur = 0;
  const std::function<void()>* body = nullptr;
};
static Runner* g_runner = nullptr;

// Fiber stacks: one allocation reused by every CTA (a fresh 64 - 128 MB vector per CTA spent the time in page faults).
static char* stack_pool(size_t bytes) {
  static char* pool = nullptr;
  static size_t size = 0;
  if (bytes > size) {
    std::free(pool);
    pool = static_cast<char*>(std::malloc(bytes));
    size = bytes;
  }
  return pool;
}

static inline void to_scheduler(Runner* r) {
#if EMUL_FAST_SWITCH
  emul_switch_stack(&r->sp[r->cur], r->main_sp);
#else
  swapcontext(&r->ctx[r->cur], &r->main);
#endif
}

static void trampoline() {
  Runner* r = g_runner;
  (*r->body)();
  r->done[r->cur] = 1;
  to_scheduler(r);                 // never resumed
  abort();
}

static inline void barrier() { to_scheduler(g_runner); }

// One CTA of `nthreads` threads running `body` (the kernel call) to completion.
static void run_block(int nthreads, const std::function<void()>& body) {
  Runner r;
  r.stacks = stack_pool((size_t)nthreads * kStackBytes + 64);
  r.done.assign(nthreads, 0);
  r.body = &body;
  g_runner = &r;
  blockDim = {(unsigned)nthreads, 1, 1};
#if EMUL_FAST_SWITCH
  r.sp.resize(nthreads);
  for (int t = 0; t < nthreads; ++t) {
    // initial frame, low to high: r15 r14 r13 r12 rbx rbp | return address = trampoline | a null caller; the top is
    // 16-byte aligned, so the trampoline starts with rsp = top - 8 like any function entered by a call
    uintptr_t top = reinterpret_cast<uintptr_t>(r.stacks + (size_t)(t + 1) * kStackBytes) & ~(uintptr_t)15;
    void** frame = reinterpret_cast<void**>(top) - 8;
    for (int i = 0; i < 6; ++i) frame[i] = nullptr;
    frame[6] = reinterpret_cast<void*>(&trampoline);
    frame[7] = nullptr;
    r.sp[t] = frame;
  }
#else
  r.ctx.resize(nthreads);
  for (int t = 0; t < nthreads; ++t) {
    getcontext(&r.ctx[t]);
    r.ctx[t].uc_stack.ss_sp = r.stacks + (size_t)t * kStackBytes;
    r.ctx[t].uc_stack.ss_size = kStackBytes;
    r.ctx[t].uc_link = &r.main;
    makecontext(&r.ctx[t], trampoline, 0);
  }
#endif
  bool alive = true;
  while (alive) {
    alive = false;
    for (int i = 0; i < nthreads; ++i) {
      const int t = descending ? nthreads - 1 - i : i;
      if (r.done[t]) continue;
      r.cur = t;
      threadIdx = {(unsigned)t, 0, 0};
#if EMUL_FAST_SWITCH
      emul_switch_stack(&r.main_sp, r.sp[t]);
#else
      swapcontext(&r.main, &r.ctx[t]);
#endif
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
