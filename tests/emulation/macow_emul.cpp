// TEST INFRASTRUCTURE: compiles csrc/macow_kernel.cuh with g++ as a ONE-THREAD-PER-CTA emulation (threadIdx = 0,
// blockDim = 1, __syncthreads a no-op), so the index arithmetic of the pack and unit kernels -- tap offsets, weight
// transposes, tile order, stage sequencing -- can be checked against the oracle on the CPU-only build container.
// Never linked into libipoke_inn.so; built by tests/test_macow_emulation.py into tests/emulation/_build/.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

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
static inline void __syncthreads() {}
template <typename T> static inline T __ldg(const T* p) { return *p; }
static inline float __shfl_xor_sync(unsigned, float, int) { return 0.f; }   // the other lanes do not exist
using std::min;
using std::max;
namespace macow { alignas(16) float smem[1 << 20]; }

#include "../../ipoke-main-opticalflow_b200/csrc/macow_kernel.cuh"

template <bool kRev, int kQC, bool kWS>
static void run_ws(int k1, const macow::UnitParams& p) {
  for (unsigned b = 0; b < (unsigned)p.B; ++b) {
    blockIdx.x = b;
    if (k1 == 1) macow::macow_unit_kernel<kRev, 1, kQC, kWS>(p);
    else if (k1 == 3) macow::macow_unit_kernel<kRev, 3, kQC, kWS>(p);
    else macow::macow_unit_kernel<kRev, 5, kQC, kWS>(p);
  }
}

template <bool kRev, int kQC>
static void run_k1(int k1, const macow::UnitParams& p, int ws) {
  if (ws) run_ws<kRev, kQC, true>(k1, p); else run_ws<kRev, kQC, false>(k1, p);
}

extern "C" {

long long emul_unit_floats(int C, int Hd, int k0, int k1) { return macow::unit_floats(C, Hd, k0, k1); }

void emul_pack(const float* const* params, float* packed, int C, int Hd, int k0, int k1) {
  macow::PackParams p;
  for (int i = 0; i < 20; ++i) p.params[i] = params[i];
  p.packed = packed; p.C = C; p.Hd = Hd; p.k0 = k0; p.k1 = k1;
  p.hc = 0; p.wh = nullptr; p.wh_ld = 0; p.wh_col0 = 0;
  blockDim.x = 1; gridDim.x = 1; blockIdx.x = 0; threadIdx.x = 0;
  macow::macow_pack_kernel(p);
}

void emul_run(const float* packed, int n_units, const float* in, float* out, float* logdet, int accumulate, int B,
              int C, int Hd, int H, int W, int k0, int k1, int act, float alpha, int s0, int s1, int tile_lines,
              int reverse, int qc, int ksplit, int ws) {
  macow::UnitParams p;
  p.packed = packed; p.in = in; p.out = out; p.logdet = logdet;
  p.B = B; p.C = C; p.Hd = Hd; p.H = H; p.W = W; p.k0 = k0; p.k1 = k1; p.alpha = alpha;
  p.stage_begin = s0; p.stage_end = s1; p.tile_lines = tile_lines; p.accumulate = accumulate;
  p.unit_floats = macow::unit_floats(C, Hd, k0, k1);
  p.in_stride = p.out_stride = (long long)C * H * W;
  p.hbias = nullptr; p.hb_ld = 0;
  blockDim.x = 1; gridDim.x = B; threadIdx.x = 0;
  p.act = act;
  p.ksplit = ksplit;
  p.hid_floats = ((reverse ? 1 : tile_lines) * ksplit * (H > W ? H : W) * Hd + 3) & ~3;
  if (reverse) {
    if (qc == 8) run_k1<true, 8>(k1, p, ws); else run_k1<true, 4>(k1, p, ws);
  } else {
    if (qc == 8) run_k1<false, 8>(k1, p, ws); else run_k1<false, 4>(k1, p, ws);
  }
}
}
