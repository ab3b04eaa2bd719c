// C ABI of the MaCow unit (include/ipoke_macow.h): argument checks, launch geometry, kernel dispatch.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>

#include "ipoke_macow.h"
#include "macow_kernel.cuh"

namespace {

constexpr int kMaxSmem = 227 * 1024;

struct Geometry {
  int threads, tile_lines, smem_bytes;
  int qc;         // pixels per phase-1 work item: 8 when a line has more than 4 pixels
  int ksplit;     // inverse: thread groups sharing one line's input-channel loop
  int hid_floats; // hidden-activation area
  int wsmem;      // 1: one flow's weights are staged in shared memory
  int carveout;   // preferred shared-memory carve-out in percent, -1 = driver default
};

int check_desc(const ipoke_macow_desc* d) {
  if (!d) return IPOKE_INN_E_INVALID;
  if (d->C < 1 || d->hidden < 4 || (d->hidden & 3) || d->k0 < 1) return IPOKE_INN_E_UNSUPPORTED;
  if (d->k1 != 1 && d->k1 != 3 && d->k1 != 5) return IPOKE_INN_E_UNSUPPORTED;     // ShiftedConv2d asserts odd (macow_utils.py:454)
  if (d->activation < 0 || d->activation > 2) return IPOKE_INN_E_INVALID;
  return IPOKE_INN_OK;
}

int geometry(const ipoke_macow_desc* d, int B, int H, int W, bool reverse, Geometry* g) {
  if (H < 1 || W < 1) return IPOKE_INN_E_INVALID;
  // measured on B200 (profiles/r02_macow_sweep.txt): one CTA per SM with the weights in shared memory; 512 threads up to
  // 8 x 8 latents, 256 from 16 x 16 on (more work items per thread, fewer barrier stalls)
  (void)B;
  int threads = std::max(H, W) > 8 ? 256 : 512;
  if (const char* e = std::getenv("IPOKE_MACOW_THREADS")) threads = std::atoi(e);
  threads = std::min(512, std::max(32, threads)) & ~31;
  int hid_kb = 48;                                                     // forward: hidden tile budget
  if (const char* e = std::getenv("IPOKE_MACOW_HID_KB")) hid_kb = std::max(1, std::atoi(e));
  g->qc = std::max(H, W) > 4 ? 8 : 4;
  if (const char* e = std::getenv("IPOKE_MACOW_QC")) g->qc = std::atoi(e) == 8 ? 8 : 4;
  const long long img = 2 * (((long long)d->C * H * W + 3) & ~3LL);      // row-major and column-major copies
  const long long line = (long long)std::max(H, W) * d->hidden;         // hidden floats of one line
  int tile = 1, ksplit = 1;
  if (!reverse) {
    tile = (int)std::max(1LL, std::min<long long>(std::max(H, W), (long long)hid_kb * 256 / line));
    while (tile > 1 && (img + tile * line) * 4 > kMaxSmem) --tile;
  } else {
    // one line offers hidden * ceil(Q / qc) work items: share the input-channel loop until every thread has one
    const int items = (d->hidden / 2) * ((std::min(H, W) + g->qc - 1) / g->qc);
    ksplit = std::max(1, std::min(std::min(4, d->C), threads / std::max(items, 1)));   // 8 groups measured slower than 4
    if (const char* e = std::getenv("IPOKE_MACOW_KSPLIT")) ksplit = std::max(1, std::min(std::min(8, d->C), std::atoi(e)));
    while (ksplit > 1 && (img + ksplit * line) * 4 > kMaxSmem) --ksplit;
  }
  const long long hid = ((long long)tile * ksplit * line + 3) & ~3LL;
  long long smem = (img + hid) * 4;
  if (smem > kMaxSmem) return IPOKE_INN_E_UNSUPPORTED;                   // two images + one hidden line must fit one SM
  const long long wfl = (macow::flow_floats(d->C, d->hidden, d->k0, d->k1) + 3) & ~3LL;
  g->wsmem = smem + wfl * 4 <= kMaxSmem;                                 // the flow's weights fit beside them
  if (const char* e = std::getenv("IPOKE_MACOW_WSMEM")) g->wsmem = g->wsmem && std::atoi(e) != 0;
  if (g->wsmem) smem += wfl * 4;
  g->hid_floats = (int)hid;
  g->ksplit = ksplit;
  g->carveout = -1;
  if (const char* e = std::getenv("IPOKE_MACOW_CARVEOUT")) g->carveout = std::min(100, std::max(-1, std::atoi(e)));
  g->threads = threads;
  g->tile_lines = tile;
  g->smem_bytes = (int)smem;
  return IPOKE_INN_OK;
}

template <bool kReverse, int kK1, int kQC, bool kWS>
int launch_w(const macow::UnitParams& p, const Geometry& g, cudaStream_t s) {
  auto kern = macow::macow_unit_kernel<kReverse, kK1, kQC, kWS>;
  if (g.smem_bytes > 32 * 1024) {   // the 48 KB default covers dynamic + static shared memory together
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, g.smem_bytes);
    if (e != cudaSuccess) return (int)e;
  }
  if (g.carveout >= 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, g.carveout);
    if (e != cudaSuccess) return (int)e;
  }
  kern<<<p.B, g.threads, g.smem_bytes, s>>>(p);
  return (int)cudaGetLastError();
}

template <bool kReverse, int kK1, int kQC>
int launch_t(const macow::UnitParams& p, const Geometry& g, cudaStream_t s) {
  return g.wsmem ? launch_w<kReverse, kK1, kQC, true>(p, g, s) : launch_w<kReverse, kK1, kQC, false>(p, g, s);
}

template <bool kReverse, int kK1>
int launch_q(const macow::UnitParams& p, const Geometry& g, cudaStream_t s) {
  return g.qc == 8 ? launch_t<kReverse, kK1, 8>(p, g, s) : launch_t<kReverse, kK1, 4>(p, g, s);
}

template <bool kReverse>
int launch(const ipoke_macow_desc* d, const macow::UnitParams& p, const Geometry& g, cudaStream_t s) {
  switch (d->k1) {
    case 1: return launch_q<kReverse, 1>(p, g, s);
    case 3: return launch_q<kReverse, 3>(p, g, s);
    default: return launch_q<kReverse, 5>(p, g, s);
  }
}

int run(const ipoke_macow_desc* d, const float* packed, int n_units, const float* in, float* out, float* logdet,
        int accumulate, int B, int H, int W, int s0, int s1, bool reverse, void* stream) {
  if (int e = check_desc(d)) return e;
  if (!packed || !in || !out || n_units < 1 || B < 0) return IPOKE_INN_E_INVALID;
  if (s0 < 0 || s1 > n_units * IPOKE_MACOW_STAGES_PER_UNIT || s0 > s1) return IPOKE_INN_E_INVALID;
  if (B == 0) return IPOKE_INN_OK;
  Geometry g;
  if (int e = geometry(d, B, H, W, reverse, &g)) return e;
  macow::UnitParams p;
  p.packed = packed;
  p.in = in;
  p.out = out;
  p.logdet = logdet;
  p.B = B; p.C = d->C; p.Hd = d->hidden; p.H = H; p.W = W; p.k0 = d->k0; p.k1 = d->k1;
  p.alpha = d->alpha;
  p.stage_begin = s0; p.stage_end = s1;
  p.tile_lines = g.tile_lines;
  p.accumulate = accumulate;
  p.act = d->activation;
  p.ksplit = g.ksplit;
  p.hid_floats = g.hid_floats;
  p.unit_floats = macow::unit_floats(d->C, d->hidden, d->k0, d->k1);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return reverse ? launch<true>(d, p, g, s) : launch<false>(d, p, g, s);
}

}  // namespace

extern "C" {

size_t ipoke_macow_unit_floats(const ipoke_macow_desc* d) {
  if (check_desc(d)) return 0;
  return (size_t)macow::unit_floats(d->C, d->hidden, d->k0, d->k1);
}

int ipoke_macow_pack_unit(const ipoke_macow_desc* d, const void* const* params, float* packed, void* stream) {
  if (int e = check_desc(d)) return e;
  if (!params || !packed) return IPOKE_INN_E_INVALID;
  macow::PackParams p;
  for (int i = 0; i < 20; ++i) {
    if (!params[i]) return IPOKE_INN_E_INVALID;
    p.params[i] = static_cast<const float*>(params[i]);
  }
  p.packed = packed;
  p.C = d->C; p.Hd = d->hidden; p.k0 = d->k0; p.k1 = d->k1;
  macow::macow_pack_kernel<<<32, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return (int)cudaGetLastError();
}

int ipoke_macow_forward(const ipoke_macow_desc* d, const float* packed, int n_units, const float* x, float* out,
                        float* logdet, int accumulate, int B, int H, int W, int stage_begin, int stage_end,
                        void* stream) {
  return run(d, packed, n_units, x, out, logdet, accumulate, B, H, W, stage_begin, stage_end, false, stream);
}

int ipoke_macow_reverse(const ipoke_macow_desc* d, const float* packed, int n_units, const float* z, float* out,
                        int B, int H, int W, int stage_begin, int stage_end, void* stream) {
  return run(d, packed, n_units, z, out, nullptr, 0, B, H, W, stage_begin, stage_end, true, stream);
}

int ipoke_macow_launch_info(const ipoke_macow_desc* d, int H, int W, int reverse, int* smem_bytes, int* threads,
                            int* tile_lines) {
  if (int e = check_desc(d)) return e;
  Geometry g;
  if (int e = geometry(d, 1, H, W, reverse != 0, &g)) return e;
  if (smem_bytes) *smem_bytes = g.smem_bytes;
  if (threads) *threads = g.threads;
  if (tile_lines) *tile_lines = g.tile_lines;
  return IPOKE_INN_OK;
}

}  // extern "C"
