// Host side of the MaCow flow kernels (macow_flow_kernel.cuh): argument checks, pass sizing, buffer carving and the
// launch sequence, written over a Launcher so that the very same code drives the CUDA kernels (macow_capi.cu) and the
// multi-thread CPU emulation of the tests (tests/emulation/macow_flow_emul.cpp).
#pragma once
#include <algorithm>
#include <cstdlib>

#include "ipoke_macow.h"
#include "macow_flow_kernel.cuh"

namespace macow {

constexpr int kHostMaxSmem = 227 * 1024;

// ---- MaCow units (macow_kernel.cuh)
struct Geometry {
  int threads, tile_lines, smem_bytes;
  int qc;         // pixels per phase-1 work item: 8 when a line has more than 4 pixels
  int ksplit;     // inverse: thread groups sharing one line's input-channel loop
  int hid_floats; // hidden-activation area
  int wsmem;      // 1: one flow's weights are staged in shared memory
  int carveout;   // preferred shared-memory carve-out in percent, -1 = driver default
};

inline int check_desc(const ipoke_macow_desc* d) {
  if (!d) return IPOKE_INN_E_INVALID;
  if (d->C < 1 || d->hidden < 4 || (d->hidden & 3) || d->k0 < 1) return IPOKE_INN_E_UNSUPPORTED;
  if (d->k1 != 1 && d->k1 != 3 && d->k1 != 5) return IPOKE_INN_E_UNSUPPORTED;     // ShiftedConv2d asserts odd (macow_utils.py:454)
  if (d->activation < 0 || d->activation > 2 || d->h_channels < 0) return IPOKE_INN_E_INVALID;
  return IPOKE_INN_OK;
}

inline int cond_ld(const ipoke_macow_desc* d, int n_units) { return round4(n_units * 8 * d->C); }

inline int geometry(const ipoke_macow_desc* d, int B, int H, int W, bool reverse, Geometry* g) {
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
    while (tile > 1 && (img + tile * line) * 4 > kHostMaxSmem) --tile;
  } else {
    // one line offers hidden * ceil(Q / qc) work items: share the input-channel loop until every thread has one
    const int items = (d->hidden / 2) * ((std::min(H, W) + g->qc - 1) / g->qc);
    ksplit = std::max(1, std::min(std::min(4, d->C), threads / std::max(items, 1)));   // 8 groups measured slower than 4
    if (const char* e = std::getenv("IPOKE_MACOW_KSPLIT")) ksplit = std::max(1, std::min(std::min(8, d->C), std::atoi(e)));
    while (ksplit > 1 && (img + ksplit * line) * 4 > kHostMaxSmem) --ksplit;
  }
  const long long hid = ((long long)tile * ksplit * line + 3) & ~3LL;
  long long smem = (img + hid) * 4;
  if (smem > kHostMaxSmem) return IPOKE_INN_E_UNSUPPORTED;                   // two images + one hidden line must fit one SM
  const long long wfl = (macow::flow_floats(d->C, d->hidden, d->k0, d->k1) + 3) & ~3LL;
  g->wsmem = smem + wfl * 4 <= kHostMaxSmem;                                 // the flow's weights fit beside them
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

template <class Launcher>
int units_run(Launcher& L, const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb, const float* in,
              long long in_stride, float* out, long long out_stride, float* logdet, int accumulate, int B, int H, int W,
              int s0, int s1, bool reverse) {
  if (int e = check_desc(d)) return e;
  if (!packed || !in || !out || n_units < 1 || B < 0) return IPOKE_INN_E_INVALID;
  if ((d->h_channels > 0) != (hb != nullptr)) return IPOKE_INN_E_INVALID;      // conditional units need their hb, others none
  if (in_stride < (long long)d->C * H * W || out_stride < (long long)d->C * H * W) return IPOKE_INN_E_INVALID;
  if (s0 < 0 || s1 > n_units * IPOKE_MACOW_STAGES_PER_UNIT || s0 > s1) return IPOKE_INN_E_INVALID;
  if (B == 0) return IPOKE_INN_OK;
  Geometry g;
  if (int e = geometry(d, B, H, W, reverse, &g)) return e;
  UnitParams p;
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
  p.unit_floats = unit_floats(d->C, d->hidden, d->k0, d->k1);
  p.in_stride = in_stride;
  p.out_stride = out_stride;
  p.hbias = hb;
  p.hb_ld = hb ? cond_ld(d, n_units) : 0;
  return L.units(reverse, d->k1, p, g);
}

template <class Launcher>
int pack_unit_run(Launcher& L, const ipoke_macow_desc* d, const void* const* params, float* packed, float* wh, int n_units,
                  int unit_index) {
  if (int e = check_desc(d)) return e;
  if (!params || !packed) return IPOKE_INN_E_INVALID;
  if ((d->h_channels > 0) != (wh != nullptr)) return IPOKE_INN_E_INVALID;
  if (wh && (n_units < 1 || unit_index < 0 || unit_index >= n_units)) return IPOKE_INN_E_INVALID;
  PackParams p;
  for (int i = 0; i < 20; ++i) {
    if (!params[i]) return IPOKE_INN_E_INVALID;
    p.params[i] = static_cast<const float*>(params[i]);
  }
  p.packed = packed;
  p.C = d->C; p.Hd = d->hidden; p.k0 = d->k0; p.k1 = d->k1;
  p.hc = d->h_channels; p.wh = wh;
  p.wh_ld = wh ? cond_ld(d, n_units) : 0;
  p.wh_col0 = unit_index * 8 * d->C;
  return L.pack_unit(p);
}

// hb = act(h) x Wh for a chain of conditional units (one GEMM over rows = (sample, pixel); the activation is applied
// while the NCHW conditioning is gathered into the A tile).
template <class Launcher>
int cond_bias_run(Launcher& L, const ipoke_macow_desc* d, const float* wh, int n_units, const float* h, long long h_stride,
                  float* hb, int B, int H, int W) {
  if (int e = check_desc(d)) return e;
  if (d->h_channels < 1 || !wh || !h || !hb || n_units < 1 || B < 0 || H < 1 || W < 1) return IPOKE_INN_E_INVALID;
  if (h_stride < (long long)d->h_channels * H * W) return IPOKE_INN_E_INVALID;
  if (B == 0) return IPOKE_INN_OK;
  const int ldn = cond_ld(d, n_units);
  GemmParams p{};
  p.A = h; p.sample_stride = h_stride; p.lda = 0;
  p.Wt = wh; p.ldw = ldn;
  p.Y = hb; p.ldy = ldn;
  p.M = B * H * W; p.N = ldn; p.K = d->h_channels;
  p.H = H; p.W = W;
  p.g = NiceGeom{0, 0, 1, 0, 0, 1};
  p.act = d->activation;
  p.k_split = 1 << 30; p.y_split_stride = 0;
  return L.template gemm<SRC_NCHW_ACT, EPI_STORE, 128, 64, 8, 4>(false, (p.M + 127) / 128, (ldn + 63) / 64, 1, p);
}

// ---- coupling nets and per-channel glue (macow_flow_kernel.cuh)
inline int check_nice(const ipoke_macow_nice_desc* d) {
  if (!d) return IPOKE_INN_E_INVALID;
  if (d->c_in < 1 || d->c_t < 1 || d->hidden < 1 || d->z_stride < 1 || d->p_stride < 1 || d->z_off < 0 || d->p_off < 0)
    return IPOKE_INN_E_INVALID;
  if (d->activation < 0 || d->activation > 2) return IPOKE_INN_E_INVALID;
  return IPOKE_INN_OK;
}

// Workspace floats of one sample of a coupling pass: two pixel-major hidden-activation buffers [HW][Hd] and the nine
// tap outputs of the last convolution [HW][9 * round4(2 c_t)].
inline long long nice_sample_floats(const ipoke_macow_nice_desc* d, int H, int W) {
  return (long long)H * W * (2LL * d->hidden + 9LL * round4(2 * d->c_t));
}

// Samples per pass of the coupling net: the buffers of a pass should stay in L2 (48 MB budget, IPOKE_MACOW_NICE_MB
// overrides), bounded by the workspace the caller brought.
inline int nice_chunk_samples(const ipoke_macow_nice_desc* d, int B, int H, int W, size_t ws_floats) {
  long long mb = 48;
  if (const char* e = std::getenv("IPOKE_MACOW_NICE_MB")) mb = std::max(1, std::atoi(e));
  const long long per_sample = nice_sample_floats(d, H, W);
  long long chunk = std::max(1LL, mb * 1024 * 1024 / 4 / per_sample);
  chunk = std::min<long long>(chunk, (long long)(ws_floats / (size_t)per_sample));
  return (int)std::max(1LL, std::min<long long>(chunk, B));
}

// K slices of the tap GEMM: enough CTAs for two per SM when the batch is small, at most 8, and only as many as fit the
// dead first activation buffer (slices * ldt <= Hd).  IPOKE_MACOW_TAP_SPLITS overrides the CTA-count rule.
inline int nice_tap_splits(int ctas, int Hd, int ldt) {
  int s = ctas >= 148 ? 1 : std::min(8, (296 + ctas - 1) / ctas);
  if (const char* e = std::getenv("IPOKE_MACOW_TAP_SPLITS")) s = std::max(1, std::atoi(e));
  s = std::min(s, Hd / ldt);
  s = std::min(s, (Hd + kGemmBK - 1) / kGemmBK);
  return std::max(s, 1);
}

template <class Launcher>
int pack_nice_run(Launcher& L, const ipoke_macow_nice_desc* d, const void* const* params, float* packed) {
  if (int e = check_nice(d)) return e;
  if (!params || !packed) return IPOKE_INN_E_INVALID;
  PackNiceParams p;
  for (int i = 0; i < 5; ++i) {
    if (!params[i]) return IPOKE_INN_E_INVALID;
    p.params[i] = static_cast<const float*>(params[i]);
  }
  p.packed = packed;
  p.c_in = d->c_in; p.c_t = d->c_t; p.Hd = d->hidden;
  return L.pack_nice(p, (2 * d->c_t + 63) / 64, 64);
}

template <class Launcher>
int nice_run(Launcher& L, const ipoke_macow_nice_desc* d, const float* packed, float* x, long long x_stride,
             float* logdet, int reverse, int B, int H, int W, float* ws, size_t ws_floats) {
  if (int e = check_nice(d)) return e;
  if (!packed || !x || !ws || B < 0 || H < 1 || W < 1) return IPOKE_INN_E_INVALID;
  const int HW = H * W, Hd = d->hidden;
  const int c_hi = std::max(d->z_off + d->z_stride * (d->c_in - 1), d->p_off + d->p_stride * (d->c_t - 1)) + 1;
  if (x_stride < (long long)c_hi * HW) return IPOKE_INN_E_INVALID;
  if (ws_floats < (size_t)nice_sample_floats(d, H, W)) return IPOKE_INN_E_INVALID;      // at least one sample per pass
  if (B == 0) return IPOKE_INN_OK;
  const int chunk = nice_chunk_samples(d, B, H, W, ws_floats);
  bool async = true;                                         // double-buffered cp.async K slices (IPOKE_MACOW_GEMM_ASYNC=0:
  if (const char* e = std::getenv("IPOKE_MACOW_GEMM_ASYNC")) async = std::atoi(e) != 0;   // single buffer through registers)
  const int ld = round4(Hd), ld3 = round4(2 * d->c_t), ldt = 9 * ld3;
  const float* w1 = packed;
  const float* w2 = w1 + 9LL * d->c_in * ld;
  const float* w3 = w2 + (long long)Hd * ld;
  const float* b3 = w3 + 9LL * Hd * ld3;
  float* a1 = ws;
  float* a2 = a1 + (size_t)chunk * HW * Hd;
  float* taps = a2 + (size_t)chunk * HW * Hd;
  for (int b0 = 0; b0 < B; b0 += chunk) {
    const int nb = std::min(chunk, B - b0), M = nb * HW;
    float* img = x + (long long)b0 * x_stride;
    GemmParams p{};
    p.M = M; p.H = H; p.W = W;
    p.sample_stride = x_stride;
    p.g = NiceGeom{d->c_in, d->z_off, d->z_stride, d->c_t, d->p_off, d->p_stride};
    p.act = d->activation;
    p.k_split = 1 << 30; p.y_split_stride = 0;
    const int gx = (M + 127) / 128;
    // conv1: 3x3 over the net's input channels of the image -> activation -> a1 [M, Hd]
    p.A = img; p.lda = 0; p.Wt = w1; p.ldw = ld; p.K = 9 * d->c_in; p.N = Hd; p.Y = a1; p.ldy = Hd;
    if (int e = L.template gemm<SRC_IMAGE, EPI_ACT, 128, 128, 8, 8>(async, gx, (Hd + 127) / 128, 1, p)) return e;
    // conv2: 1x1 -> activation -> a2
    p.A = a1; p.lda = Hd; p.Wt = w2; p.K = Hd; p.Y = a2;
    if (int e = L.template gemm<SRC_ROWS, EPI_ACT, 128, 128, 8, 8>(async, gx, (Hd + 127) / 128, 1, p)) return e;
    // conv3 (weight-normed 3x3) as nine 1x1 tap products side by side: taps [M, 9 ld3].  A small batch gives this
    // narrow product only a handful of CTAs: K is then split across grid.z and the partial products (parked in a1, which
    // is dead by now) are added in slice order by the tap kernel.
    const int splits = nice_tap_splits(gx * ((ldt + 63) / 64), Hd, ldt);
    p.A = a2; p.Wt = w3; p.ldw = ldt; p.N = ldt; p.ldy = ldt;
    p.Y = splits > 1 ? a1 : taps;
    if (splits > 1) {
      p.k_split = ((Hd + splits - 1) / splits + kGemmBK - 1) / kGemmBK * kGemmBK;
      p.y_split_stride = (long long)M * ldt;
    }
    const int gz = splits > 1 ? (Hd + p.k_split - 1) / p.k_split : 1;
    if (int e = L.template gemm<SRC_ROWS, EPI_STORE, 128, 64, 8, 4>(async, gx, (ldt + 63) / 64, gz, p)) return e;
    // shift-add of the taps + bias -> (shift, scale) -> affine map of the transformed channels in place, log-det
    AffineParams q;
    q.T = p.Y; q.splits = gz; q.split_stride = p.y_split_stride; q.bias = b3; q.x = img; q.logdet = (!reverse && logdet) ? logdet + b0 : nullptr;
    q.H = H; q.W = W; q.ldt = ldt; q.ld3 = ld3; q.sample_stride = x_stride; q.g = p.g; q.alpha = d->alpha;
    q.reverse = reverse != 0;
    if (int e = L.affine(nb, 256, q)) return e;
  }
  return IPOKE_INN_OK;
}

template <class Launcher>
int chanmap_run(Launcher& L, float* x, long long x_stride, int B, int H, int W, const long long* perm, int n_perm,
                const float* log_scale, const float* bias, int norm_c0, int norm_n, int norm_inverse, int norm_first,
                float* logdet) {
  if (!x || B < 0 || H < 1 || W < 1 || n_perm < 0 || norm_n < 0 || norm_c0 < 0) return IPOKE_INN_E_INVALID;
  if ((norm_n > 0) != (log_scale != nullptr) || (norm_n > 0) != (bias != nullptr)) return IPOKE_INN_E_INVALID;
  if ((n_perm > 0) != (perm != nullptr)) return IPOKE_INN_E_INVALID;
  ChanMapParams p;
  p.x = x; p.sample_stride = x_stride; p.HW = H * W;
  p.n_stage = std::max(n_perm, norm_n > 0 ? norm_c0 + norm_n : 0);
  if (x_stride < (long long)p.n_stage * p.HW) return IPOKE_INN_E_INVALID;
  if (B == 0 || p.n_stage == 0) return IPOKE_INN_OK;
  p.perm = perm; p.n_perm = n_perm;
  p.log_scale = log_scale; p.bias = bias;
  p.norm_c0 = norm_c0; p.norm_n = norm_n; p.norm_inverse = norm_inverse != 0; p.norm_first = norm_first != 0;
  p.logdet = logdet;
  const long long smem = (long long)p.n_stage * p.HW * 4;
  if (smem > kHostMaxSmem) return IPOKE_INN_E_UNSUPPORTED;
  return L.chanmap(B, 256, (size_t)smem, p);
}

}  // namespace macow
