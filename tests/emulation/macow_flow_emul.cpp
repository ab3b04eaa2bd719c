// TEST INFRASTRUCTURE: the MaCow C ABI (include/ipoke_macow.h) on the CPU.  The device code of csrc/macow_kernel.cuh and
// csrc/macow_flow_kernel.cuh is compiled by g++ and run by the fiber emulation of cuda_emul.h with the REAL launch
// geometry (threads per CTA, grids, shared-memory carving) that csrc/macow_flow_host.h computes -- the same host code
// the CUDA library uses, driven through an emulation Launcher instead of <<< >>>.  Pointers are host pointers.
// Built by tests/test_macow_flow_emulation.py into tests/emulation/_build/; never linked into libipoke_inn.so and never
// loaded by the package (the tests hand it to the Python host code by monkeypatching).
#include "cuda_emul.h"

namespace macow { alignas(16) float smem[1 << 20]; }

#include "../../ipoke-main-opticalflow_b200/csrc/macow_flow_host.h"

namespace {

struct EmulLauncher {
  template <int kSrc, int kEpi, int BM, int BN, int TM, int TN>
  int gemm(bool async, int gx, int gy, int gz, const macow::GemmParams& p) {
    for (int z = 0; z < gz; ++z) {
      emul::grid_z = z;
      bool done = false;
      if constexpr (kSrc != macow::SRC_NCHW_ACT) {
        if (async) {
          emul::launch(gx, gy, macow::kGemmThreads, [&] { macow::nice_gemm_kernel<kSrc, kEpi, BM, BN, TM, TN, true>(p); });
          done = true;
        }
      }
      if (!done) emul::launch(gx, gy, macow::kGemmThreads, [&] { macow::nice_gemm_kernel<kSrc, kEpi, BM, BN, TM, TN, false>(p); });
    }
    emul::grid_z = 0;
    return 0;
  }
  int affine(int blocks, int threads, const macow::AffineParams& p) {
    emul::launch(blocks, 1, threads, [&] { macow::nice_affine_kernel(p); });
    return 0;
  }
  int chanmap(int B, int threads, size_t smem, const macow::ChanMapParams& p) {
    if (smem > sizeof(macow::smem)) return IPOKE_INN_E_UNSUPPORTED;
    emul::launch(B, 1, threads, [&] { macow::macow_chanmap_kernel(p); });
    return 0;
  }
  int pack_unit(const macow::PackParams& p) {
    emul::launch(2, 1, 64, [&] { macow::macow_pack_kernel(p); });
    return 0;
  }
  int pack_nice(const macow::PackNiceParams& p, int row_blocks, int row_threads) {
    emul::launch(3, 1, 64, [&] { macow::macow_pack_nice_kernel(p); });
    emul::launch(row_blocks, 1, row_threads, [&] { macow::macow_pack_nice_rows_kernel(p); });
    return 0;
  }
  template <bool kRev, int kK1, int kQC>
  void units_w(const macow::UnitParams& p, const macow::Geometry& g) {
    if (g.wsmem) emul::launch(p.B, 1, g.threads, [&] { macow::macow_unit_kernel<kRev, kK1, kQC, true>(p); });
    else emul::launch(p.B, 1, g.threads, [&] { macow::macow_unit_kernel<kRev, kK1, kQC, false>(p); });
  }
  template <bool kRev, int kK1>
  void units_q(const macow::UnitParams& p, const macow::Geometry& g) {
    if (g.qc == 8) units_w<kRev, kK1, 8>(p, g); else units_w<kRev, kK1, 4>(p, g);
  }
  template <bool kRev>
  void units_k(int k1, const macow::UnitParams& p, const macow::Geometry& g) {
    if (k1 == 1) units_q<kRev, 1>(p, g); else if (k1 == 3) units_q<kRev, 3>(p, g); else units_q<kRev, 5>(p, g);
  }
  int units(bool reverse, int k1, const macow::UnitParams& p, const macow::Geometry& g) {
    if ((size_t)g.smem_bytes > sizeof(macow::smem)) return IPOKE_INN_E_UNSUPPORTED;
    if (reverse) units_k<true>(k1, p, g); else units_k<false>(k1, p, g);
    return 0;
  }
};

}  // namespace

extern "C" {

void emul_set_descending(int on) { emul::descending = on != 0; }

size_t ipoke_macow_unit_floats(const ipoke_macow_desc* d) {
  if (macow::check_desc(d)) return 0;
  return (size_t)macow::unit_floats(d->C, d->hidden, d->k0, d->k1);
}

int ipoke_macow_pack_unit(const ipoke_macow_desc* d, const void* const* params, float* packed, void*) {
  EmulLauncher L;
  return macow::pack_unit_run(L, d, params, packed, nullptr, 0, 0);
}

size_t ipoke_macow_cond_floats(const ipoke_macow_desc* d, int n_units) {
  if (macow::check_desc(d) || d->h_channels < 1 || n_units < 1) return 0;
  return (size_t)d->h_channels * macow::cond_ld(d, n_units);
}

int ipoke_macow_pack_unit_cond(const ipoke_macow_desc* d, const void* const* params, float* packed, float* wh, int n_units,
                               int unit_index, void*) {
  EmulLauncher L;
  return macow::pack_unit_run(L, d, params, packed, wh, n_units, unit_index);
}

int ipoke_macow_cond_bias(const ipoke_macow_desc* d, const float* wh, int n_units, const float* h, long long h_stride,
                          float* hb, int B, int H, int W, void*) {
  EmulLauncher L;
  return macow::cond_bias_run(L, d, wh, n_units, h, h_stride, hb, B, H, W);
}

int ipoke_macow_units_forward_cond(const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb,
                                   const float* x, long long x_stride, float* out, long long out_stride, float* logdet,
                                   int accumulate, int B, int H, int W, int s0, int s1, void*) {
  EmulLauncher L;
  return macow::units_run(L, d, packed, n_units, hb, x, x_stride, out, out_stride, logdet, accumulate, B, H, W, s0, s1, false);
}

int ipoke_macow_units_reverse_cond(const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb,
                                   const float* z, long long z_stride, float* out, long long out_stride, int B, int H,
                                   int W, int s0, int s1, void*) {
  EmulLauncher L;
  return macow::units_run(L, d, packed, n_units, hb, z, z_stride, out, out_stride, nullptr, 0, B, H, W, s0, s1, true);
}

int ipoke_macow_units_forward(const ipoke_macow_desc* d, const float* packed, int n_units, const float* x,
                              long long x_stride, float* out, long long out_stride, float* logdet, int accumulate,
                              int B, int H, int W, int s0, int s1, void*) {
  EmulLauncher L;
  return macow::units_run(L, d, packed, n_units, nullptr, x, x_stride, out, out_stride, logdet, accumulate, B, H, W, s0, s1, false);
}

int ipoke_macow_units_reverse(const ipoke_macow_desc* d, const float* packed, int n_units, const float* z,
                              long long z_stride, float* out, long long out_stride, int B, int H, int W, int s0, int s1,
                              void*) {
  EmulLauncher L;
  return macow::units_run(L, d, packed, n_units, nullptr, z, z_stride, out, out_stride, nullptr, 0, B, H, W, s0, s1, true);
}

int ipoke_macow_forward(const ipoke_macow_desc* d, const float* packed, int n_units, const float* x, float* out,
                        float* logdet, int accumulate, int B, int H, int W, int s0, int s1, void* stream) {
  const long long n = d ? (long long)d->C * H * W : 0;
  return ipoke_macow_units_forward(d, packed, n_units, x, n, out, n, logdet, accumulate, B, H, W, s0, s1, stream);
}

int ipoke_macow_reverse(const ipoke_macow_desc* d, const float* packed, int n_units, const float* z, float* out, int B,
                        int H, int W, int s0, int s1, void* stream) {
  const long long n = d ? (long long)d->C * H * W : 0;
  return ipoke_macow_units_reverse(d, packed, n_units, z, n, out, n, B, H, W, s0, s1, stream);
}

size_t ipoke_macow_nice_floats(const ipoke_macow_nice_desc* d) {
  if (macow::check_nice(d)) return 0;
  return (size_t)macow::nice_floats(d->c_in, d->c_t, d->hidden);
}

int ipoke_macow_pack_nice(const ipoke_macow_nice_desc* d, const void* const* params, float* packed, void*) {
  EmulLauncher L;
  return macow::pack_nice_run(L, d, params, packed);
}

size_t ipoke_macow_nice_workspace_floats(const ipoke_macow_nice_desc* d, int B, int H, int W) {
  if (macow::check_nice(d) || B < 1 || H < 1 || W < 1) return 0;
  return (size_t)macow::nice_chunk_samples(d, B, H, W, ~(size_t)0) * (size_t)macow::nice_sample_floats(d, H, W);
}

int ipoke_macow_nice(const ipoke_macow_nice_desc* d, const float* packed, float* x, long long x_stride, float* logdet,
                     int reverse, int B, int H, int W, float* ws, size_t ws_floats, void*) {
  EmulLauncher L;
  return macow::nice_run(L, d, packed, x, x_stride, logdet, reverse, B, H, W, ws, ws_floats);
}

int ipoke_macow_chanmap(float* x, long long x_stride, int B, int H, int W, const long long* perm, int n_perm,
                        const float* log_scale, const float* bias, int norm_c0, int norm_n, int norm_inverse,
                        int norm_first, float* logdet, void*) {
  EmulLauncher L;
  return macow::chanmap_run(L, x, x_stride, B, H, W, perm, n_perm, log_scale, bias, norm_c0, norm_n, norm_inverse,
                            norm_first, logdet);
}

int ipoke_macow_launch_info(const ipoke_macow_desc* d, int H, int W, int reverse, int* smem_bytes, int* threads,
                            int* tile_lines) {
  macow::Geometry g;
  if (int e = macow::geometry(d, 1, H, W, reverse != 0, &g)) return e;
  if (smem_bytes) *smem_bytes = g.smem_bytes;
  if (threads) *threads = g.threads;
  if (tile_lines) *tile_lines = g.tile_lines;
  return 0;
}

}  // extern "C"
