// C ABI of the MaCow conv flow (include/ipoke_macow.h): the CUDA launcher behind the argument checks, pass sizing and
// launch sequences of macow_flow_host.h (shared with the CPU emulation of the tests), and the extern "C" entry points.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>

#include "ipoke_macow.h"
#include "macow_kernel.cuh"
#include "macow_flow_kernel.cuh"
#include "macow_flow_host.h"

namespace {

using macow::Geometry;

// Launches of the flow kernels on a stream (the Launcher of macow_flow_host.h).
struct CudaLauncher {
  cudaStream_t s;
  template <int kSrc, int kEpi, int BM, int BN, int TM, int TN>
  int gemm(bool async, int gx, int gy, int gz, const macow::GemmParams& p) {
    if constexpr (kSrc != macow::SRC_NCHW_ACT) {
      if (async) {
        macow::nice_gemm_kernel<kSrc, kEpi, BM, BN, TM, TN, true><<<dim3(gx, gy, gz), macow::kGemmThreads, 0, s>>>(p);
        return (int)cudaGetLastError();
      }
    }
    macow::nice_gemm_kernel<kSrc, kEpi, BM, BN, TM, TN, false><<<dim3(gx, gy, gz), macow::kGemmThreads, 0, s>>>(p);
    return (int)cudaGetLastError();
  }
  int affine(int blocks, int threads, const macow::AffineParams& p) {
    macow::nice_affine_kernel<<<blocks, threads, 0, s>>>(p);
    return (int)cudaGetLastError();
  }
  int chanmap(int B, int threads, size_t smem, const macow::ChanMapParams& p) {
    if (smem > 32 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(macow::macow_chanmap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return (int)e;
    }
    macow::macow_chanmap_kernel<<<B, threads, smem, s>>>(p);
    return (int)cudaGetLastError();
  }
  int units(bool reverse, int k1, const macow::UnitParams& p, const Geometry& g);
  int pack_unit(const macow::PackParams& p) {
    macow::macow_pack_kernel<<<32, 256, 0, s>>>(p);
    return (int)cudaGetLastError();
  }
  int pack_nice(const macow::PackNiceParams& p, int row_blocks, int row_threads) {
    macow::macow_pack_nice_kernel<<<148, 256, 0, s>>>(p);
    macow::macow_pack_nice_rows_kernel<<<row_blocks, row_threads, 0, s>>>(p);
    return (int)cudaGetLastError();
  }
};

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
int launch(int k1, const macow::UnitParams& p, const Geometry& g, cudaStream_t s) {
  switch (k1) {
    case 1: return launch_q<kReverse, 1>(p, g, s);
    case 3: return launch_q<kReverse, 3>(p, g, s);
    default: return launch_q<kReverse, 5>(p, g, s);
  }
}

int CudaLauncher::units(bool reverse, int k1, const macow::UnitParams& p, const Geometry& g) {
  return reverse ? launch<true>(k1, p, g, s) : launch<false>(k1, p, g, s);
}

int run(const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb, const float* in, long long in_stride,
        float* out, long long out_stride, float* logdet, int accumulate, int B, int H, int W, int s0, int s1, bool reverse,
        void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::units_run(L, d, packed, n_units, hb, in, in_stride, out, out_stride, logdet, accumulate, B, H, W, s0, s1,
                          reverse);
}

}  // namespace

extern "C" {

size_t ipoke_macow_unit_floats(const ipoke_macow_desc* d) {
  if (macow::check_desc(d)) return 0;
  return (size_t)macow::unit_floats(d->C, d->hidden, d->k0, d->k1);
}

int ipoke_macow_pack_unit(const ipoke_macow_desc* d, const void* const* params, float* packed, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::pack_unit_run(L, d, params, packed, nullptr, 0, 0);
}

size_t ipoke_macow_cond_floats(const ipoke_macow_desc* d, int n_units) {
  if (macow::check_desc(d) || d->h_channels < 1 || n_units < 1) return 0;
  return (size_t)d->h_channels * macow::cond_ld(d, n_units);
}

int ipoke_macow_pack_unit_cond(const ipoke_macow_desc* d, const void* const* params, float* packed, float* wh, int n_units,
                               int unit_index, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::pack_unit_run(L, d, params, packed, wh, n_units, unit_index);
}

int ipoke_macow_cond_bias(const ipoke_macow_desc* d, const float* wh, int n_units, const float* h, long long h_stride,
                          float* hb, int B, int H, int W, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::cond_bias_run(L, d, wh, n_units, h, h_stride, hb, B, H, W);
}

int ipoke_macow_units_forward_cond(const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb,
                                   const float* x, long long x_stride, float* out, long long out_stride, float* logdet,
                                   int accumulate, int B, int H, int W, int stage_begin, int stage_end, void* stream) {
  return run(d, packed, n_units, hb, x, x_stride, out, out_stride, logdet, accumulate, B, H, W, stage_begin, stage_end,
             false, stream);
}

int ipoke_macow_units_reverse_cond(const ipoke_macow_desc* d, const float* packed, int n_units, const float* hb,
                                   const float* z, long long z_stride, float* out, long long out_stride, int B, int H,
                                   int W, int stage_begin, int stage_end, void* stream) {
  return run(d, packed, n_units, hb, z, z_stride, out, out_stride, nullptr, 0, B, H, W, stage_begin, stage_end, true,
             stream);
}

int ipoke_macow_forward(const ipoke_macow_desc* d, const float* packed, int n_units, const float* x, float* out,
                        float* logdet, int accumulate, int B, int H, int W, int stage_begin, int stage_end,
                        void* stream) {
  const long long n = d ? (long long)d->C * H * W : 0;
  return run(d, packed, n_units, nullptr, x, n, out, n, logdet, accumulate, B, H, W, stage_begin, stage_end, false, stream);
}

int ipoke_macow_reverse(const ipoke_macow_desc* d, const float* packed, int n_units, const float* z, float* out,
                        int B, int H, int W, int stage_begin, int stage_end, void* stream) {
  const long long n = d ? (long long)d->C * H * W : 0;
  return run(d, packed, n_units, nullptr, z, n, out, n, nullptr, 0, B, H, W, stage_begin, stage_end, true, stream);
}

int ipoke_macow_units_forward(const ipoke_macow_desc* d, const float* packed, int n_units, const float* x,
                              long long x_stride, float* out, long long out_stride, float* logdet, int accumulate,
                              int B, int H, int W, int stage_begin, int stage_end, void* stream) {
  return run(d, packed, n_units, nullptr, x, x_stride, out, out_stride, logdet, accumulate, B, H, W, stage_begin,
             stage_end, false, stream);
}

int ipoke_macow_units_reverse(const ipoke_macow_desc* d, const float* packed, int n_units, const float* z,
                              long long z_stride, float* out, long long out_stride, int B, int H, int W,
                              int stage_begin, int stage_end, void* stream) {
  return run(d, packed, n_units, nullptr, z, z_stride, out, out_stride, nullptr, 0, B, H, W, stage_begin, stage_end, true,
             stream);
}

size_t ipoke_macow_nice_floats(const ipoke_macow_nice_desc* d) {
  if (macow::check_nice(d)) return 0;
  return (size_t)macow::nice_floats(d->c_in, d->c_t, d->hidden);
}

int ipoke_macow_pack_nice(const ipoke_macow_nice_desc* d, const void* const* params, float* packed, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::pack_nice_run(L, d, params, packed);
}

size_t ipoke_macow_nice_workspace_floats(const ipoke_macow_nice_desc* d, int B, int H, int W) {
  if (macow::check_nice(d) || B < 1 || H < 1 || W < 1) return 0;
  return (size_t)macow::nice_chunk_samples(d, B, H, W, ~(size_t)0) * (size_t)macow::nice_sample_floats(d, H, W);
}

int ipoke_macow_nice(const ipoke_macow_nice_desc* d, const float* packed, float* x, long long x_stride, float* logdet,
                     int reverse, int B, int H, int W, float* ws, size_t ws_floats, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::nice_run(L, d, packed, x, x_stride, logdet, reverse, B, H, W, ws, ws_floats);
}

int ipoke_macow_chanmap(float* x, long long x_stride, int B, int H, int W, const long long* perm, int n_perm,
                        const float* log_scale, const float* bias, int norm_c0, int norm_n, int norm_inverse,
                        int norm_first, float* logdet, void* stream) {
  CudaLauncher L{static_cast<cudaStream_t>(stream)};
  return macow::chanmap_run(L, x, x_stride, B, H, W, perm, n_perm, log_scale, bias, norm_c0, norm_n, norm_inverse,
                            norm_first, logdet);
}

int ipoke_macow_launch_info(const ipoke_macow_desc* d, int H, int W, int reverse, int* smem_bytes, int* threads,
                            int* tile_lines) {
  if (int e = macow::check_desc(d)) return e;
  Geometry g;
  if (int e = macow::geometry(d, 1, H, W, reverse != 0, &g)) return e;
  if (smem_bytes) *smem_bytes = g.smem_bytes;
  if (threads) *threads = g.threads;
  if (tile_lines) *tile_lines = g.tile_lines;
  return IPOKE_INN_OK;
}

}  // extern "C"
