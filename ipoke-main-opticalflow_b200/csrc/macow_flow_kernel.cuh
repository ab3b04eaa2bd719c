// The rest of the MaCow flow around the units (macow_kernel.cuh), sm_100a, float32 like the reference:
//   * NICE2d coupling with the convolutional net (macow2.py:291-437, macow_utils.py:253-338): 3x3 conv -> activation ->
//     1x1 conv -> activation -> weight-normed 3x3 conv -> Affine (macow_utils.py:42-66), forward and inverse;
//   * the per-channel glue of MaCowStep / MultiScalePrior / MultiScaleInternal (macow2.py:543-594, 821-918, 999-1117):
//     ActNorm2dFlow (macow2.py:502-521) and the Shuffle permutation (flow_blocks.py:314-326), in place on a channel
//     prefix of the image (the multi-scale split / unsplit is then no operation at all: a level works on the first
//     channels of every sample and leaves the factored-out channels where the final concatenation wants them).
//
// Design.  The coupling net is three GEMMs over rows = (sample, pixel):  [M, 9 c_in] x [9 c_in, Hd] with the 3x3 window
// gathered from the NCHW image while the A tile is loaded (implicit im2col),  [M, Hd] x [Hd, Hd],  and the last 3x3
// convolution as nine 1x1 "tap" GEMMs in one  [M, Hd] x [Hd, 9 * 2 c_t]  product whose taps a small per-sample kernel
// shift-adds (K stays Hd and N grows ninefold: full tiles and enough CTAs, where the im2col formulation had 9 Hd deep
// K loops on a handful of CTAs -- measured: 68 % of the flow's time).  Activations are fused into the epilogues of the
// first two, bias + affine map + log-determinant into the tap kernel.  Hidden activations are pixel-major [M, Hd] so
// every layer reads contiguous K.  One tiled float32 kernel (128-row tiles, 16-deep K slices in shared memory, 8 x 8 or
// 8 x 4 register tiles, 256 threads) serves all three.  The hidden width of the reference's configurations
// (1536 - 2048) makes the middle GEMM tensor-core work; this float32 version is the parity path (DESIGN.md section 4b
// says what the tcgen05 version changes).
#pragma once
#include "macow_kernel.cuh"

namespace macow {

// Channel geometry of one NICE2d: the net reads image channels z_off + z_stride * i (i < c_in) and transforms
// channels p_off + p_stride * i (i < c_t)  (continuous / skip split, order up / down: macow2.py:363-389, 410-414).
struct NiceGeom {
  int c_in, z_off, z_stride, c_t, p_off, p_stride;
};

__host__ __device__ inline int round4(int v) { return (v + 3) & ~3; }

// Packed coupling net: Wt1 [9 c_in][ld] | Wt2 [Hd][ld] | Wt3 [Hd][9 ld3] | bias3 [ld3], ld = round4(Hd), ld3 = round4(2 c_t);
// Wt3 holds the nine taps of the last convolution side by side (column tap * ld3 + slot), slots interleaved like bias3:
// 2 i = shift of transformed channel i, 2 i + 1 = its log-scale.
__host__ __device__ inline long long nice_floats(int c_in, int c_t, int Hd) {
  const long long ld = round4(Hd), ld3 = round4(2 * c_t);
  return 9LL * c_in * ld + (long long)Hd * ld + 9LL * Hd * ld3 + ld3;
}

enum { SRC_IMAGE = 0, SRC_ROWS = 1, SRC_NCHW_ACT = 2 };   // NCHW_ACT: A(m, k) = act(h[b][k][pixel]) (register path only)
enum { EPI_ACT = 0, EPI_STORE = 1 };

struct GemmParams {
  const float* A;      // SRC_ROWS: activations [M][lda];  SRC_IMAGE: the image (first sample of the pass)
  const float* Wt;     // [K][ldw], ldw a multiple of 4, columns >= N zero
  float* Y;            // [M][ldy]
  int M, N, K, lda, ldw, ldy;
  int H, W;
  long long sample_stride;
  NiceGeom g;
  int act;
  int k_split;               // K elements per grid.z slice (a multiple of 16); slice z writes Y + z * y_split_stride
  long long y_split_stride;  // (split-K partial products, added in a fixed order by the consumer)
};

constexpr int kGemmBK = 16;
constexpr int kGemmThreads = 256;

// Address of A(m, k), or null where the operand is zero (outside the matrix / the image: zero padding of the 3x3 window).
template <int kSrc>
__device__ __forceinline__ const float* gemm_addr_a(const GemmParams& p, int m, int k, int py, int px) {
  if (m >= p.M || k >= p.K) return nullptr;
  if (kSrc == SRC_ROWS) return p.A + (long long)m * p.lda + k;
  const int HW = p.H * p.W;
  if (kSrc == SRC_NCHW_ACT) return p.A + (m / HW) * p.sample_stride + (long long)k * HW + py * p.W + px;
  const int c = k / 9, tap = k - c * 9;                      // SRC_IMAGE: k = c * 9 + tap
  const int dy = tap / 3 - 1, dx = tap - (tap / 3) * 3 - 1;
  const int yy = py + dy, xx = px + dx;
  if (yy < 0 || yy >= p.H || xx < 0 || xx >= p.W) return nullptr;
  const int b = m / HW;
  return p.A + b * p.sample_stride + (long long)(p.g.z_off + p.g.z_stride * c) * HW + yy * p.W + xx;
}

// Asynchronous global -> shared copies (LDGSTS): 4 bytes for the transposed A tile, 16 bytes for the B tile; a null
// source zero-fills (src-size 0).  The emulation copies at once.
#ifdef MACOW_HOST_EMULATION
inline void cp_async4(float* dst, const float* src, const float*) { *dst = src ? *src : 0.f; }
inline void cp_async16(float* dst, const float* src, const float*) {
  for (int i = 0; i < 4; ++i) dst[i] = src ? src[i] : 0.f;
}
inline void cp_async_commit() {}
template <int N> inline void cp_async_wait() {}
#else
__device__ __forceinline__ void cp_async4(float* dst, const float* src, const float* valid_base) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  const int bytes = src ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src ? src : valid_base), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async16(float* dst, const float* src, const float* valid_base) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  const int bytes = src ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src ? src : valid_base), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
#endif

// Y[m][n] = epi(sum_k A(m, k) Wt[k][n]); grid = (row tiles, column tiles, K slices).  A thread owns TM consecutive rows
// and TN columns in groups of 4: group g covers columns g * (BN / (TN / 4)) + tx * 4 .. + 3, so that the lanes of a warp
// read consecutive 16-byte words of a Bs row (no bank conflicts) and store 16-byte pieces of consecutive columns.
// kAsync: the K slices are double-buffered in shared memory, slice t + 1 in flight (cp.async) while slice t is
// multiplied; otherwise one buffer, loads through registers (the first version, kept for A/B and as the fallback).
template <int kSrc, int kEpi, int BM, int BN, int TM, int TN, bool kAsync>
__global__ void __launch_bounds__(kGemmThreads) nice_gemm_kernel(const GemmParams p) {
  static_assert((BM / TM) * (BN / TN) == kGemmThreads, "one register tile per thread");
  static_assert(!(kAsync && kSrc == SRC_NCHW_ACT), "the activation on load needs the register path");
  static_assert(BM % 16 == 0 && TN % 4 == 0 && TM % 4 == 0, "tile shapes");
  constexpr int BK = kGemmBK;
  constexpr int LDA_S = BM + 4;                              // As[k][row], rows 16-byte aligned
  constexpr int NG = TN / 4;                                 // column groups per thread
  constexpr int GW = BN / NG;                                // columns of one group across the CTA
  constexpr int NBUF = kAsync ? 2 : 1;
  MACOW_SHARED __align__(16) float As[NBUF * BK * LDA_S];
  MACOW_SHARED __align__(16) float Bs[NBUF * BK * BN];
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int HW = p.H * p.W;
  // A-tile loader: thread -> (kk = tid % 16, rows r0 + 16 i)
  constexpr int RPT = BM / 16;
  const int a_kk = tid & 15, a_r0 = tid >> 4;
  int a_py[RPT], a_px[RPT];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int pix = (m0 + a_r0 + 16 * i) % HW;
    a_py[i] = pix / p.W;
    a_px[i] = pix - a_py[i] * p.W;
  }
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i) {
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  }
  const int k_begin = blockIdx.z * p.k_split, k_end = min(p.K, k_begin + p.k_split);

  auto load_tile = [&](int k0, int buf) {
    float* as = As + buf * BK * LDA_S;
    float* bs = Bs + buf * BK * BN;
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const int r = a_r0 + 16 * i;
      const float* src = gemm_addr_a<kSrc>(p, m0 + r, k0 + a_kk, a_py[i], a_px[i]);
      if (kAsync) {
        cp_async4(as + a_kk * LDA_S + r, src, p.A);
      } else {
        float v = src ? __ldg(src) : 0.f;
        if (kSrc == SRC_NCHW_ACT) v = activate(v, p.act);    // act(0) = 0 for relu / elu / leaky_relu: padding stays zero
        as[a_kk * LDA_S + r] = v;
      }
    }
    for (int v = tid; v < BK * (BN / 4); v += kGemmThreads) {
      const int kk = v / (BN / 4), n4 = (v - kk * (BN / 4)) * 4;
      const bool ok = k0 + kk < p.K && n0 + n4 < p.ldw;
      const float* src = ok ? p.Wt + (long long)(k0 + kk) * p.ldw + n0 + n4 : nullptr;
      if (kAsync) {
        cp_async16(bs + kk * BN + n4, src, p.Wt);
      } else {
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) w = __ldg(reinterpret_cast<const float4*>(src));
        *reinterpret_cast<float4*>(bs + kk * BN + n4) = w;
      }
    }
  };
  auto multiply_tile = [&](int buf) {
    const float* as = As + buf * BK * LDA_S;
    const float* bs = Bs + buf * BK * BN;
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; i += 4) {
        const float4 v = *reinterpret_cast<const float4*>(as + kk * LDA_S + ty * TM + i);
        a[i] = v.x; a[i + 1] = v.y; a[i + 2] = v.z; a[i + 3] = v.w;
      }
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        const float4 v = *reinterpret_cast<const float4*>(bs + kk * BN + g * GW + tx * 4);
        b[4 * g] = v.x; b[4 * g + 1] = v.y; b[4 * g + 2] = v.z; b[4 * g + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i) {
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
    }
  };

  if (kAsync) {
    if (k_begin < k_end) {
      load_tile(k_begin, 0);
      cp_async_commit();
    }
    int buf = 0;
    for (int k0 = k_begin; k0 < k_end; k0 += BK, buf ^= 1) {
      if (k0 + BK < k_end) {                                 // next slice into the other buffer (free since the barrier
        load_tile(k0 + BK, buf ^ 1);                         // that ended the previous iteration)
        cp_async_commit();
        cp_async_wait<1>();                                  // this thread's copies of the current slice have landed
      } else {
        cp_async_wait<0>();
      }
      __syncthreads();                                       // ... and everybody else's
      multiply_tile(buf);
      __syncthreads();                                       // all reads of this buffer done before it is refilled
    }
  } else {
    for (int k0 = k_begin; k0 < k_end; k0 += BK) {
      load_tile(k0, 0);
      __syncthreads();
      multiply_tile(0);
      __syncthreads();
    }
  }
  const bool vec_ok = (p.ldy & 3) == 0 && (p.y_split_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(p.Y) & 15) == 0;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      const int n = n0 + g * GW + tx * 4;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = kEpi == EPI_ACT ? activate(acc[i][4 * g + j], p.act) : acc[i][4 * g + j];
      float* y = p.Y + blockIdx.z * p.y_split_stride + (long long)m * p.ldy + n;
      if (vec_ok && n + 3 < p.N) {
        *reinterpret_cast<float4*>(y) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (n + j < p.N) y[j] = v[j];
        }
      }
    }
  }
}

// Last step of a coupling.  The weight-normed 3x3 convolution was computed as nine 1x1 "tap" GEMMs,
// T[m][tap * ld3 + 2 i (+ 1)] = a2[m] . w3[tap][:, shift (log-scale) of transformed channel i]; here every output pixel
// adds the taps of its in-image neighbours (zero padding) and the bias, turns the log-scale into
// scale = tanh(ls / 2) * alpha + 1 (macow_utils.py:49-52) and maps the transformed channels in place: forward
// scale * x + shift with logdet[b] += sum log(scale), inverse (x - shift) / (scale + 1e-12) (macow_utils.py:54-66).
// One CTA per sample; the log-determinant is reduced in a fixed order (thread-strided partial sums, shuffle tree,
// warp sums added by thread 0).
struct AffineParams {
  const float* T;      // [splits][M][ldt] of the pass (split-K partial products of the tap GEMM)
  int splits;
  long long split_stride;
  const float* bias;   // [ld3]: (shift, log-scale) pairs
  float* x;            // image, first sample of the pass
  float* logdet;       // [samples of the pass] or null
  int H, W, ldt, ld3;
  long long sample_stride;
  NiceGeom g;
  float alpha;
  int reverse;
};

__global__ void __launch_bounds__(256) nice_affine_kernel(const AffineParams p) {
  const int HW = p.H * p.W, c_t = p.g.c_t, n = HW * c_t;
  const int b = blockIdx.x;
  float* img = p.x + b * p.sample_stride;
  const float* Tb = p.T + (long long)b * HW * p.ldt;
  float lsum = 0.f;
  for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
    const int pix = idx / c_t, ch = idx - pix * c_t;
    const int py = pix / p.W, px = pix - py * p.W;
    float mu = __ldg(p.bias + 2 * ch), ls = __ldg(p.bias + 2 * ch + 1);
    for (int sp = 0; sp < p.splits; ++sp) {
      const float* Ts = Tb + sp * p.split_stride;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3 - 1, dx = tap % 3 - 1;
        const int yy = py + dy, xx = px + dx;
        if (yy < 0 || yy >= p.H || xx < 0 || xx >= p.W) continue;
        const float* t = Ts + (long long)(yy * p.W + xx) * p.ldt + tap * p.ld3 + 2 * ch;
        mu += t[0];
        ls += t[1];
      }
    }
    const float scale = tanhf(0.5f * ls) * p.alpha + 1.0f;
    float* xp = img + (long long)(p.g.p_off + p.g.p_stride * ch) * HW + pix;
    if (p.reverse) {
      *xp = (*xp - mu) / (scale + 1e-12f);
    } else {
      *xp = fmaf(scale, *xp, mu);
      lsum += logf(scale);
    }
  }
  if (!p.reverse && p.logdet != nullptr) {
    MACOW_SHARED float red[8];
    float v = lsum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) s += red[w];
      p.logdet[b] += s;
    }
  }
}

// Per-channel glue, in place, one CTA per sample with the touched channel prefix staged in shared memory:
//   optional ActNorm2dFlow (forward x * exp(ls) + b, inverse (x - b) / (exp(ls) + 1e-8)) on channels [norm_c0, norm_c0 + norm_n),
//   optional permutation out[c] = in[perm[c]] of the first n_perm channels, in either order (norm_first).
struct ChanMapParams {
  float* x;
  long long sample_stride;
  int HW, n_stage;                 // channels staged = max(n_perm, norm_c0 + norm_n)
  const long long* perm;           // int64 [n_perm] (the module's buffer) or null
  int n_perm;
  const float* log_scale;          // [norm_n] or null
  const float* bias;
  int norm_c0, norm_n, norm_inverse, norm_first;
  float* logdet;                   // forward ActNorm: += HW * sum(log_scale); nullable
};

__device__ __forceinline__ float chanmap_norm(const ChanMapParams& p, float v, int c) {
  const float e = expf(__ldg(p.log_scale + c)), b = __ldg(p.bias + c);
  return p.norm_inverse ? (v - b) / (e + 1e-8f) : fmaf(v, e, b);
}

__global__ void macow_chanmap_kernel(const ChanMapParams p) {
  extern __shared__ __align__(16) float smem[];
  const int HW = p.HW, n = p.n_stage * HW;
  float* img = p.x + blockIdx.x * p.sample_stride;
  const bool has_norm = p.log_scale != nullptr && p.norm_n > 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float v = img[i];
    const int c = i / HW;
    if (has_norm && p.norm_first && c >= p.norm_c0 && c < p.norm_c0 + p.norm_n) v = chanmap_norm(p, v, c - p.norm_c0);
    smem[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i / HW, pix = i - c * HW;
    const int src = (p.perm != nullptr && c < p.n_perm) ? (int)p.perm[c] : c;
    float v = smem[src * HW + pix];
    if (has_norm && !p.norm_first && c >= p.norm_c0 && c < p.norm_c0 + p.norm_n) v = chanmap_norm(p, v, c - p.norm_c0);
    img[i] = v;
  }
  if (has_norm && !p.norm_inverse && p.logdet != nullptr && threadIdx.x == 0) {
    float s = 0.f;
    for (int c = 0; c < p.norm_n; ++c) s += __ldg(p.log_scale + c);
    p.logdet[blockIdx.x] += s * (float)HW;
  }
}

// Pack one coupling net.  params: conv1.weight [Hd, c_in, 3, 3], conv2.weight [Hd, Hd, 1, 1], conv3.conv.bias [2 c_t],
// conv3.conv.weight_g [2 c_t], conv3.conv.weight_v [2 c_t, Hd, 3, 3]  (macow_utils.py:270-283, 211-233).
struct PackNiceParams {
  const float* params[5];
  float* packed;
  int c_in, c_t, Hd;
};

__global__ void macow_pack_nice_kernel(const PackNiceParams p) {
  const int c_in = p.c_in, c_t = p.c_t, Hd = p.Hd;
  const int ld = round4(Hd), ld3 = round4(2 * c_t);
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (long long)gridDim.x * blockDim.x;
  float* w1 = p.packed;
  float* w2 = w1 + 9LL * c_in * ld;
  float* w3 = w2 + (long long)Hd * ld;
  for (long long i = tid; i < 9LL * c_in * ld; i += nthreads) {           // Wt1[(c * 9 + tap)][j] = conv1.weight[j][c][tap]
    const int j = (int)(i % ld), k = (int)(i / ld);
    w1[i] = j < Hd ? p.params[0][(long long)j * c_in * 9 + k] : 0.f;
  }
  for (long long i = tid; i < (long long)Hd * ld; i += nthreads) {        // Wt2[jin][jout] = conv2.weight[jout][jin]
    const int jo = (int)(i % ld), ji = (int)(i / ld);
    w2[i] = jo < Hd ? p.params[1][(long long)jo * Hd + ji] : 0.f;
  }
  for (long long i = tid; i < 9LL * Hd * ld3 + ld3; i += nthreads) w3[i] = 0.f;   // padded columns of Wt3 / b3
}

// Second launch of the pack (same stream, after the zero fill above): the weight-normed last convolution
// (torch _weight_norm, dim 0: row o of v scaled by g[o] / ||v[o]||), one thread per output row, and its bias.
__global__ void macow_pack_nice_rows_kernel(const PackNiceParams p) {
  const int c_t = p.c_t, Hd = p.Hd;
  const int ld = round4(Hd), ld3 = round4(2 * c_t);
  float* w3 = p.packed + 9LL * p.c_in * ld + (long long)Hd * ld;
  float* b3 = w3 + 9LL * Hd * ld3;
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= 2 * c_t) return;
  const float* v = p.params[4] + (long long)o * Hd * 9;
  float ss = 0.f;
  for (int i = 0; i < Hd * 9; ++i) ss = fmaf(v[i], v[i], ss);
  const float mul = p.params[3][o] / sqrtf(ss);
  const int slot = o < c_t ? 2 * o : 2 * (o - c_t) + 1;
  for (int j = 0; j < Hd; ++j) {
    for (int tap = 0; tap < 9; ++tap) w3[((long long)j * 9 + tap) * ld3 + slot] = v[j * 9 + tap] * mul;
  }
  b3[slot] = p.params[2][o];
}

}  // namespace macow
