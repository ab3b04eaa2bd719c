// MaCow unit (masked-convolution autoregressive flow) for sm_100a: one CTA per sample, the whole chain of
// stages of one or more units in ONE launch with the image resident in shared memory.
//
// Reference (altmann144/ipoke-main-opticalflow, paths relative to models/modules/INN):
//   MaCowUnit.forward            macow2.py:963-996   conv1 (A) -> conv2 (B) -> actnorm1 -> conv3 (C) -> conv4 (D) -> actnorm2
//   MaskedConvFlow.forward       macow2.py:116-119   params = MCFBlock(x); out = scale * x + mu; logdet = sum log scale
//   MaskedConvFlow inverse       macow2.py:222-289   one row (A top-down, B bottom-up) / column (C, D) per step:
//                                                    F.conv2d + activation + 1x1 conv + affine inverse PER LINE, i.e.
//                                                    ~10 launches per line, 4 * 16 lines per unit at 16 x 16
//   MCFBlock.forward             macow_utils.py:429-436   shifted conv (no bias) -> activation -> weight-normed 1x1 conv
//   ShiftedConv2d                macow_utils.py:460-497   pad + cut so that the kH x kW window lies strictly before /
//                                                    after the output pixel along the masked axis, centred in the other
//   Affine                       macow_utils.py:49-66     scale = tanh(log_scale / 2) * alpha + 1; inverse (z - mu) / (scale + 1e-12)
//   ActNorm2dFlow                macow2.py:502-521   x * exp(ls) + b, logdet = H W sum(ls); inverse (x - b) / (exp(ls) + 1e-8)
//
// Design: the inverse is a chain of L dependent line steps per flow, each a tiny (Hd x C k0 k1) x Q product followed
// by a (2C x Hd) x Q product: launch- and latency-bound in the reference.  Here a CTA owns one sample; a line step
// is two phases separated by __syncthreads(): (1) hidden[q][j] = act(sum w[c][dl][dq][j] * X[c][l +- dl][q + dq])
// into shared memory, (2) mu / log_scale per (channel, pixel), the affine map applied in place.  The image never
// leaves shared memory between the 6 * n_units stages; weights stream from the packed buffer through L1 (every CTA
// of the grid reads the same few 10 KB, lanes over the output-channel index so each warp load is one 128 B line).
// In unified coordinates every flow is "k0 taps along the masked (line) axis, k1 centred taps along the line":
// rows are lines for A / B, columns for C / D (whose kernels are (k1, k0), macow2.py:948-956); the pack kernel
// transposes the C / D weights accordingly.
#pragma once
#ifndef MACOW_HOST_EMULATION   // tests/emulation/macow_emul.cpp compiles this header with g++ behind single-thread shims
#include <cuda_runtime.h>
#endif
#include <stdint.h>

namespace macow {

enum Act { ACT_RELU = 0, ACT_ELU = 1, ACT_LEAKY = 2 };

struct UnitParams {
  const float* packed;   // n_units packed units, unit_floats apart
  const float* in;       // [B, C, H, W]
  float* out;            // [B, C, H, W]
  float* logdet;         // [B] or null (forward only)
  int B, C, Hd, H, W, k0, k1;
  float alpha;
  int stage_begin, stage_end;   // global stage range [begin, end) over n_units * 6, forward order
  int tile_lines;               // lines per forward tile (the inverse always steps one line)
  int accumulate;               // logdet += instead of =
  int act;                      // Act
  long long unit_floats;
};

__host__ __device__ inline long long flow_floats(int C, int Hd, int k0, int k1) {
  return (long long)C * k0 * k1 * Hd + (long long)Hd * 2 * C + 2 * C;
}
__host__ __device__ inline long long unit_floats(int C, int Hd, int k0, int k1) {
  return 4 * flow_floats(C, Hd, k0, k1) + 4 * C;
}

__device__ __forceinline__ float activate(float h, int act) {
  if (act == ACT_RELU) return fmaxf(h, 0.f);
  if (act == ACT_ELU) return h > 0.f ? h : expm1f(h);
  return h > 0.f ? h : 0.1f * h;
}

// One masked-convolution flow over the image in shared memory, in place.
// kQC = pixels along the line per phase-1 work item (one thread: one hidden channel, kQC pixels, all taps).
template <bool kReverse, int kK1, int kQC>
__device__ __forceinline__ void mcf_stage(float* __restrict__ X, float* __restrict__ hid, const float* __restrict__ fw,
                                          const UnitParams& p, int flow, float& logdet_acc) {
  const int C = p.C, Hd = p.Hd, H = p.H, W = p.W, k0 = p.k0, HW = H * W;
  constexpr int k1 = kK1;
  const bool rows = flow < 2;                    // A, B: lines are rows; C, D: lines are columns
  const bool before = (flow & 1) == 0;           // A, C: the window lies before the line; B, D: after it
  const int L = rows ? H : W, Q = rows ? W : H;
  const int sl = rows ? W : 1, sq = rows ? 1 : W;
  const int cq = (k1 - 1) >> 1;
  const float* __restrict__ wT = fw;                                   // [C][k0][k1][Hd]
  const float* __restrict__ w1T = fw + (long long)C * k0 * k1 * Hd;     // [Hd][2C]
  const float* __restrict__ b1 = w1T + (long long)Hd * 2 * C;           // [2C]
  const int T = kReverse ? 1 : p.tile_lines;
  const int ntiles = (L + T - 1) / T;
  const bool ascending = kReverse ? before : !before;
  const int nqc = (Q + kQC - 1) / kQC;

  for (int it = 0; it < ntiles; ++it) {
    const int ti = ascending ? it : ntiles - 1 - it;
    const int l0 = ti * T, tl = min(T, L - l0);
    // ---- phase 1: hidden activations of the tile's pixels, hid[pixel][j]
    const int items1 = Hd * tl * nqc;
    for (int item = threadIdx.x; item < items1; item += blockDim.x) {
      const int j = item % Hd;
      const int r = item / Hd;
      const int qc = r % nqc, t = r / nqc;
      const int l = l0 + t, q0 = qc * kQC;
      float acc[kQC];
#pragma unroll
      for (int i = 0; i < kQC; ++i) acc[i] = 0.f;
      for (int dl = 0; dl < k0; ++dl) {
        const int ls = before ? l - k0 + dl : l + 1 + dl;
        if (ls < 0 || ls >= L) continue;                               // zero padding along the masked axis
        const float* xrow = X + ls * sl;
        const float* wrow = wT + (long long)dl * k1 * Hd + j;
        for (int c = 0; c < C; ++c) {
          float xs[kQC + k1 - 1];
#pragma unroll
          for (int i = 0; i < kQC + k1 - 1; ++i) {
            const int q = q0 - cq + i;
            xs[i] = (q >= 0 && q < Q) ? xrow[c * HW + q * sq] : 0.f;
          }
          const float* wc = wrow + (long long)c * k0 * k1 * Hd;
#pragma unroll
          for (int dq = 0; dq < k1; ++dq) {
            const float w = __ldg(wc + dq * Hd);
#pragma unroll
            for (int i = 0; i < kQC; ++i) acc[i] = fmaf(w, xs[i + dq], acc[i]);
          }
        }
      }
#pragma unroll
      for (int i = 0; i < kQC; ++i)
        if (q0 + i < Q) hid[(long long)(t * Q + q0 + i) * Hd + j] = activate(acc[i], p.act);
    }
    __syncthreads();
    // ---- phase 2: mu / log_scale per (channel, pixel) and the affine map, in place
    const int items2 = C * tl * Q;
    for (int item = threadIdx.x; item < items2; item += blockDim.x) {
      const int c = item % C;
      const int pix = item / C;
      const int t = pix / Q, q = pix - t * Q;
      const float* h = hid + (long long)pix * Hd;
      float mu = __ldg(b1 + c), lsc = __ldg(b1 + C + c);
      const float* wm = w1T + c;
      for (int j = 0; j < Hd; j += 4) {
        const float4 hv = *reinterpret_cast<const float4*>(h + j);
        mu = fmaf(__ldg(wm + (j + 0) * 2 * C), hv.x, mu);
        lsc = fmaf(__ldg(wm + (j + 0) * 2 * C + C), hv.x, lsc);
        mu = fmaf(__ldg(wm + (j + 1) * 2 * C), hv.y, mu);
        lsc = fmaf(__ldg(wm + (j + 1) * 2 * C + C), hv.y, lsc);
        mu = fmaf(__ldg(wm + (j + 2) * 2 * C), hv.z, mu);
        lsc = fmaf(__ldg(wm + (j + 2) * 2 * C + C), hv.z, lsc);
        mu = fmaf(__ldg(wm + (j + 3) * 2 * C), hv.w, mu);
        lsc = fmaf(__ldg(wm + (j + 3) * 2 * C + C), hv.w, lsc);
      }
      const float scale = tanhf(0.5f * lsc) * p.alpha + 1.0f;
      float* xp = X + c * HW + (l0 + t) * sl + q * sq;
      if (kReverse) {
        *xp = (*xp - mu) / (scale + 1e-12f);
      } else {
        *xp = fmaf(scale, *xp, mu);
        logdet_acc += logf(scale);
      }
    }
    __syncthreads();
  }
}

template <bool kReverse>
__device__ __forceinline__ void actnorm_stage(float* __restrict__ X, const float* __restrict__ an, const UnitParams& p,
                                              float& logdet_acc) {
  const int HW = p.H * p.W, n = p.C * HW;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i / HW;
    const float e = expf(__ldg(an + c)), b = __ldg(an + p.C + c);
    X[i] = kReverse ? (X[i] - b) / (e + 1e-8f) : fmaf(X[i], e, b);
  }
  if (!kReverse && threadIdx.x == 0) {
    float s = 0.f;
    for (int c = 0; c < p.C; ++c) s += __ldg(an + c);
    logdet_acc += s * (float)HW;
  }
  __syncthreads();
}

// grid = B, dynamic smem = (C H W + hidden tile) floats.  Hd must be a multiple of 4 (float4 reads of hid);
// kK1 = kernel_size[1], the centred taps along the line (odd; the reference's configs use 3).
template <bool kReverse, int kK1, int kQC>
__global__ void __launch_bounds__(256) macow_unit_kernel(const UnitParams p) {
  extern __shared__ __align__(16) float smem[];
  const int n = p.C * p.H * p.W;
  float* X = smem;
  float* hid = smem + ((n + 3) & ~3);
  const int b = blockIdx.x;
  const float* src = p.in + (long long)b * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) X[i] = __ldg(src + i);
  __syncthreads();
  float logdet_acc = 0.f;
  const int ns = p.stage_end - p.stage_begin;
  for (int k = 0; k < ns; ++k) {
    const int s = kReverse ? p.stage_end - 1 - k : p.stage_begin + k;
    const int unit = s / 6, st = s - unit * 6;
    const float* up = p.packed + unit * p.unit_floats;
    const long long ff = flow_floats(p.C, p.Hd, p.k0, p.k1);
    if (st == 2 || st == 5) {
      actnorm_stage<kReverse>(X, up + 4 * ff + (st == 5 ? 2 * p.C : 0), p, logdet_acc);
    } else {
      const int flow = st < 2 ? st : st - 1;
      mcf_stage<kReverse, kK1, kQC>(X, hid, up + flow * ff, p, flow, logdet_acc);
    }
  }
  float* dst = p.out + (long long)b * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = X[i];
  if (!kReverse && p.logdet != nullptr) {
    // deterministic block reduction (fixed item -> thread mapping, fixed tree)
    __shared__ float red[8];
    float v = logdet_acc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) s += red[w];
      p.logdet[b] = p.accumulate ? p.logdet[b] + s : s;
    }
  }
}

// Pack one unit's parameters (reference state_dict layouts) into the kernel layout.
// params[f*4 + {0,1,2,3}] = flow f: shift_conv.weight [Hd, C, kH, kW], conv1x1.conv.bias [2C], weight_g [2C], weight_v [2C, Hd];
// params[16..19] = actnorm1.log_scale, actnorm1.bias, actnorm2.log_scale, actnorm2.bias ([C] each).
struct PackParams {
  const float* params[20];
  float* packed;
  int C, Hd, k0, k1;
};

__global__ void macow_pack_kernel(const PackParams p) {
  const int C = p.C, Hd = p.Hd, k0 = p.k0, k1 = p.k1;
  const long long ff = flow_floats(C, Hd, k0, k1);
  const int nshift = C * k0 * k1 * Hd;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  for (int f = 0; f < 4; ++f) {
    float* dst = p.packed + f * ff;
    const float* w = p.params[f * 4 + 0];
    const bool rows = f < 2;
    // dst[((c*k0 + dl)*k1 + dq)*Hd + j] = w[j][c][dy][dx], (dy, dx) = (dl, dq) for A / B (kernel (k0, k1)),
    // (dq, dl) for C / D (kernel (k1, k0), macow2.py:948-956)
    for (int i = tid; i < nshift; i += nthreads) {
      const int j = i % Hd;
      int r = i / Hd;
      const int dq = r % k1; r /= k1;
      const int dl = r % k0;
      const int c = r / k0;
      const int src = rows ? ((j * C + c) * k0 + dl) * k1 + dq : ((j * C + c) * k1 + dq) * k0 + dl;
      dst[i] = w[src];
    }
    // weight norm of the 1x1 convolution (macow_utils.py:233, torch _weight_norm: v * (g / ||v||) per output row):
    // one thread per output row (2C rows of <= 512 elements; this kernel runs once per parameter change)
    const float* bias = p.params[f * 4 + 1];
    const float* g = p.params[f * 4 + 2];
    const float* v = p.params[f * 4 + 3];
    float* w1T = dst + nshift;
    float* b1 = w1T + (long long)Hd * 2 * C;
    for (int o = tid; o < 2 * C; o += nthreads) {
      float ss = 0.f;
      for (int j = 0; j < Hd; ++j) ss = fmaf(v[o * Hd + j], v[o * Hd + j], ss);
      const float mul = g[o] / sqrtf(ss);
      for (int j = 0; j < Hd; ++j) w1T[(long long)j * 2 * C + o] = v[o * Hd + j] * mul;
      b1[o] = bias[o];
    }
  }
  float* an = p.packed + 4 * ff;
  for (int i = tid; i < 4 * C; i += nthreads) an[i] = p.params[16 + i / C][i % C];
}

}  // namespace macow
