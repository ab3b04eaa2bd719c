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

#ifdef MACOW_HOST_EMULATION
#define MACOW_FAST_EXP(v) expf(v)
#define MACOW_SHARED static               // function-scope shared arrays: one copy for all emulated threads
#else
#define MACOW_FAST_EXP(v) __expf(v)
#define MACOW_SHARED __shared__
#endif

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
  int hid_floats;               // floats of the hidden-activation area (tile x ksplit), 4-float aligned
  int ksplit;                   // inverse: groups of threads sharing the input-channel loop of one line (>= 1)
  long long unit_floats;
  long long in_stride, out_stride;  // floats between consecutive samples of in / out (C * H * W when dense)
  const float* hbias;               // conditional units: [B * H * W][hb_ld], column (unit * 4 + flow) * 2C + 2c = the
  int hb_ld;                        // conditioning's contribution to (shift, log-scale) of channel c; null = unconditional
};

__host__ __device__ inline long long flow_floats(int C, int Hd, int k0, int k1) {
  return (long long)C * k0 * k1 * Hd + (long long)Hd * 2 * C + 2 * C;
}
__host__ __device__ inline long long unit_floats(int C, int Hd, int k0, int k1) {
  return 4 * flow_floats(C, Hd, k0, k1) + 4 * C;
}

__device__ __forceinline__ float activate(float h, int act) {
  if (act == ACT_RELU) return fmaxf(h, 0.f);
  // ELU: exp(h) - 1 with the fast exponential (absolute error < 5e-7 for h < 0, inside the stated 2e-5 tolerance);
  // expm1f costs ~35 instructions per value, 7 % of the forward kernel
  if (act == ACT_ELU) return h > 0.f ? h : MACOW_FAST_EXP(h) - 1.0f;
  return h > 0.f ? h : 0.1f * h;
}

// One masked-convolution flow over the image in shared memory, in place.
// The image is held twice, row-major (XA[c][y][x]) and column-major (XB[c][x][y]), so that every flow reads its lines
// contiguously: Xl = the copy whose rows are this flow's lines ([c][l][q]), Xo = the other one ([c][q][l]).
// kQC = pixels along the line per phase-1 work item (one thread: one hidden channel, kQC pixels, all taps of its
// share of the input channels).  p.ksplit > 1 (inverse: a line offers only Hd * Q / kQC work items) splits the input
// channels over ksplit groups of threads; their partial sums meet in shared memory before the activation.
// kWS: the flow's packed weights (flow_floats() floats) are first copied into shared memory (wsm) and every line step
// reads them from there; otherwise they stream from global memory through L1 (units too wide for one SM's shared memory).
template <bool kWS>
__device__ __forceinline__ float ldw(const float* ptr) {
  if (kWS) return *ptr;
  return __ldg(ptr);
}

template <bool kWS>
__device__ __forceinline__ float2 ldw2(const float* ptr) {          // ptr is 8-byte aligned (even float index)
  if (kWS) return *reinterpret_cast<const float2*>(ptr);
  return __ldg(reinterpret_cast<const float2*>(ptr));
}

template <bool kReverse, int kK1, int kQC, bool kWS>
__device__ __forceinline__ void mcf_stage(float* __restrict__ XA, float* __restrict__ XB, float* __restrict__ hid,
                                          float* __restrict__ wsm, const float* __restrict__ fw_global,
                                          const UnitParams& p, int flow, const float* __restrict__ hb,
                                          float& logdet_acc) {
  const int C = p.C, Hd = p.Hd, k0 = p.k0, HW = p.H * p.W;
  constexpr int k1 = kK1;
  constexpr int cq = (k1 - 1) >> 1;
  constexpr int kJT = 2;                         // hidden channels per phase-1 item
  constexpr int kPT = kReverse ? 1 : 4;          // pixels per phase-2 item (the inverse has one line's worth of items)
  const bool rows = flow < 2;                    // A, B: lines are rows; C, D: lines are columns
  const bool before = (flow & 1) == 0;           // A, C: the window lies before the line; B, D: after it
  const int L = rows ? p.H : p.W, Q = rows ? p.W : p.H;
  float* __restrict__ Xl = rows ? XA : XB;
  float* __restrict__ Xo = rows ? XB : XA;
  const float* fw = fw_global;
  if (kWS) {
    const int nf = (int)flow_floats(C, Hd, k0, k1);
    int done = 0;
    if ((reinterpret_cast<uintptr_t>(fw_global) & 15) == 0) {          // 16-byte copies where the unit's offset allows
      const float4* src4 = reinterpret_cast<const float4*>(fw_global);
      float4* dst4 = reinterpret_cast<float4*>(wsm);
      const int n4 = nf >> 2;
#pragma unroll 8
      for (int i = threadIdx.x; i < n4; i += blockDim.x) dst4[i] = __ldg(src4 + i);
      done = n4 << 2;
    }
    for (int i = done + threadIdx.x; i < nf; i += blockDim.x) wsm[i] = __ldg(fw_global + i);
    __syncthreads();
    fw = wsm;
  }
  const float* __restrict__ wT = fw;                                   // [C][k0][k1][Hd]
  const float* __restrict__ w1T = fw + (long long)C * k0 * k1 * Hd;     // [Hd][C][2]
  const float* __restrict__ b1 = w1T + (long long)Hd * 2 * C;           // [C][2]
  const int T = kReverse ? 1 : p.tile_lines;
  const int S = kReverse ? p.ksplit : 1;
  const int cpk = (C + S - 1) / S;
  const int ntiles = (L + T - 1) / T;
  const bool ascending = kReverse ? before : !before;
  const int nqc = (Q + kQC - 1) / kQC;
  const bool vec = (Q & 3) == 0;                 // float4 reads of a line: every line starts 16-byte aligned
  const int part_stride = T * Q * Hd;            // floats between the partial-sum buffers of two channel groups

  for (int it = 0; it < ntiles; ++it) {
    const int ti = ascending ? it : ntiles - 1 - it;
    const int l0 = ti * T, tl = min(T, L - l0);
    // ---- phase 1: hidden activations of the tile's pixels, hid[pixel][j]; one item = kJT adjacent hidden channels
    const int HJ = Hd / kJT;
    const int items1 = HJ * tl * nqc * S;
    for (int item = threadIdx.x; item < items1; item += blockDim.x) {
      const int j = (item % HJ) * kJT;
      int r = item / HJ;
      const int qc = r % nqc;
      r /= nqc;
      const int t = r % tl, ks = r / tl;
      const int l = l0 + t, q0 = qc * kQC;
      const int c0 = ks * cpk, c1 = min(C, c0 + cpk);
      float acc[kJT][kQC];
#pragma unroll
      for (int i = 0; i < kQC; ++i) {
#pragma unroll
        for (int jt = 0; jt < kJT; ++jt) acc[jt][i] = 0.f;
      }
      for (int dl = 0; dl < k0; ++dl) {
        const int ls = before ? l - k0 + dl : l + 1 + dl;
        if (ls < 0 || ls >= L) continue;                               // zero padding along the masked axis
        const float* xline = Xl + ls * Q;
        const float* wrow = wT + (long long)dl * k1 * Hd + j;
#pragma unroll(kWS ? 2 : 4)
        for (int c = c0; c < c1; ++c) {
          const float* xr = xline + c * HW;
          float xs[kQC + k1 - 1];
          if (vec) {
#pragma unroll
            for (int v = 0; v < kQC / 4; ++v) {
              float4 x4 = make_float4(0.f, 0.f, 0.f, 0.f);
              if (q0 + 4 * v < Q) x4 = *reinterpret_cast<const float4*>(xr + q0 + 4 * v);
              xs[cq + 4 * v + 0] = x4.x;
              xs[cq + 4 * v + 1] = x4.y;
              xs[cq + 4 * v + 2] = x4.z;
              xs[cq + 4 * v + 3] = x4.w;
            }
#pragma unroll
            for (int i = 0; i < cq; ++i) {
              const int ql = q0 - cq + i, qr = q0 + kQC + i;
              xs[i] = ql >= 0 ? xr[ql] : 0.f;
              xs[cq + kQC + i] = qr < Q ? xr[qr] : 0.f;
            }
          } else {
#pragma unroll
            for (int i = 0; i < kQC + k1 - 1; ++i) {
              const int q = q0 - cq + i;
              xs[i] = (q >= 0 && q < Q) ? xr[q] : 0.f;
            }
          }
          const float* wc = wrow + (long long)c * k0 * k1 * Hd;
#pragma unroll
          for (int dq = 0; dq < k1; ++dq) {
            if (kJT == 2) {
              const float2 w = ldw2<kWS>(wc + dq * Hd);
#pragma unroll
              for (int i = 0; i < kQC; ++i) {
                acc[0][i] = fmaf(w.x, xs[i + dq], acc[0][i]);
                acc[kJT - 1][i] = fmaf(w.y, xs[i + dq], acc[kJT - 1][i]);
              }
            } else {
              const float w = ldw<kWS>(wc + dq * Hd);
#pragma unroll
              for (int i = 0; i < kQC; ++i) acc[0][i] = fmaf(w, xs[i + dq], acc[0][i]);
            }
          }
        }
      }
      float* dst = hid + ks * part_stride + (long long)(t * Q + q0) * Hd + j;
#pragma unroll
      for (int i = 0; i < kQC; ++i) {
        if (q0 + i < Q) {
#pragma unroll
          for (int jt = 0; jt < kJT; ++jt) dst[i * Hd + jt] = S == 1 ? activate(acc[jt][i], p.act) : acc[jt][i];
        }
      }
    }
    __syncthreads();
    if (S > 1) {                                   // add the channel groups' partial sums in a fixed order, activate
      const int nh = tl * Q * Hd;
      for (int i = threadIdx.x; i < nh; i += blockDim.x) {
        float s = hid[i];
        for (int ks = 1; ks < S; ++ks) s += hid[ks * part_stride + i];
        hid[i] = activate(s, p.act);
      }
      __syncthreads();
    }
    // ---- phase 2: mu / log_scale per (channel, group of kPT pixels) and the affine map, in place in both copies
    const int npix = tl * Q, ngrp = (npix + kPT - 1) / kPT;
    const int items2 = C * ngrp;
    for (int item = threadIdx.x; item < items2; item += blockDim.x) {
      const int c = item % C;
      const int pix0 = (item / C) * kPT;
      const float* hp[kPT];
      float mu[kPT], lg[kPT];
      const float2 bias = ldw2<kWS>(b1 + 2 * c);
#pragma unroll
      for (int k = 0; k < kPT; ++k) {
        hp[k] = hid + (long long)min(pix0 + k, npix - 1) * Hd;          // tail groups re-read the last pixel, unused
        mu[k] = bias.x;
        lg[k] = bias.y;
      }
      const float* wm = w1T + 2 * c;                                    // [Hd][C][2]: (mu, log-scale) weights of channel c
#pragma unroll 2
      for (int j = 0; j < Hd; j += 4) {
        const float* w = wm + (long long)j * 2 * C;
        const float2 w0 = ldw2<kWS>(w), w1 = ldw2<kWS>(w + 2 * C), w2 = ldw2<kWS>(w + 4 * C), w3 = ldw2<kWS>(w + 6 * C);
#pragma unroll
        for (int k = 0; k < kPT; ++k) {
          const float4 h = *reinterpret_cast<const float4*>(hp[k] + j);
          mu[k] = fmaf(w0.x, h.x, mu[k]); lg[k] = fmaf(w0.y, h.x, lg[k]);
          mu[k] = fmaf(w1.x, h.y, mu[k]); lg[k] = fmaf(w1.y, h.y, lg[k]);
          mu[k] = fmaf(w2.x, h.z, mu[k]); lg[k] = fmaf(w2.y, h.z, lg[k]);
          mu[k] = fmaf(w3.x, h.w, mu[k]); lg[k] = fmaf(w3.y, h.w, lg[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < kPT; ++k) {
        const int pix = pix0 + k;
        if (pix < npix) {
          const int t = pix / Q, q = pix - t * Q, l = l0 + t;
          if (hb != nullptr) {                                          // conditioning term of this image pixel
            const float2 hv = __ldg(reinterpret_cast<const float2*>(hb + (long long)(rows ? l * Q + q : q * L + l) * p.hb_ld + 2 * c));
            mu[k] += hv.x;
            lg[k] += hv.y;
          }
          const float scale = tanhf(0.5f * lg[k]) * p.alpha + 1.0f;
          float* xl = Xl + c * HW + l * Q + q;
          float v = *xl;
          if (kReverse) {
            v = (v - mu[k]) / (scale + 1e-12f);
          } else {
            v = fmaf(scale, v, mu[k]);
            logdet_acc += logf(scale);
          }
          *xl = v;
          Xo[c * HW + q * L + l] = v;
        }
      }
    }
    __syncthreads();
  }
}

template <bool kReverse>
__device__ __forceinline__ void actnorm_stage(float* __restrict__ XA, float* __restrict__ XB, const float* __restrict__ an,
                                              const UnitParams& p, float& logdet_acc) {
  const int H = p.H, W = p.W, HW = H * W, n = p.C * HW;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i / HW, rem = i - c * HW;
    const int y = rem / W, x = rem - y * W;
    const float e = expf(__ldg(an + c)), b = __ldg(an + p.C + c);
    const float v = kReverse ? (XA[i] - b) / (e + 1e-8f) : fmaf(XA[i], e, b);
    XA[i] = v;
    XB[c * HW + x * H + y] = v;
  }
  if (!kReverse && threadIdx.x == 0) {
    float s = 0.f;
    for (int c = 0; c < p.C; ++c) s += __ldg(an + c);
    logdet_acc += s * (float)HW;
  }
  __syncthreads();
}

// grid = B, dynamic smem = (2 images + hidden tile x ksplit [+ one flow's weights]) floats.  Hd must be a multiple of 4 (float4 reads of hid);
// kK1 = kernel_size[1], the centred taps along the line (odd; the reference's configs use 3).
template <bool kReverse, int kK1, int kQC, bool kWS>
__global__ void __launch_bounds__(512, 1) macow_unit_kernel(const UnitParams p) {
  extern __shared__ __align__(16) float smem[];
  const int H = p.H, W = p.W, HW = H * W, n = p.C * HW;
  const int n4 = (n + 3) & ~3;
  float* XA = smem;
  float* XB = smem + n4;
  float* hid = smem + 2 * n4;
  float* wsm = hid + p.hid_floats;          // kWS only: one flow's weights
  const int b = blockIdx.x;
  const float* src = p.in + (long long)b * p.in_stride;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i / HW, rem = i - c * HW;
    const int y = rem / W, x = rem - y * W;
    const float v = __ldg(src + i);
    XA[i] = v;
    XB[c * HW + x * H + y] = v;
  }
  __syncthreads();
  float logdet_acc = 0.f;
  const int ns = p.stage_end - p.stage_begin;
  for (int k = 0; k < ns; ++k) {
    const int s = kReverse ? p.stage_end - 1 - k : p.stage_begin + k;
    const int unit = s / 6, st = s - unit * 6;
    const float* up = p.packed + unit * p.unit_floats;
    const long long ff = flow_floats(p.C, p.Hd, p.k0, p.k1);
    if (st == 2 || st == 5) {
      actnorm_stage<kReverse>(XA, XB, up + 4 * ff + (st == 5 ? 2 * p.C : 0), p, logdet_acc);
    } else {
      const int flow = st < 2 ? st : st - 1;
      const float* hb = p.hbias == nullptr ? nullptr
                                           : p.hbias + (long long)b * HW * p.hb_ld + (unit * 4 + flow) * 2 * p.C;
      mcf_stage<kReverse, kK1, kQC, kWS>(XA, XB, hid, wsm, up + flow * ff, p, flow, hb, logdet_acc);
    }
  }
  float* dst = p.out + (long long)b * p.out_stride;
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = XA[i];
  if (!kReverse && p.logdet != nullptr) {
    // deterministic block reduction (fixed item -> thread mapping, fixed tree)
    MACOW_SHARED float red[16];
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
  int hc;            // conditioning channels: weight_v rows are [Hd + hc] wide, the last hc columns go to wh
  float* wh;         // [hc][wh_ld]: the chain's conditioning matrix (null when hc == 0)
  int wh_ld, wh_col0;  // this unit's first column = unit_index * 8 C
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
    const int vw = Hd + p.hc;                                         // row width of weight_v
    for (int o = tid; o < 2 * C; o += nthreads) {
      float ss = 0.f;
      for (int j = 0; j < vw; ++j) ss = fmaf(v[o * vw + j], v[o * vw + j], ss);
      const float mul = g[o] / sqrtf(ss);
      const int slot = o < C ? 2 * o : 2 * (o - C) + 1;               // output o: mu of channel o, or log-scale of o - C
      for (int j = 0; j < Hd; ++j) w1T[(long long)j * 2 * C + slot] = v[o * vw + j] * mul;
      for (int k = 0; k < p.hc; ++k) p.wh[(long long)k * p.wh_ld + p.wh_col0 + f * 2 * C + slot] = v[o * vw + Hd + k] * mul;
      b1[slot] = bias[o];
    }
  }
  float* an = p.packed + 4 * ff;
  for (int i = tid; i < 4 * C; i += nthreads) an[i] = p.params[16 + i / C][i % C];
}

}  // namespace macow
