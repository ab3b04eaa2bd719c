// Fused Adam / AMSGrad step over every parameter tensor of the flow in ONE launch (multi-tensor apply).
// The reference trains the flat INN with torch.optim.Adam(lr, betas=(0.9, 0.999), weight_decay, amsgrad=True)
// (models/second_stage_video_fc.py:685-687; torch/optim/adam.py `_single_tensor_adam`): per parameter 8+ eager
// elementwise kernels over 360 tensors.  Here one kernel reads p, g, m, v, vmax once and writes p, m, v, vmax once:
// 36 bytes per parameter, HBM-bound.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ipoke_inn.h"
#include "ptx.cuh"

namespace inn {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_CHUNK = 8192;          // elements per CTA: 8 float4 per thread

struct AdamParams {
  const ipoke_inn_adam_job* jobs;         // device table, one entry per parameter tensor
  const int* chunk_job;                   // chunk -> job
  const long long* chunk_off;             // chunk -> first element inside the job
  int n_chunks;
  float beta1, beta2, eps, weight_decay;
  float step_size;                        // lr / (1 - beta1^t)
  float inv_sqrt_bc2;                     // 1 / sqrt(1 - beta2^t)
  float grad_scale;                       // gradients are multiplied by this first (1 / world_size after a sum all-reduce)
  int amsgrad;
};

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float& vmax, const AdamParams& a) {
  g = g * a.grad_scale;
  g = fmaf(a.weight_decay, p, g);                          // grad.add(param, alpha=weight_decay)
  m = m + (g - m) * (1.0f - a.beta1);                      // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf((1.0f - a.beta2) * g, g, v * a.beta2);          // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  float vv = v;
  if (a.amsgrad) {
    vmax = fmaxf(vmax, v);                                 // torch.maximum(max_exp_avg_sq, exp_avg_sq)
    vv = vmax;
  }
  const float denom = sqrtf(vv) * a.inv_sqrt_bc2 + a.eps;  // (sqrt / bias_correction2_sqrt).add_(eps)
  p = p - a.step_size * (m / denom);                       // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(ADAM_THREADS) inn_adam_kernel(const AdamParams a) {
  const int c = blockIdx.x;
  const ipoke_inn_adam_job j = a.jobs[a.chunk_job[c]];
  const long long off = a.chunk_off[c];
  const long long left = j.n - off;
  const int n = left < ADAM_CHUNK ? static_cast<int>(left) : ADAM_CHUNK;
  float* p = j.p + off;
  const float* g = j.g + off;
  float* m = j.m + off;
  float* v = j.v + off;
  float* vm = a.amsgrad ? j.vmax + off : nullptr;
  const uintptr_t al = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                       reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(vm);
  if ((al & 15u) == 0) {
    const int n4 = n >> 2;
    // all loads of the chunk first (8 x 5 independent 16-byte loads per thread), then math and stores
    for (int i0 = threadIdx.x; i0 < n4; i0 += 2 * ADAM_THREADS) {
      float4 P[2], G[2], M[2], V[2], X[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = i0 + u * ADAM_THREADS;
        if (i < n4) {
          P[u] = reinterpret_cast<const float4*>(p)[i];
          G[u] = __ldcs(reinterpret_cast<const float4*>(g) + i);      // gradients are read once: streaming
          M[u] = reinterpret_cast<const float4*>(m)[i];
          V[u] = reinterpret_cast<const float4*>(v)[i];
          X[u] = a.amsgrad ? reinterpret_cast<const float4*>(vm)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = i0 + u * ADAM_THREADS;
        if (i < n4) {
          adam_update(P[u].x, G[u].x, M[u].x, V[u].x, X[u].x, a);
          adam_update(P[u].y, G[u].y, M[u].y, V[u].y, X[u].y, a);
          adam_update(P[u].z, G[u].z, M[u].z, V[u].z, X[u].z, a);
          adam_update(P[u].w, G[u].w, M[u].w, V[u].w, X[u].w, a);
          reinterpret_cast<float4*>(p)[i] = P[u];
          reinterpret_cast<float4*>(m)[i] = M[u];
          reinterpret_cast<float4*>(v)[i] = V[u];
          if (a.amsgrad) reinterpret_cast<float4*>(vm)[i] = X[u];
        }
      }
    }
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += ADAM_THREADS) {
      float x = a.amsgrad ? vm[i] : 0.f;
      adam_update(p[i], g[i], m[i], v[i], x, a);
      if (a.amsgrad) vm[i] = x;
    }
  } else {
    for (int i = threadIdx.x; i < n; i += ADAM_THREADS) {
      float x = a.amsgrad ? vm[i] : 0.f;
      adam_update(p[i], g[i], m[i], v[i], x, a);
      if (a.amsgrad) vm[i] = x;
    }
  }
}

// ---- sharded / packed-master step (ShardedAdam) ------------------------------------------------------------------
// One contiguous span of the weight arena owned by this rank: the fp32 master, its gradient (fp32, or bf16 off the
// wire), the moments of the span, and the bf16 hi / lo images of the span inside the packed buffer the GEMMs read.
// One pass: read p, g, m, v, vmax (20 B, 18 B with bf16 gradients), write p, m, v, vmax, hi (+ lo) (18 - 20 B).
// The separate fp32 -> bf16 repack pass of the replicated optimizer (read 4 B, write 2 - 4 B per weight) is gone.
struct AdamShardParams {
  float* p; const float* g; const __nv_bfloat16* g16;   // exactly one of g / g16
  float* m; float* v; float* vmax;
  __nv_bfloat16* hi; __nv_bfloat16* lo;                  // lo nullable (single-pass bf16 mode); hi nullable (no images)
  long long n;                                           // multiple of 4, all pointers 16-byte aligned
  AdamParams a;
};

__device__ __forceinline__ void split4(const float4& P, uint2& H, uint2& L) {
  __nv_bfloat16 h[4], l[4];
  split_bf16(P.x, h[0], l[0]);
  split_bf16(P.y, h[1], l[1]);
  split_bf16(P.z, h[2], l[2]);
  split_bf16(P.w, h[3], l[3]);
  H = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
  L = make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
}

__global__ void __launch_bounds__(ADAM_THREADS) inn_adam_shard_kernel(const AdamShardParams q) {
  const long long n4 = q.n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * ADAM_THREADS;
  for (long long i0 = static_cast<long long>(blockIdx.x) * ADAM_THREADS + threadIdx.x; i0 < n4; i0 += 2 * stride) {
    float4 P[2], G[2], M[2], V[2], X[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i < n4) {
        P[u] = reinterpret_cast<const float4*>(q.p)[i];
        if (q.g) {
          G[u] = __ldcs(reinterpret_cast<const float4*>(q.g) + i);
        } else {
          const uint2 r = __ldcs(reinterpret_cast<const uint2*>(q.g16) + i);
          G[u] = make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u),
                             __uint_as_float(r.y << 16), __uint_as_float(r.y & 0xffff0000u));
        }
        M[u] = reinterpret_cast<const float4*>(q.m)[i];
        V[u] = reinterpret_cast<const float4*>(q.v)[i];
        X[u] = q.a.amsgrad ? reinterpret_cast<const float4*>(q.vmax)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i < n4) {
        adam_update(P[u].x, G[u].x, M[u].x, V[u].x, X[u].x, q.a);
        adam_update(P[u].y, G[u].y, M[u].y, V[u].y, X[u].y, q.a);
        adam_update(P[u].z, G[u].z, M[u].z, V[u].z, X[u].z, q.a);
        adam_update(P[u].w, G[u].w, M[u].w, V[u].w, X[u].w, q.a);
        reinterpret_cast<float4*>(q.p)[i] = P[u];
        reinterpret_cast<float4*>(q.m)[i] = M[u];
        reinterpret_cast<float4*>(q.v)[i] = V[u];
        if (q.a.amsgrad) reinterpret_cast<float4*>(q.vmax)[i] = X[u];
        if (q.hi) {                              // NULL: parameters without packed images (biases, ActNorm)
          uint2 H, L;
          split4(P[u], H, L);
          reinterpret_cast<uint2*>(q.hi)[i] = H;
          if (q.lo) reinterpret_cast<uint2*>(q.lo)[i] = L;
        }
      }
    }
  }
}

// fp32 master span -> bf16 hi / lo images (after load_state_dict or other out-of-band edits of the master)
__global__ void __launch_bounds__(ADAM_THREADS) inn_master_split_kernel(const float* __restrict__ p, __nv_bfloat16* hi,
                                                                        __nv_bfloat16* lo, long long n) {
  const long long n4 = n >> 2;
  for (long long i = static_cast<long long>(blockIdx.x) * ADAM_THREADS + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * ADAM_THREADS) {
    uint2 H, L;
    split4(reinterpret_cast<const float4*>(p)[i], H, L);
    reinterpret_cast<uint2*>(hi)[i] = H;
    if (lo) reinterpret_cast<uint2*>(lo)[i] = L;
  }
}

// fp32 gradient span -> bf16 wire format (scaled first, e.g. by 1 / world_size so that the sum on the wire is the mean)
__global__ void __launch_bounds__(ADAM_THREADS) inn_grad_to_bf16_kernel(const float* __restrict__ g, __nv_bfloat16* out,
                                                                        long long n, float scale) {
  const long long n4 = n >> 2;
  for (long long i = static_cast<long long>(blockIdx.x) * ADAM_THREADS + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * ADAM_THREADS) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(g) + i);
    const __nv_bfloat162 a = __floats2bfloat162_rn(v.x * scale, v.y * scale);
    const __nv_bfloat162 b = __floats2bfloat162_rn(v.z * scale, v.w * scale);
    reinterpret_cast<uint2*>(out)[i] = make_uint2(*reinterpret_cast<const unsigned*>(&a), *reinterpret_cast<const unsigned*>(&b));
  }
}

}  // namespace inn
