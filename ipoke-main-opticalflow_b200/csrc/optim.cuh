// Fused Adam / AMSGrad step over every parameter tensor of the flow in ONE launch (multi-tensor apply).
// The reference trains the flat INN with torch.optim.Adam(lr, betas=(0.9, 0.999), weight_decay, amsgrad=True)
// (models/second_stage_video_fc.py:685-687; torch/optim/adam.py `_single_tensor_adam`): per parameter 8+ eager
// elementwise kernels over 360 tensors.  Here one kernel reads p, g, m, v, vmax once and writes p, m, v, vmax once:
// 36 bytes per parameter, HBM-bound.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ipoke_inn.h"

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

}  // namespace inn
