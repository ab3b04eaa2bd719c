// HBM-bound fp32 kernels around the coupling GEMMs: block boundaries (Shuffle gather + ActNorm +
// InvLeakyRelu and their inverses), parameter packing, ActNorm statistics, FlowLoss.
#pragma once
#include "ptx.cuh"

namespace inn {

// ----------------------------------------------------------------------------- boundary
// out[b,p] = post_p( pre_q( in[b,q] ) ),  q = src[p]
//   forward, start of block k : src = Shuffle_{k-1} (composed with the half swap of the coupling output),
//                               post = ActNorm_k then InvLeakyRelu         (flow_blocks.py:257-260, 322-324;
//                                                                           modules.py:292; flow_blocks.py:380-381)
//   reverse, start of block k : pre = InvLeakyRelu^-1 then ActNorm_{k+1}^-1 (flow_blocks.py:385-386; modules.py:312),
//                               src = Shuffle_k^-1                          (flow_blocks.py:326)
// and writes the bf16 (hi/lo) copy of the half the next subnetwork reads, the conditioning columns of the
// MLP input buffer, and the ActNorm log-det term.
struct BoundaryParams {
  const float* in;  int ld_in;
  float* out;       int ld_out;        // nullable
  float* raw_out;   int ld_raw;        // nullable: value after gather+pre, before post
  float* raw_out2;                      // nullable second copy (same ld)
  const int* src;                       // nullable = identity
  const float* pre_loc;  const float* pre_scale;  int pre_leaky;     // pre_scale == nullptr -> no pre
  const float* post_loc; const float* post_scale; int post_leaky;    // post_scale == nullptr -> no post
  float alpha;
  __nv_bfloat16* u_hi; __nv_bfloat16* u_lo; int ld_u;                // nullable
  int u_begin, u_count, u_pad;          // columns p in [u_begin, u_begin+u_count) -> U col p-u_begin; zero to u_pad
  const float* cond; int E; int ld_cond; int u_cond_off; int e_pad;  // cond nullable
  int cond_rows;                        // > 0: cond has cond_rows rows and row b reads row b % cond_rows (multi-draw calls)
  __nv_bfloat16* u2_hi; __nv_bfloat16* u2_lo;                         // nullable: second buffer that gets the cond columns
  float* logdet; const float* an_logdet; int logdet_init;            // logdet nullable
  float* block_logdet;                                                // nullable snapshot before the add
  float* block_logdet2;                                               // nullable second snapshot
  int B, C;
  int fast_rows;                        // 1: warp-per-row path (C % 4 == 0, all row strides % 4 == 0, u_begin / u_count % 4 == 0,
                                        //    dynamic smem = 13 * C floats): coalesced row load -> smem -> permuted read ->
                                        //    coalesced float4 stores, instead of one uncoalesced gather per element
};

__global__ void __launch_bounds__(256) inn_boundary_kernel(const BoundaryParams p) {
  grid_dep_launch();      // PDL: the next GEMM may start its prologue / weight prefetch
  grid_dep_wait();        // previous kernel's state is complete and visible
  const long long total = static_cast<long long>(p.B) * p.C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long tid0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (p.fast_rows) {
    extern __shared__ float bsm[];
    const int C = p.C;
    float* c_pre_s = bsm;                 // indexed by SOURCE column q
    float* c_pre_l = bsm + C;
    float* c_post_s = bsm + 2 * C;        // indexed by destination column c
    float* c_post_l = bsm + 3 * C;
    int* c_src = reinterpret_cast<int*>(bsm + 4 * C);
    float* rowbuf = bsm + 5 * C + (threadIdx.x >> 5) * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      c_src[c] = p.src ? __ldg(p.src + c) : c;
      c_pre_s[c] = p.pre_scale ? __ldg(p.pre_scale + c) : 1.f;
      c_pre_l[c] = p.pre_scale ? __ldg(p.pre_loc + c) : 0.f;
      c_post_s[c] = p.post_scale ? __ldg(p.post_scale + c) : 1.f;
      c_post_l[c] = p.post_scale ? __ldg(p.post_loc + c) : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const float an_ld = (p.logdet && p.an_logdet) ? __ldg(p.an_logdet) : 0.f;
    for (int b = blockIdx.x * wpb + (threadIdx.x >> 5); b < p.B; b += gridDim.x * wpb) {
      const float4* src4 = reinterpret_cast<const float4*>(p.in + static_cast<size_t>(b) * p.ld_in);
      for (int g = lane; g < (C >> 2); g += 32) reinterpret_cast<float4*>(rowbuf)[g] = __ldg(src4 + g);
      __syncwarp();
      for (int g = lane; g < (C >> 2); g += 32) {
        const int c = g << 2;
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int q = c_src[c + e];
          float x = rowbuf[q];
          if (p.pre_scale) {
            if (p.pre_leaky) x = x / (x >= 0.f ? 1.0f : p.alpha);      // same operations as the element path
            x = x / c_pre_s[q] - c_pre_l[q];
          }
          v[e] = x;
        }
        if (p.raw_out) *reinterpret_cast<float4*>(p.raw_out + static_cast<size_t>(b) * p.ld_raw + c) = make_float4(v[0], v[1], v[2], v[3]);
        if (p.raw_out2) *reinterpret_cast<float4*>(p.raw_out2 + static_cast<size_t>(b) * p.ld_raw + c) = make_float4(v[0], v[1], v[2], v[3]);
        if (p.post_scale) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float x = c_post_s[c + e] * (v[e] + c_post_l[c + e]);
            if (p.post_leaky) x = x * (x >= 0.f ? 1.0f : p.alpha);
            v[e] = x;
          }
        }
        if (p.out) *reinterpret_cast<float4*>(p.out + static_cast<size_t>(b) * p.ld_out + c) = make_float4(v[0], v[1], v[2], v[3]);
        if (p.u_hi && c >= p.u_begin && c < p.u_begin + p.u_count) {
          __nv_bfloat16 h[4], l[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) split_bf16(v[e], h[e], l[e]);
          const size_t o = static_cast<size_t>(b) * p.ld_u + (c - p.u_begin);
          *reinterpret_cast<uint2*>(p.u_hi + o) = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
          if (p.u_lo) *reinterpret_cast<uint2*>(p.u_lo + o) = make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
        }
      }
      if (lane == 0 && p.logdet) {
        const float cur = p.logdet_init ? 0.f : p.logdet[b];
        if (p.block_logdet) p.block_logdet[b] = cur;
        if (p.block_logdet2) p.block_logdet2[b] = cur;
        p.logdet[b] = cur + an_ld;
      }
      __syncwarp();                                           // rowbuf is rewritten by the next row
    }
  } else
  for (long long idx = tid0; idx < total; idx += stride) {
    const int b = static_cast<int>(idx / p.C);
    const int c = static_cast<int>(idx - static_cast<long long>(b) * p.C);
    const int q = p.src ? __ldg(p.src + c) : c;
    float v = __ldg(p.in + static_cast<size_t>(b) * p.ld_in + q);
    if (p.pre_scale) {
      if (p.pre_leaky) v = v / (v >= 0.f ? 1.0f : p.alpha);
      v = v / __ldg(p.pre_scale + q) - __ldg(p.pre_loc + q);
    }
    if (p.raw_out) p.raw_out[static_cast<size_t>(b) * p.ld_raw + c] = v;
    if (p.raw_out2) p.raw_out2[static_cast<size_t>(b) * p.ld_raw + c] = v;
    if (p.post_scale) {
      v = __ldg(p.post_scale + c) * (v + __ldg(p.post_loc + c));
      if (p.post_leaky) v = v * (v >= 0.f ? 1.0f : p.alpha);
    }
    if (p.out) p.out[static_cast<size_t>(b) * p.ld_out + c] = v;
    if (p.u_hi && c >= p.u_begin && c < p.u_begin + p.u_count) {
      __nv_bfloat16 h, l;
      split_bf16(v, h, l);
      const size_t o = static_cast<size_t>(b) * p.ld_u + (c - p.u_begin);
      p.u_hi[o] = h;
      if (p.u_lo) p.u_lo[o] = l;
    }
    if (c == 0 && p.logdet) {
      const float cur = p.logdet_init ? 0.f : p.logdet[b];
      if (p.block_logdet) p.block_logdet[b] = cur;
      if (p.block_logdet2) p.block_logdet2[b] = cur;
      p.logdet[b] = p.an_logdet ? cur + __ldg(p.an_logdet) : cur;
    }
  }
  // zero the K padding of the x part of the MLP input buffer
  if (p.u_hi && p.u_pad > p.u_count) {
    const int w = p.u_pad - p.u_count;
    const long long tot = static_cast<long long>(p.B) * w;
    for (long long idx = tid0; idx < tot; idx += stride) {
      const int b = static_cast<int>(idx / w);
      const size_t o = static_cast<size_t>(b) * p.ld_u + p.u_count + static_cast<int>(idx - static_cast<long long>(b) * w);
      p.u_hi[o] = __float2bfloat16_rn(0.f);
      if (p.u_lo) p.u_lo[o] = __float2bfloat16_rn(0.f);
      if (p.u2_hi) {
        p.u2_hi[o] = __float2bfloat16_rn(0.f);
        if (p.u2_lo) p.u2_lo[o] = __float2bfloat16_rn(0.f);
      }
    }
  }
  // conditioning columns (written once per pass)
  if (p.cond && p.u_hi) {
    const long long tot = static_cast<long long>(p.B) * p.e_pad;
    for (long long idx = tid0; idx < tot; idx += stride) {
      const int b = static_cast<int>(idx / p.e_pad);
      const int e = static_cast<int>(idx - static_cast<long long>(b) * p.e_pad);
      const int cb = p.cond_rows > 0 ? b % p.cond_rows : b;
      const float v = e < p.E ? __ldg(p.cond + static_cast<size_t>(cb) * p.ld_cond + e) : 0.f;
      __nv_bfloat16 h, l;
      split_bf16(v, h, l);
      const size_t o = static_cast<size_t>(b) * p.ld_u + p.u_cond_off + e;
      p.u_hi[o] = h;
      if (p.u_lo) p.u_lo[o] = l;
      if (p.u2_hi) {
        p.u2_hi[o] = h;
        if (p.u2_lo) p.u2_lo[o] = l;
      }
    }
  }
}

// ----------------------------------------------------------------------------- backward (elementwise parts)
// Gradients are those of sum(dz*z) + sum(dlogdet*logdet); the reference has no backward code of its own
// (torch.autograd over flow_blocks.py:125-143, 255-267 and modules.py:292-301).

// Start of a backward call: sum_b dlogdet[b] (ActNorm's log|scale| term, modules.py:298-301), zero the
// conditioning-gradient accumulator and the ActNorm partial-sum scratch.
__global__ void __launch_bounds__(256) inn_bwd_init_kernel(const float* __restrict__ dlogdet, int B, float* dld_sum,
                                                           float* zero0, long long n0, float* zero1, long long n1) {
  grid_dep_launch();
  grid_dep_wait();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long tid0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (zero0) for (long long i = tid0; i < n0; i += stride) zero0[i] = 0.f;
  if (zero1) for (long long i = tid0; i < n1; i += stride) zero1[i] = 0.f;
  if (blockIdx.x == 0) {
    __shared__ double red[256];
    double a = 0.0;
    for (int b = threadIdx.x; b < B; b += blockDim.x) a += static_cast<double>(dlogdet[b]);
    red[threadIdx.x] = a;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) *dld_sum = static_cast<float>(red[0]);
  }
}

// Backward of one affine half-step  y = keep * exp(s) + t,  logdet += sum_j s_j,  s = tanh(s_pre)
// (flow_blocks.py:138-142):  dt = g,  ds_pre = (g * keep * exp(s) + dlogdet[b]) * (1 - s^2),  dkeep = g * exp(s).
// Writes the bf16 (hi/lo) operand [ds_pre | dt] of the tail dgrad / wgrad GEMMs and dkeep in place.
struct TailBwdParams {
  float* dstate; int ld_state; int keep_off;      // gradient of the coupling state [B, C]; kept half at keep_off
  const float* s_save; const float* keep_save; int ld_save;   // [B, C/2] from the training forward
  const float* dlogdet;                            // [B]
  __nv_bfloat16* dt_hi; __nv_bfloat16* dt_lo; int ld_dt; int npk;   // [B, 2*npk]: cols [0,npk) = ds_pre, [npk, 2npk) = dt
  int B, C2;
};

__global__ void __launch_bounds__(256) inn_tail_bwd_kernel(const TailBwdParams p) {
  grid_dep_launch();
  grid_dep_wait();
  const long long total = static_cast<long long>(p.B) * p.npk;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int b = static_cast<int>(idx / p.npk);
    const int c = static_cast<int>(idx - static_cast<long long>(b) * p.npk);
    float ds_pre = 0.f, dt = 0.f;
    if (c < p.C2) {
      float* gp = p.dstate + static_cast<size_t>(b) * p.ld_state + p.keep_off + c;
      const float g = *gp;
      const float s = p.s_save[static_cast<size_t>(b) * p.ld_save + c];
      const float keep = p.keep_save[static_cast<size_t>(b) * p.ld_save + c];
      const float es = expf(s);
      dt = g;
      ds_pre = (g * keep * es + __ldg(p.dlogdet + b)) * (1.f - s * s);
      *gp = g * es;
    }
    __nv_bfloat16 h, l;
    const size_t o = static_cast<size_t>(b) * p.ld_dt + c;
    split_bf16(ds_pre, h, l);
    p.dt_hi[o] = h;
    if (p.dt_lo) p.dt_lo[o] = l;
    split_bf16(dt, h, l);
    p.dt_hi[o + p.npk] = h;
    if (p.dt_lo) p.dt_lo[o + p.npk] = l;
  }
}

// Backward of the block boundary: InvLeakyRelu (flow_blocks.py:380-381), ActNorm (modules.py:292-301) and the
// previous block's Shuffle (flow_blocks.py:322-324) in one pass:
//   a = scale * (h0 + loc);  da = g * (a >= 0 ? 1 : alpha);  dh0 = da * scale
//   dloc_c = sum_b da * scale_c;  dscale_c = sum_b da * (h0 + loc_c) + (sum_b dlogdet[b]) / scale_c
//   dst[b, map[c]] = dh0   (map = previous block's forward gather map; identity for the first block)
// Channel sums: per-CTA partials -> atomics into scratch -> the last CTA to finish writes the final values.
struct BoundaryBwdParams {
  const float* g; int ld_g;              // gradient w.r.t. the block's post-activation state [B, C]
  const float* h0; int ld_h0;            // the block's input (before ActNorm) saved by the forward pass
  const float* loc; const float* scale;  // packed copies [C]
  int leaky; float alpha;
  float* dst; int ld_dst;                // nullable
  const int* map;                        // nullable = identity
  float* acc;                            // scratch [2*C] zeroed at the start of the backward call
  unsigned int* counter;                 // zeroed likewise
  const float* dld_sum;
  float* gloc; float* gscale;            // final [C] each
  int B, C;
};

__global__ void __launch_bounds__(256) inn_boundary_bwd_kernel(const BoundaryBwdParams p) {
  grid_dep_launch();
  grid_dep_wait();
  __shared__ float red[2][8][33];
  __shared__ bool is_last;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int rows_per = (p.B + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(p.B, r0 + rows_per);
  float a_loc = 0.f, a_scale = 0.f;
  if (c < p.C) {
    const float sc = __ldg(p.scale + c), lc = __ldg(p.loc + c);
    const int q = p.map ? __ldg(p.map + c) : c;
    for (int r = r0 + ty; r < r1; r += 8) {
      const float xin = __ldg(p.h0 + static_cast<size_t>(r) * p.ld_h0 + c) + lc;
      float da = __ldg(p.g + static_cast<size_t>(r) * p.ld_g + c);
      if (p.leaky && sc * xin < 0.f) da *= p.alpha;
      a_scale += da * xin;
      const float dh = da * sc;
      a_loc += dh;
      if (p.dst) p.dst[static_cast<size_t>(r) * p.ld_dst + q] = dh;
    }
  }
  red[0][ty][tx] = a_loc;
  red[1][ty][tx] = a_scale;
  __syncthreads();
  if (ty == 0 && c < p.C) {
    float sl = 0.f, ss = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { sl += red[0][k][tx]; ss += red[1][k][tx]; }
    atomicAdd(p.acc + c, sl);
    atomicAdd(p.acc + p.C + c, ss);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(p.counter, 1u);
    is_last = done == gridDim.x * gridDim.y - 1;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    const float dls = *p.dld_sum;
    for (int i = threadIdx.x; i < p.C; i += blockDim.x) {
      p.gloc[i] = __ldcg(p.acc + i);
      p.gscale[i] = __ldcg(p.acc + p.C + i) + dls / __ldg(p.scale + i);
    }
  }
}

// ----------------------------------------------------------------------------- packing
struct PackJob {
  const float* w;      // [N, K] row-major (nn.Linear.weight)
  const float* b;      // [N]
  int N, K;
  int ld_dst;          // elements per packed row
  long long dst_hi;    // element offsets into the packed buffer viewed as bf16
  long long dst_lo;    // < 0: not stored
  int row0;            // first destination row
  int col_split;       // source columns >= col_split are shifted by col_shift
  int col_shift;
  long long dstT_hi;   // transposed copy [K(+shift) rows, ld_dstT] for dgrad; < 0: none
  long long dstT_lo;
  int ld_dstT;
  long long bias_off;  // float offset into the packed buffer viewed as fp32
};

// bias_only = 1: refresh the fp32 biases only (packed-master training: the weights are written by the optimizer)
__global__ void __launch_bounds__(256) inn_pack_linear_kernel(const PackJob* __restrict__ jobs, void* packed, int bias_only) {
  const PackJob j = jobs[blockIdx.y];
  __nv_bfloat16* pk = reinterpret_cast<__nv_bfloat16*>(packed);
  float* pf = reinterpret_cast<float*>(packed);
  const long long total = bias_only ? 0 : static_cast<long long>(j.N) * j.K;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(idx / j.K);
    const int k = static_cast<int>(idx - static_cast<long long>(n) * j.K);
    const int kd = k < j.col_split ? k : k + j.col_shift;
    __nv_bfloat16 h, l;
    split_bf16(__ldg(j.w + idx), h, l);
    const long long o = static_cast<long long>(j.row0 + n) * j.ld_dst + kd;
    pk[j.dst_hi + o] = h;
    if (j.dst_lo >= 0) pk[j.dst_lo + o] = l;
    if (j.dstT_hi >= 0) {
      const long long ot = static_cast<long long>(kd) * j.ld_dstT + j.row0 + n;
      pk[j.dstT_hi + ot] = h;
      if (j.dstT_lo >= 0) pk[j.dstT_lo + ot] = l;
    }
  }
  if (blockIdx.x == 0)
    for (int n = threadIdx.x; n < j.N; n += blockDim.x) pf[j.bias_off + j.row0 + n] = __ldg(j.b + n);
}

struct PackBlockJob {
  const float* loc; const float* scale;       // [C]
  const long long* perm_fwd; const long long* perm_bwd;   // int64 [C]
  long long loc_off, scale_off, an_logdet_off;            // float offsets
  long long map_fwd_off, map_rev_off, map_fwd_inv_off, map_rev_inv_off;    // int32 offsets
};

// One CTA per flow block: copy loc/scale, sum log|scale| (modules.py:298-299), build the int32 gather maps
//   map_fwd[p] = (perm_fwd[p] + C/2) % C     (coupling output is stored [xa' | xb'], logical order is [xb' | xa'])
//   map_rev[p] = perm_bwd[(p + C/2) % C]
//   map_fwd_inv[map_fwd[p]] = p            (gradient of the gather: dS[b, q] = g[b, map_fwd_inv[q]]; folded forward scatter)
//   map_rev_inv[map_rev[p]] = p            (folded reverse scatter)
__global__ void __launch_bounds__(256) inn_pack_block_kernel(const PackBlockJob* __restrict__ jobs, void* packed, int C) {
  const PackBlockJob j = jobs[blockIdx.x];
  float* pf = reinterpret_cast<float*>(packed);
  int* pi = reinterpret_cast<int*>(packed);
  __shared__ float red[256];
  float acc = 0.f;
  const int half = C / 2;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float s = __ldg(j.scale + c);
    pf[j.loc_off + c] = __ldg(j.loc + c);
    pf[j.scale_off + c] = s;
    acc += logf(fabsf(s));
    pi[j.map_fwd_off + c] = static_cast<int>((j.perm_fwd[c] + half) % C);
    pi[j.map_rev_off + c] = static_cast<int>(j.perm_bwd[(c + half) % C]);
    pi[j.map_fwd_inv_off + static_cast<int>((j.perm_fwd[c] + half) % C)] = c;   // inverse of map_fwd (backward un-shuffle)
    pi[j.map_rev_inv_off + static_cast<int>(j.perm_bwd[(c + half) % C])] = c;     // inverse of map_rev (folded reverse scatter)
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) pf[j.an_logdet_off] = red[0];
}

// ----------------------------------------------------------------------------- ActNorm init
__global__ void __launch_bounds__(256) inn_actnorm_stats_kernel(const float* __restrict__ x, int B, int C,
                                                                double* __restrict__ moments) {
  // each thread owns one channel of a row slab: coalesced across channels
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int rows_per = (B + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(B, r0 + rows_per);
  double s = 0.0, ss = 0.0;
  for (int r = r0; r < r1; ++r) {
    const double v = static_cast<double>(__ldg(x + static_cast<size_t>(r) * C + c));
    s += v;
    ss += v * v;
  }
  if (r1 > r0) {
    atomicAdd(moments + c, s);
    atomicAdd(moments + C + c, ss);
  }
}

__global__ void inn_actnorm_finalize_kernel(const double* __restrict__ moments, double count, int C,
                                            float* __restrict__ loc, float* __restrict__ scale) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double mean = moments[c] / count;
  double var = (moments[C + c] - count * mean * mean) / (count - 1.0);     // unbiased, torch.std default
  var = var > 0.0 ? var : 0.0;
  loc[c] = static_cast<float>(-mean);
  scale[c] = static_cast<float>(1.0 / (sqrt(var) + 1e-6));
}

// ----------------------------------------------------------------------------- FlowLoss
// single CTA: B*C is small next to the GEMMs; deterministic tree reduction
// `ref` (nullable): the N(0, I) sample the reference draws for its logged reference NLL (loss.py:27, 32); its mean NLL goes
// to out3[3] in the same launch.  Up to FLOW_LOSS_MAX_BLOCKS CTAs: every CTA reduces its slice with a fixed-order tree and
// parks three fp64 partials in `scratch`; the CTA that takes the last ticket adds them in block order (deterministic),
// writes the results and resets the ticket counter for the next call.  scratch = FLOW_LOSS_SCRATCH_BYTES, zeroed once by
// the caller; NULL = single CTA (small batches).
constexpr int FLOW_LOSS_MAX_BLOCKS = 32;
constexpr int FLOW_LOSS_SCRATCH_BYTES = 8 * (3 * FLOW_LOSS_MAX_BLOCKS + 1);
__global__ void __launch_bounds__(1024) inn_flow_loss_kernel(const float* __restrict__ z, const float* __restrict__ logdet,
                                                             int B, int C, float nll_w, float ld_w, float* out3,
                                                             float* dz, float* dlogdet, const float* __restrict__ ref,
                                                             double* scratch) {
  __shared__ double r_a[1024];
  __shared__ double r_b[1024];
  __shared__ double r_c[1024];
  __shared__ unsigned int s_ticket;
  double a = 0.0, l = 0.0, r = 0.0;
  const long long total = static_cast<long long>(B) * C;
  const float inv_b = 1.0f / static_cast<float>(B);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const float v = z[i];
    a += 0.5 * static_cast<double>(v) * v;
    if (dz) dz[i] = nll_w * v * inv_b;
    if (ref) {
      const float w = ref[i];
      r += 0.5 * static_cast<double>(w) * w;
    }
  }
  for (long long b = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; b < B; b += stride) {
    l += logdet[b];
    if (dlogdet) dlogdet[b] = -ld_w * inv_b;
  }
  r_a[threadIdx.x] = a;
  r_b[threadIdx.x] = l;
  r_c[threadIdx.x] = r;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      r_a[threadIdx.x] += r_a[threadIdx.x + s];
      r_b[threadIdx.x] += r_b[threadIdx.x + s];
      r_c[threadIdx.x] += r_c[threadIdx.x + s];
    }
    __syncthreads();
  }
  double sa = r_a[0], sl = r_b[0], sr = r_c[0];
  if (gridDim.x > 1) {
    if (threadIdx.x == 0) {
      scratch[3 * blockIdx.x] = sa;
      scratch[3 * blockIdx.x + 1] = sl;
      scratch[3 * blockIdx.x + 2] = sr;
      __threadfence();
      s_ticket = atomicAdd(reinterpret_cast<unsigned int*>(scratch + 3 * FLOW_LOSS_MAX_BLOCKS), 1u);
    }
    __syncthreads();
    if (s_ticket != gridDim.x - 1) return;
    if (threadIdx.x == 0) {
      __threadfence();
      sa = sl = sr = 0.0;
      for (unsigned int k = 0; k < gridDim.x; ++k) {
        sa += scratch[3 * k];
        sl += scratch[3 * k + 1];
        sr += scratch[3 * k + 2];
      }
      *reinterpret_cast<unsigned int*>(scratch + 3 * FLOW_LOSS_MAX_BLOCKS) = 0u;      // ready for the next call
    }
  }
  if (threadIdx.x == 0) {
    const double nll = sa / B, nld = -sl / B;
    out3[0] = static_cast<float>(nll_w * nll + ld_w * nld);
    out3[1] = static_cast<float>(nll);
    out3[2] = static_cast<float>(nld);
    if (ref) out3[3] = static_cast<float>(sr / B);
  }
}

}  // namespace inn
