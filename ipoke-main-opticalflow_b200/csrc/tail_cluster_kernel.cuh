// Coupling tail for small batches (one wave of tiles): the last Linear of s_i and t_i as a thread-block cluster.
//
// The persistent tail of gemm_kernel.cuh gives one CTA both accumulators of a 128 x BN tile over the whole K: at
// batch <= ~512 that is a few dozen CTAs, each pulling 2 x 128 x H activations through one SM's L2 port (~128 GB/s)
// and then issuing the folded scatter (one 32-byte sector per element) from one SM.  Here a cluster of
// cs = 2 * KS CTAs shares a 128 x 16 tile: CTA (net, kh) accumulates net in {s, t} over the kh-th K slice in TMEM,
// then the partial accumulators are reduce-scattered by rows through distributed shared memory -- CTA j receives
// the 128 / cs rows [j * R, (j + 1) * R) of every partial -- and each CTA runs the coupling epilogue
// (bias, tanh, exp, affine update, log-det, Shuffle/ActNorm/InvLeakyRelu fold: flow_blocks.py:138-142, 151, 255-274)
// on its R x 16 elements.  Operand traffic per CTA drops by cs and the scattered stores spread over cs times more SMs.
#pragma once
#include "gemm_kernel.cuh"

namespace inn {

constexpr int TC_BN = 16;                 // columns of one tile (UMMA N)
constexpr int TC_THREADS = 256;           // warps 0-3: TMA / MMA / TMEM alloc / idle; warps 4-7: epilogue
constexpr int TC_EPI_THREADS = 128;
constexpr int TC_W_TILE_BYTES = TC_BN * BK * 2;
constexpr int TC_RECV_BYTES = BM * TC_BN * 4;       // [cs][128 / cs][16] fp32 partials addressed to this CTA

__host__ inline int tail_cluster_stage_bytes(int ksplit) { return ksplit * (A_TILE_BYTES + TC_W_TILE_BYTES); }
__host__ inline int tail_cluster_smem_bytes(int ksplit, int stages) {
  return 1024 + stages * tail_cluster_stage_bytes(ksplit) + TC_RECV_BYTES + SMEM_BARRIER_BYTES;
}

__device__ __forceinline__ void cluster_arrive_release() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait_acquire() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `local_smem` at the same offset in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t dsmem_addr(const void* local_smem, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(local_smem)), "r"(rank));
  return r;
}
__device__ __forceinline__ void dsmem_st4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// GemmParams fields used: M, num_n_tiles, num_k_blocks (whole K), ks, stages, w_prefetch, a_off[2], w_row0[2], bias, bias2 and
// every EPI_TAIL_* field (state .. ld_extra).  Grid = m_tiles * num_n_tiles * cs CTAs, cluster (cs, 1, 1).
template <int kSplit, bool kRev>
__global__ void __launch_bounds__(TC_THREADS, 1)
inn_tail_cluster_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                        const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                        const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  unsigned long long* const tr =
      (p.trace != nullptr && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? p.trace + (blockIdx.x == 0 ? 0 : 16) : nullptr;
  if (tr != nullptr && threadIdx.x == 0) tr[0] = globaltimer_ns();
  cluster_arrive_release();                  // #1: "my CTA runs" -- waited on just before the first remote store

  const int KS = p.ks;
  const int cs = 2 * KS;
  const int rank = static_cast<int>(cluster_ctarank());
  const int net = rank / KS;                 // 0: s subnetwork, 1: t subnetwork
  const int kh = rank - net * KS;            // K slice
  const int tile = blockIdx.x / cs;
  const int m_tile = tile / p.num_n_tiles;
  const int n_tile = tile - m_tile * p.num_n_tiles;
  int a_row0 = m_tile * BM;                  // computed in the prologue, off the dependent-launch critical path
  if (p.a_m_wrap > 0) a_row0 = (m_tile % p.a_m_wrap) * BM;
  const int nkb = p.num_k_blocks / KS;
  const int kb0 = kh * nkb;
  const int stages = p.stages;
  constexpr int stage_bytes = kSplit * (A_TILE_BYTES + TC_W_TILE_BYTES);

  float* recv = reinterpret_cast<float*>(smem + stages * stage_bytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(recv) + TC_RECV_BYTES);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    tma_prefetch_desc(&tmW0);
    if (kSplit == 2) {
      tma_prefetch_desc(&tmA1);
      tma_prefetch_desc(&tmW1);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tr != nullptr && threadIdx.x == 0) tr[1] = globaltimer_ns();
  if (threadIdx.x == 0) grid_dep_launch();

  if (warp == 0) {
    // ============================ TMA producer ============================
    const int w_row = p.w_row0[net] + n_tile * TC_BN;
    const int pre = p.w_prefetch ? (nkb < stages ? nkb : stages) : 0;
    for (int ki = 0; ki < pre; ++ki) {       // weights do not depend on the previous kernel
      mbar_arrive_expect_tx_el(&full_bar[ki], stage_bytes);
      uint8_t* sbase = smem + ki * stage_bytes;
#pragma unroll
      for (int sp = 0; sp < kSplit; ++sp)
        tma_load_2d_el(sbase + kSplit * A_TILE_BYTES + sp * TC_W_TILE_BYTES, sp == 0 ? &tmW0 : &tmW1, &full_bar[ki],
                       (kb0 + ki) * BK, w_row);
    }
    grid_dep_wait();
    if (tr != nullptr && lane == 0) tr[2] = globaltimer_ns();
    int stage = 0;
    uint32_t phase = 0;
    for (int ki = 0; ki < nkb; ++ki) {
      uint8_t* sbase = smem + stage * stage_bytes;
      const bool w_done = ki < pre;
      if (!w_done) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 11);
        mbar_arrive_expect_tx_el(&full_bar[stage], stage_bytes);
      }
#pragma unroll
      for (int sp = 0; sp < kSplit; ++sp) {
        tma_load_2d_el(sbase + sp * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, &full_bar[stage],
                       p.a_off[net] + (kb0 + ki) * BK, a_row0);
        if (!w_done)
          tma_load_2d_el(sbase + kSplit * A_TILE_BYTES + sp * TC_W_TILE_BYTES, sp == 0 ? &tmW0 : &tmW1,
                         &full_bar[stage], (kb0 + ki) * BK, w_row);
      }
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ==============================
    const uint32_t idesc = umma_idesc_bf16(BM, TC_BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    for (int ki = 0; ki < nkb; ++ki) {
      mbar_wait(&full_bar[stage], phase, 13);
      tc_fence_after();
      if (tr != nullptr && lane == 0 && ki == 0) tr[3] = globaltimer_ns();
      const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
      const uint64_t da_hi = umma_desc_k_sw128(sbase);
      const uint64_t da_lo = umma_desc_k_sw128(sbase + (kSplit - 1) * A_TILE_BYTES);
      const uint64_t dw_hi = umma_desc_k_sw128(sbase + kSplit * A_TILE_BYTES);
      const uint64_t dw_lo = umma_desc_k_sw128(sbase + kSplit * A_TILE_BYTES + (kSplit - 1) * TC_W_TILE_BYTES);
#pragma unroll
      for (int kk = 0; kk < BK / UMMA_K; ++kk) {
        const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 2) >> 4);
        const uint32_t accf = (ki > 0 || kk > 0) ? 1u : 0u;
        if (kSplit == 2) {
          umma_bf16_ss_el(tmem_base, da_lo + koff, dw_hi + koff, idesc, accf);
          umma_bf16_ss_el(tmem_base, da_hi + koff, dw_lo + koff, idesc, 1u);
          umma_bf16_ss_el(tmem_base, da_hi + koff, dw_hi + koff, idesc, 1u);
        } else {
          umma_bf16_ss_el(tmem_base, da_hi + koff, dw_hi + koff, idesc, accf);
        }
      }
      umma_commit_el(&empty_bar[stage]);
      if (ki == nkb - 1) {
        umma_commit_el(tfull_bar);
        if (tr != nullptr && lane == 0) tr[4] = globaltimer_ns();
      }
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  }

  // ---- rows / columns this CTA finishes after the reduce-scatter
  const int R = BM / cs;                                   // rows per CTA: 64, 32 or 16
  const int tid_e = threadIdx.x - 128;                     // epilogue thread (warps 4-7), else negative
  const int c4 = (tid_e & 3) << 2;                         // 4 consecutive columns of the tile
  const int gcol = n_tile * TC_BN + c4;                    // column inside the kept half
  const int nv = p.n_valid - gcol < 4 ? p.n_valid - gcol : 4;   // <= 0: padded columns
  float kp[2][4];                                          // kept values of up to 2 rows per thread (R = 64)
  float bs[4], bt[4], csc[4], clo[4];
  int cdst[4];
  if (warp >= 4) {
    // ---- per-column constants (parameters: independent of the previous kernel)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      bs[e] = __ldg(p.bias + gcol + e);
      bt[e] = __ldg(p.bias2 + gcol + e);
      csc[e] = 1.f;
      clo[e] = 0.f;
      cdst[e] = p.keep_off + gcol + e;
      if (p.f_out != nullptr && e < nv) {
        const int src = p.keep_off + gcol + e;
        const int dst = p.f_dst != nullptr ? __ldg(p.f_dst + src) : src;
        cdst[e] = dst;
        if (p.f_mode != 0) {
          const int pi = p.f_mode == 1 ? dst : src;        // forward: next block's ActNorm; reverse: this block's
          const float sc = __ldg(p.f_scale + pi);
          csc[e] = p.f_mode == 2 ? 1.0f / sc : sc;
          clo[e] = __ldg(p.f_loc + pi);
        }
      }
    }
    grid_dep_wait();
    // ---- kept half of my rows (written by the previous tail / boundary kernel)
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int rr = (tid_e >> 2) + 32 * i;
      const int row = m_tile * BM + rank * R + rr;
      kp[i][0] = kp[i][1] = kp[i][2] = kp[i][3] = 0.f;
      if (rr < R && row < p.M && nv > 0) {
        const float* krow = p.state + static_cast<size_t>(row) * p.ld_state + p.keep_off + gcol;
        if (p.vec4) {
          const float4 v = *reinterpret_cast<const float4*>(krow);
          kp[i][0] = v.x; kp[i][1] = v.y; kp[i][2] = v.z; kp[i][3] = v.w;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e < nv) kp[i][e] = krow[e];
        }
      }
    }
    // ---- my partial accumulator: TMEM lane = tile row; its 16 columns go to the CTA that owns the row
    const int q = warp & 3;
    const int r_loc = q * 32 + lane;
    mbar_wait(tfull_bar, 0, 14);
    tc_fence_after();
    if (tr != nullptr && tid_e == 0) tr[5] = globaltimer_ns();
    float v[16];
    tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16), v);
    tmem_ld_wait();
    tc_fence_before();
    cluster_wait_acquire();                                // #1: every CTA of the cluster is running
    const int dst_rank = r_loc / R;
    const uint32_t ra = dsmem_addr(recv + (rank * R + (r_loc - dst_rank * R)) * TC_BN, static_cast<uint32_t>(dst_rank));
#pragma unroll
    for (int j = 0; j < 4; ++j) dsmem_st4(ra + 16 * j, v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  } else {
    cluster_wait_acquire();                                // #1
  }
  cluster_arrive_release();                                // #2: partials delivered
  cluster_wait_acquire();
  if (tr != nullptr && threadIdx.x == 128) tr[9] = globaltimer_ns();

  if (warp >= 4) {
    const float neg = p.f_leaky ? (p.f_mode == 2 ? 1.0f / p.f_alpha : p.f_alpha) : 1.0f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int rr = (tid_e >> 2) + 32 * i;
      if (rr >= R) break;                                  // warp-uniform (R is a multiple of 8 rows per warp)
      const int row = m_tile * BM + rank * R + rr;
      float s4[4] = {0.f, 0.f, 0.f, 0.f}, t4[4] = {0.f, 0.f, 0.f, 0.f};
      for (int h = 0; h < KS; ++h) {
        const float4 a = *reinterpret_cast<const float4*>(recv + (h * R + rr) * TC_BN + c4);
        const float4 b = *reinterpret_cast<const float4*>(recv + ((KS + h) * R + rr) * TC_BN + c4);
        s4[0] += a.x; s4[1] += a.y; s4[2] += a.z; s4[3] += a.w;
        t4[0] += b.x; t4[1] += b.y; t4[2] += b.z; t4[3] += b.w;
      }
      float ld_sum = 0.f;
      float nw[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        // bf16 mode: MUFU tanh / ex2 (error below the bf16 GEMM error); fp32-parity mode: libm-accurate tanhf / expf
        const float s = kSplit == 2 ? tanhf(s4[e] + bs[e]) : tanh_approx(s4[e] + bs[e]);
        const float t = t4[e] + bt[e];
        if (!kRev) {
          nw[e] = kp[i][e] * (kSplit == 2 ? expf(s) : ex2_approx(s * 1.4426950408889634f)) + t;
          ld_sum += s;                                     // padded columns: zero weights and biases -> s = 0
        } else {
          nw[e] = (kp[i][e] - t) * (kSplit == 2 ? expf(-s) : ex2_approx(s * -1.4426950408889634f));
        }
        s4[e] = s;
      }
      if (!kRev) {                                         // row sum over this tile's 16 columns: 4 lanes per row
        ld_sum += __shfl_xor_sync(0xffffffffu, ld_sum, 1);
        ld_sum += __shfl_xor_sync(0xffffffffu, ld_sum, 2);
      }
      if (row >= p.M || nv <= 0) continue;
      if (!kRev && (tid_e & 3) == 0) {
        if (p.ld_extra != nullptr && n_tile == 0) ld_sum += __ldg(p.ld_extra);
        atomicAdd(p.logdet + row, ld_sum);
      }
      if (p.s_save != nullptr) {                           // training forward keeps s and the old kept half
        float* ss = p.s_save + static_cast<size_t>(row) * p.ld_save + gcol;
        float* ks = p.keep_save + static_cast<size_t>(row) * p.ld_save + gcol;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e < nv) { ss[e] = s4[e]; ks[e] = kp[i][e]; }
      }
      if (p.u_hi != nullptr) {                             // bf16 copy for this block's second half-step
        __nv_bfloat16 h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) split_bf16(nw[e], h[e], l[e]);
        __nv_bfloat16* uh = p.u_hi + static_cast<size_t>(row) * p.ld_u + gcol;
        if (p.vec4) {
          *reinterpret_cast<uint2*>(uh) = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
          if (kSplit == 2)
            *reinterpret_cast<uint2*>(p.u_lo + static_cast<size_t>(row) * p.ld_u + gcol) =
                make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e < nv) {
              uh[e] = h[e];
              if (kSplit == 2) p.u_lo[static_cast<size_t>(row) * p.ld_u + gcol + e] = l[e];
            }
        }
      }
      if (p.f_out == nullptr) {
        float* d = p.state + static_cast<size_t>(row) * p.ld_state + p.keep_off + gcol;
        if (p.vec4) {
          *reinterpret_cast<float4*>(d) = make_float4(nw[0], nw[1], nw[2], nw[3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e < nv) d[e] = nw[e];
        }
      } else {
        // folded block boundary: every element goes to its place in the next block's state (or the caller's output)
        float* orow = p.f_out + static_cast<size_t>(row) * p.f_ld;
        const size_t uoff = static_cast<size_t>(row) * p.ld_u;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (e < nv) {
            float o = nw[e];
            if (p.f_mode == 1) {
              o = csc[e] * (o + clo[e]);
              o = o * (o >= 0.f ? 1.0f : neg);
            } else if (p.f_mode == 2) {
              o = o * (o >= 0.f ? 1.0f : neg);
              o = o * csc[e] - clo[e];
            }
            orow[cdst[e]] = o;
            const int ucol = cdst[e] - p.f_u_begin;
            if (p.f_u_hi != nullptr && ucol >= 0 && ucol < p.n_valid) {
              __nv_bfloat16 h, l;
              split_bf16(o, h, l);
              p.f_u_hi[uoff + ucol] = h;
              if (kSplit == 2) p.f_u_lo[uoff + ucol] = l;
            }
          }
        }
      }
    }
  }

  if (tr != nullptr && threadIdx.x == 128) tr[6] = globaltimer_ns();
  tc_fence_before();
  __syncthreads();
  if (tr != nullptr && threadIdx.x == 0) tr[7] = globaltimer_ns();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 32);
  }
}

}  // namespace inn
