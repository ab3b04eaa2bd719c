// hidden_cluster_kernel.cuh -- hidden Linear + bias + LeakyReLU (modules.py:171-180) for SMALL batches (one wave of
// tiles) as a split-K thread-block cluster.
//
// Why: at batch <= ~512 a hidden layer is a chain of dependent one-wave launches, each bound by how fast ONE SM can pull
// its operands through its L2 port (~100-128 GB/s: r01 device trace, 3.1 us for the 320 KB a 128 x 32 tile needs at
// K = 1024).  Here KS CTAs of a cluster share one 128 x BN tile: CTA kh accumulates the kh-th K slice in TMEM
// ((128 + BN) * K / KS * 2 bytes through its port), then the partial accumulators are reduce-scattered by rows through
// distributed shared memory (CTA j receives rows [j * R, (j + 1) * R), R = 128 / KS, of every partial) and each CTA
// finishes R x BN outputs: sum, bias, LeakyReLU, bf16 (hi / lo) store.
// Same structure as the small-batch coupling tail (tail_cluster_kernel.cuh).
#pragma once
#include "tail_cluster_kernel.cuh"

namespace inn {

constexpr int HC_THREADS = 384;            // warps 0-3: TMA / MMA / TMEM alloc / idle; warps 4-11: epilogue
constexpr int HC_EPI_THREADS = 256;

__host__ inline int hidden_cluster_stage_bytes(int ksplit, int BN) { return ksplit * (A_TILE_BYTES + BN * BK * 2); }
__host__ inline int hidden_cluster_recv_bytes(int BN) { return BM * BN * 4; }   // [KS][128 / KS][BN] fp32
__host__ inline int hidden_cluster_smem_bytes(int ksplit, int BN, int stages) {
  return 1024 + stages * hidden_cluster_stage_bytes(ksplit, BN) + hidden_cluster_recv_bytes(BN) + SMEM_BARRIER_BYTES;
}

// GemmParams fields used: M, BN, num_n_tiles, tiles_per_group, a_group_stride, num_k_blocks (whole K), ks, stages,
// w_prefetch, bias, slope, out_hi, out_lo, ld_out.  Grid = m_tiles * num_n_tiles * ks CTAs, cluster (ks, 1, 1).
template <int kSplit>
__global__ void __launch_bounds__(HC_THREADS, 1)
inn_hidden_cluster_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                          const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                          const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  cluster_arrive_release();                  // #1: "my CTA runs" -- waited on just before the first remote store

  const int KS = p.ks;
  const int BN = p.BN;
  const int kh = static_cast<int>(cluster_ctarank());       // K slice of this CTA
  const int tile = blockIdx.x / KS;
  const int m_tile = tile / p.num_n_tiles;
  const int n_tile = tile - m_tile * p.num_n_tiles;
  const int group = n_tile / p.tiles_per_group;
  const int nkb = p.num_k_blocks / KS;
  const int kb0 = kh * nkb;
  const int stages = p.stages;
  const int w_tile_bytes = BN * BK * 2;
  const int stage_bytes = kSplit * (A_TILE_BYTES + w_tile_bytes);

  float* recv = reinterpret_cast<float*>(smem + stages * stage_bytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(recv) + BM * BN * 4);
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
  int tcols = 32;
  while (tcols < BN) tcols <<= 1;
  if (warp == 2) tmem_alloc(tmem_slot, tcols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) grid_dep_launch();

  if (warp == 0) {
    // ============================ TMA producer ============================
    const int w_row = n_tile * BN;
    const int pre = p.w_prefetch ? (nkb < stages ? nkb : stages) : 0;
    for (int ki = 0; ki < pre; ++ki) {       // weights do not depend on the previous kernel
      mbar_arrive_expect_tx_el(&full_bar[ki], stage_bytes);
      uint8_t* sbase = smem + ki * stage_bytes;
#pragma unroll
      for (int sp = 0; sp < kSplit; ++sp)
        tma_load_2d_el(sbase + kSplit * A_TILE_BYTES + sp * w_tile_bytes, sp == 0 ? &tmW0 : &tmW1, &full_bar[ki],
                       (kb0 + ki) * BK, w_row);
    }
    const int a_col0 = group * p.a_group_stride + kb0 * BK;
    grid_dep_wait();
    int stage = 0;
    uint32_t phase = 0;
    for (int ki = 0; ki < nkb; ++ki) {
      uint8_t* sbase = smem + stage * stage_bytes;
      const bool w_done = ki < pre;
      if (!w_done) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 31);
        mbar_arrive_expect_tx_el(&full_bar[stage], stage_bytes);
      }
#pragma unroll
      for (int sp = 0; sp < kSplit; ++sp) {
        tma_load_2d_el(sbase + sp * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, &full_bar[stage], a_col0 + ki * BK, m_tile * BM);
        if (!w_done)
          tma_load_2d_el(sbase + kSplit * A_TILE_BYTES + sp * w_tile_bytes, sp == 0 ? &tmW0 : &tmW1, &full_bar[stage],
                         (kb0 + ki) * BK, w_row);
      }
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ==============================
    const uint32_t idesc = umma_idesc_bf16(BM, BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    for (int ki = 0; ki < nkb; ++ki) {
      mbar_wait(&full_bar[stage], phase, 33);
      tc_fence_after();
      const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
      const uint64_t da_hi = umma_desc_k_sw128(sbase);
      const uint64_t da_lo = umma_desc_k_sw128(sbase + (kSplit - 1) * A_TILE_BYTES);
      const uint64_t dw_hi = umma_desc_k_sw128(sbase + kSplit * A_TILE_BYTES);
      const uint64_t dw_lo = umma_desc_k_sw128(sbase + kSplit * A_TILE_BYTES + (kSplit - 1) * w_tile_bytes);
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
      if (ki == nkb - 1) umma_commit_el(tfull_bar);
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  }

  const int R = BM / KS;                                    // rows this CTA finishes: 64, 32 or 16
  if (warp >= 4) {
    // ---- my partial accumulator: TMEM lane = tile row; the row goes to the CTA that owns it
    const int q = warp & 3;
    const int ch = (warp - 4) >> 2;                         // column half of this warp
    const int r_loc = q * 32 + lane;
    const int dst_rank = r_loc / R;
    const int half = BN >> 1;
    mbar_wait(tfull_bar, 0, 34);
    tc_fence_after();
    cluster_wait_acquire();                                 // #1: every CTA of the cluster is running
    const uint32_t ra = dsmem_addr(recv + (kh * R + (r_loc - dst_rank * R)) * BN + ch * half, static_cast<uint32_t>(dst_rank));
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + ch * half;
    for (int c0 = 0; c0 < half; c0 += 32) {
      float v[32];
      tmem_ld16(t_row + c0, v);
      tmem_ld16(t_row + c0 + 16, v + 16);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j) dsmem_st4(ra + (c0 + 4 * j) * 4, v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    tc_fence_before();
  } else {
    cluster_wait_acquire();                                 // #1
  }
  cluster_arrive_release();                                 // #2: partials delivered
  cluster_wait_acquire();

  if (warp >= 4) {
    const int tid_e = threadIdx.x - 128;
    const int upr = BN >> 4;                                // 16-column units per row
    const int n_units = R * upr;
    for (int u = tid_e; u < n_units; u += HC_EPI_THREADS) {
      const int rr = u / upr;
      const int cu = (u - rr * upr) << 4;
      const int row = m_tile * BM + kh * R + rr;
      float acc[16];
      const float4* b4 = reinterpret_cast<const float4*>(p.bias + n_tile * BN + cu);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 b = __ldg(b4 + j);
        acc[4 * j] = b.x; acc[4 * j + 1] = b.y; acc[4 * j + 2] = b.z; acc[4 * j + 3] = b.w;
      }
      for (int h = 0; h < KS; ++h) {
        const float4* src = reinterpret_cast<const float4*>(recv + (h * R + rr) * BN + cu);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 a = src[j];
          acc[4 * j] += a.x; acc[4 * j + 1] += a.y; acc[4 * j + 2] += a.z; acc[4 * j + 3] += a.w;
        }
      }
      if (row >= p.M) continue;
      uint32_t hi[8], lo[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a = acc[2 * j], b = acc[2 * j + 1];
        a = fmaxf(a, a * p.slope);                          // LeakyReLU, 0 <= slope <= 1
        b = fmaxf(b, b * p.slope);
        if (kSplit == 1) {
          hi[j] = cvt_bf16x2(a, b);
        } else {
          __nv_bfloat16 h0, l0, h1, l1;
          split_bf16(a, h0, l0);
          split_bf16(b, h1, l1);
          hi[j] = pack_bf16x2(h0, h1);
          lo[j] = pack_bf16x2(l0, l1);
        }
      }
      const size_t off = static_cast<size_t>(row) * p.ld_out + n_tile * BN + cu;
      uint4* o = reinterpret_cast<uint4*>(p.out_hi + off);
      o[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      o[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
      if (kSplit == 2) {
        uint4* ol = reinterpret_cast<uint4*>(p.out_lo + off);
        ol[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        ol[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tcols);
  }
}

}  // namespace inn
