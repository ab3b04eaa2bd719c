// tail_pair_kernel.cuh -- the coupling tail (last Linear of s_i and t_i + tanh / exp / affine update / log-det,
// flow_blocks.py:138-142, 151) for multi-wave launches as a tcgen05 CTA-pair GEMM (cta_group::2).
//
// Why: the single-CTA tail covers a 128-row tile with N / BN column tiles and every one of them pulls both 128 x H
// activation panels (s-net and t-net) through the L2 -> SM path again: 2 MB of operands per 128 rows at cfg B
// (H = 1024, N = 256, BN = 128), 572 MB per launch at 32768 rows, and the mainloop runs at the fabric's rate, not the
// tensor pipe's (device timeline r02: 17-19 us per tile against 7 us of MMA time; ncu r01: tensor pipe 13 %).  A pair
// covers 256 rows x the FULL N: each CTA stages its own 128 rows of both panels ONCE and half of the W rows (the MMA
// reads the other half from the peer's shared memory): 1 MB per 128 rows.
//
// Unfolded programs only (large batches of wide states, and every training forward): the epilogue writes the kept half
// back in place, the bf16 copy for the block's second half-step, and what the backward pass needs.
// Roles per CTA (384 threads) as in hidden_pair_kernel.cuh: warp 0 TMA producer (both CTAs), warp 1 MMA issuer (leader),
// warp 2 TMEM allocator, warps 4-11 epilogue (each CTA on its own 128 accumulator rows, one row per thread, the two
// warp groups take alternate 16-column groups).
#pragma once
#include "hidden_pair_kernel.cuh"

namespace inn {

struct TailPairCfg {
  int num_m_pairs;        // ceil(num_m_tiles / 2)
  int BN;                 // columns of s (and of t) per tile; each CTA stages BN / 2 rows of W_s and of W_t
  int stages;
  int nacc;               // accumulator sets (s | t) in TMEM: 2 when 4 * BN <= 512
  int tmem_cols;
  int rev;                // 1: tiles from the last row tile down (see PairParams::rev)
};

__host__ inline int tail_pair_stage_bytes(int ksplit, int BN) { return ksplit * (2 * A_TILE_BYTES + 2 * (BN / 2) * BK * 2); }
__host__ inline int tail_pair_smem_bytes(int ksplit, int BN, int stages) {
  return 1024 + stages * tail_pair_stage_bytes(ksplit, BN) + SMEM_BARRIER_BYTES;
}

template <int kSplit, bool kRev>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
inn_tail_pair_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                     const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                     const GemmParams p, const TailPairCfg c) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int BN = c.BN;
  const int w_half_bytes = (BN / 2) * BK * 2;
  const int stage_bytes = kSplit * (2 * A_TILE_BYTES + 2 * w_half_bytes);
  const int stages = c.stages;
  const int nacc = c.nacc;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = c.num_m_pairs * p.num_n_tiles;

  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

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
      mbar_init(&full_bar[s], 1);                    // leader's producer arrives once (+ both CTAs' bytes)
      mbar_init(&empty_bar[s], 1);                   // one multicast commit of the leader's MMA warp
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 16);                 // 8 epilogue warps of each CTA (used in the leader only)
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, c.tmem_cols);
  tc_fence_before();
  cluster_sync_all();                                // both CTAs' barriers and TMEM exist before anyone signals
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) grid_dep_launch();           // PDL: the next kernel may start its prologue

  if (warp == 0) {
    // ============================ TMA producer (both CTAs) ============================
    grid_dep_wait();
    uint32_t leader_full[MAX_STAGES];
#pragma unroll
    for (int s = 0; s < MAX_STAGES; ++s) leader_full[s] = mapa_u32(&full_bar[s], 0);
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m_pair = (c.rev ? num_tiles - 1 - tile : tile) / p.num_n_tiles;
      const int n_tile = (c.rev ? num_tiles - 1 - tile : tile) - m_pair * p.num_n_tiles;
      const int m_tile = 2 * m_pair + static_cast<int>(rank);
      const int a_row0 = (p.a_m_wrap > 0 ? m_tile % p.a_m_wrap : m_tile) * BM;
      const int w_r = n_tile * BN + static_cast<int>(rank) * (BN / 2);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 31);
        if (rank == 0) mbar_arrive_expect_tx_el(&full_bar[stage], 2 * stage_bytes);
        uint8_t* sbase = smem + stage * stage_bytes;
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
#pragma unroll
          for (int sp = 0; sp < kSplit; ++sp) {
            const int t = sub * kSplit + sp;
            tma_load_2d_pair_el(sbase + t * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, leader_full[stage],
                                p.a_off[sub] + kb * BK, a_row0);
            tma_load_2d_pair_el(sbase + 2 * kSplit * A_TILE_BYTES + t * w_half_bytes, sp == 0 ? &tmW0 : &tmW1,
                                leader_full[stage], kb * BK, p.w_row0[sub] + w_r);
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ============================ MMA issuer (leader CTA only) ============================
    const uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 32);      // both CTAs' epilogues have drained this accumulator set
      tc_fence_after();
      const uint32_t d_base = tmem_base + acc * 2 * BN;    // [s: BN columns | t: BN columns]
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&full_bar[stage], phase, 33);
        tc_fence_after();
        const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
        const uint32_t w0 = sbase + 2 * kSplit * A_TILE_BYTES;
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          const uint64_t da_hi = umma_desc_k_sw128(sbase + (sub * kSplit) * A_TILE_BYTES);
          const uint64_t da_lo = umma_desc_k_sw128(sbase + (sub * kSplit + kSplit - 1) * A_TILE_BYTES);
          const uint64_t dw_hi = umma_desc_k_sw128(w0 + (sub * kSplit) * w_half_bytes);
          const uint64_t dw_lo = umma_desc_k_sw128(w0 + (sub * kSplit + kSplit - 1) * w_half_bytes);
          const uint32_t d = d_base + sub * BN;
#pragma unroll
          for (int kk = 0; kk < BK / UMMA_K; ++kk) {
            const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 2) >> 4);
            const uint32_t first = (kb == 0 && kk == 0) ? 0u : 1u;
            if (kSplit == 2) {                               // hi*hi + hi*lo + lo*hi (drops lo*lo, ~2^-16 relative)
              umma_bf16_ss_pair_el(d, da_lo + koff, dw_hi + koff, idesc, first);
              umma_bf16_ss_pair_el(d, da_hi + koff, dw_lo + koff, idesc, 1u);
              umma_bf16_ss_pair_el(d, da_hi + koff, dw_hi + koff, idesc, 1u);
            } else {
              umma_bf16_ss_pair_el(d, da_hi + koff, dw_hi + koff, idesc, first);
            }
          }
        }
        umma_commit_pair_el(&empty_bar[stage]);              // frees the slot in BOTH CTAs when the MMAs retire
        if (kb == p.num_k_blocks - 1) umma_commit_pair_el(&tfull_bar[acc]);
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ============================ epilogue (both CTAs) ============================
    grid_dep_wait();
    const int q = warp & 3;                          // TMEM lane quarter
    const int ch = (warp - 4) >> 2;                  // which of the alternating 16-column groups
    const uint32_t leader_tempty[2] = {mapa_u32(&tempty_bar[0], 0), mapa_u32(&tempty_bar[1], 0)};
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int m_pair = (c.rev ? num_tiles - 1 - tile : tile) / p.num_n_tiles;
      const int n_tile = (c.rev ? num_tiles - 1 - tile : tile) - m_pair * p.num_n_tiles;
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      const int row = (2 * m_pair + static_cast<int>(rank)) * BM + q * 32 + lane;
      const bool row_ok = row < p.M;
      const int col_base = n_tile * BN;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 2 * BN;
      float* krow = p.state + static_cast<size_t>(row_ok ? row : 0) * p.ld_state + p.keep_off + col_base;
      // software pipeline over this warp's 16-column groups: the kept values and the two bias slices of group g + 1 are
      // requested at the top of group g (ncu r02: the first version stalled on these loads -- long scoreboard on the
      // first use of the bias and of the kept values -- with the epilogue fully exposed when TMEM holds one set)
      float kp[16], kn[16];
      float4 bsn[4], btn[4];
      auto load_group = [&](int c0, float* kd, float4* bsd, float4* btd) {
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (row_ok && col_base + c0 + 4 * j4 < p.n_valid) v = *reinterpret_cast<const float4*>(krow + c0 + 4 * j4);
          kd[4 * j4] = v.x; kd[4 * j4 + 1] = v.y; kd[4 * j4 + 2] = v.z; kd[4 * j4 + 3] = v.w;
          bsd[j4] = __ldg(reinterpret_cast<const float4*>(p.bias + col_base + c0) + j4);
          btd[j4] = __ldg(reinterpret_cast<const float4*>(p.bias2 + col_base + c0) + j4);
        }
      };
      float4 bsc[4], btc[4];
      if (ch * 16 < BN) load_group(ch * 16, kp, bsc, btc);   // first group in flight before the accumulator is ready
      mbar_wait(&tfull_bar[acc], acc_phase, 34);
      tc_fence_after();
      float ld_sum = 0.f;
      const bool sv = p.s_save != nullptr && row_ok;
      bool released = false;
      for (int c0 = ch * 16; c0 < BN; c0 += 32) {
        float vs[16], vt[16];
        tmem_ld16(t_row + c0, vs);
        tmem_ld16(t_row + BN + c0, vt);
        if (c0 + 32 < BN) load_group(c0 + 32, kn, bsn, btn);
        tmem_ld_wait();
        if (c0 + 32 >= BN) {                         // last group of this warp: hand the accumulator set back early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(leader_tempty[acc]);
          released = true;
        }
        const int gcol = col_base + c0;
        const int nv = p.n_valid - gcol < 16 ? p.n_valid - gcol : 16;   // <= 0 for padded groups
        if (sv && nv > 0) {
          float* ks = p.keep_save + static_cast<size_t>(row) * p.ld_save + gcol;
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4)
            if (4 * j4 < nv) *reinterpret_cast<float4*>(ks + 4 * j4) = make_float4(kp[4 * j4], kp[4 * j4 + 1], kp[4 * j4 + 2], kp[4 * j4 + 3]);
        }
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float bsv[4] = {bsc[j4].x, bsc[j4].y, bsc[j4].z, bsc[j4].w};
          const float btv[4] = {btc[j4].x, btc[j4].y, btc[j4].z, btc[j4].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int j = 4 * j4 + e;
            // bf16 mode: MUFU tanh / ex2 (error below the bf16 GEMM error); fp32-parity mode: libm-accurate tanhf / expf
            const float s = kSplit == 2 ? tanhf(vs[j] + bsv[e]) : tanh_approx(vs[j] + bsv[e]);
            const float t = vt[j] + btv[e];
            if (!kRev) {
              kp[j] = kp[j] * (kSplit == 2 ? expf(s) : ex2_approx(s * 1.4426950408889634f)) + t;
              ld_sum += s;                           // padded columns: zero weights and biases -> s = 0
            } else {
              kp[j] = (kp[j] - t) * (kSplit == 2 ? expf(-s) : ex2_approx(s * -1.4426950408889634f));
            }
            vs[j] = s;
          }
        }
        if (row_ok && nv > 0) {
          if (sv) {
            float* ss = p.s_save + static_cast<size_t>(row) * p.ld_save + gcol;
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4)
              if (4 * j4 < nv) *reinterpret_cast<float4*>(ss + 4 * j4) = make_float4(vs[4 * j4], vs[4 * j4 + 1], vs[4 * j4 + 2], vs[4 * j4 + 3]);
          }
          if (p.u_hi != nullptr) {                   // bf16 copy for this block's second half-step
            __nv_bfloat16* uh = p.u_hi + static_cast<size_t>(row) * p.ld_u + gcol;
            __nv_bfloat16* ul = kSplit == 2 ? p.u_lo + static_cast<size_t>(row) * p.ld_u + gcol : nullptr;
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              if (4 * j4 < nv) {
                __nv_bfloat16 h[4], l[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) split_bf16(kp[4 * j4 + e], h[e], l[e]);
                *reinterpret_cast<uint2*>(uh + 4 * j4) = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
                if (kSplit == 2)
                  *reinterpret_cast<uint2*>(ul + 4 * j4) = make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
              }
            }
          }
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4)
            if (4 * j4 < nv)
              *reinterpret_cast<float4*>(krow + c0 + 4 * j4) = make_float4(kp[4 * j4], kp[4 * j4 + 1], kp[4 * j4 + 2], kp[4 * j4 + 3]);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) kp[j] = kn[j];
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) { bsc[j4] = bsn[j4]; btc[j4] = btn[j4]; }
      }
      if (!released) {                               // warps without a column group still release TMEM
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(leader_tempty[acc]);
      }
      if (!kRev && row_ok) atomicAdd(p.logdet + row, ld_sum);   // 2 column groups (and possibly several n-tiles) per row
    }
  }

  tc_fence_before();
  cluster_sync_all();                                // neither CTA leaves while the peer may still signal / read it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, c.tmem_cols);
  }
}

}  // namespace inn
