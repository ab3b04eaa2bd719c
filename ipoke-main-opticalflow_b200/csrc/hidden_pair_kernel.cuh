// hidden_pair_kernel.cuh -- hidden Linear + bias + LeakyReLU (modules.py:171-180) for multi-wave launches as a
// tcgen05 CTA-pair GEMM (cta_group::2): two SMs of one TPC compute one 256 x BN output tile.
//
// Why: with one CTA per 128 x 256 tile every k-block pulls 16 KB of A and 32 KB of W through the L2 -> SM path for
// 4 MMAs (0.32 us of tensor time): 150 GB/s per SM, 22 TB/s over the chip -- more than the fabric delivers (ncu r01 v7:
// 15 TB/s, tensor pipe 68 % active at cfg B, K = 1024).  In a pair each CTA stages its own 128 rows of A and only HALF of
// the W tile (the MMA reads the other half from the peer's shared memory), so the same tensor work needs 32 KB per SM
// and k-block, and the freed shared memory holds 6 pipeline stages instead of 4.
//
// Roles per CTA (384 threads): warp 0 TMA producer (both CTAs; bytes are counted on the LEADER's full barrier),
// warp 1 MMA issuer (leader = cluster rank 0 only; commits are multicast to both CTAs' barriers), warp 2 TMEM
// allocator, warps 4-11 epilogue (both CTAs, each on its own 128 accumulator rows; warp-private TMA-store slabs,
// no CTA-wide barrier anywhere in the steady state).
#pragma once
#include "gemm_kernel.cuh"

namespace inn {

struct PairParams {
  int M;                  // valid rows
  int num_m_pairs;        // ceil(M / 256)
  int num_n_tiles;        // G * Hp / BN
  int tiles_per_group;    // Hp / BN
  int num_k_blocks;       // K / 64
  int BN;                 // 128 or 256: UMMA N of the pair; each CTA stages BN / 2 rows of W
  int stages;
  int tmem_cols;          // power of two >= nacc * BN
  int nacc;               // accumulator stages in TMEM
  int epi_off;            // byte offset of the epilogue smem region = stages * stage_bytes
  int a_group_stride;     // A column offset (elements) per subnetwork group (0 for a chain's first layer)
  int a_m_wrap;           // > 0: A holds a_m_wrap 128-row tiles, tile t reads t % a_m_wrap (multi-draw 'cond' blocks)
  int wide;               // 1: 64-column store boxes (4 KB per warp and split), 0: 32-column boxes (2 KB)
  const float* bias;      // [G * Hp]
  float slope;
  // kDgrad (dX = dY * W, W read MN-major from the forward layout; epilogue = LeakyReLU' mask instead of bias + LeakyReLU)
  int w_row0;             // first W row (contraction index) of group 0
  int w_group_rows;       // W rows per group
  int out_col0;           // first output column; tile (., n) writes columns out_col0 + n * BN ...
  const __nv_bfloat16* mask_act;   // saved post-LeakyReLU activation (hi part) at the output position
  int ld_mask, mask_col0;
  int rev;                // 1: walk the tiles from the last row tile down.  Consecutive launches of a program alternate, so a
                          // layer starts on the rows its producer wrote last (the part of its input still in L2)
};

__host__ inline int pair_stage_bytes(int ksplit, int BN) { return ksplit * (A_TILE_BYTES + (BN / 2) * BK * 2); }
__host__ inline int pair_epi_bytes(int ksplit, int wide) { return ksplit * 8 * (wide ? 4096 : 2048); }
__host__ inline int pair_smem_bytes(int ksplit, int BN, int stages, int wide) {
  return 1024 + stages * pair_stage_bytes(ksplit, BN) + pair_epi_bytes(ksplit, wide) + SMEM_BARRIER_BYTES;
}

// Sign mask of a 32-row x 32-column bf16 block, fetched by the whole warp with coalesced loads: lanes 0-15 read the 64
// bytes of an even row, lanes 16-31 those of the next row (16 independent 4-byte loads per lane, all issued before the
// first use), two ballots per row pair turn the signs into bit masks, and lane L keeps row L's: bit k of `even` / `odd`
// = (element [L][2k] / [L][2k + 1] > 0).  A 16-byte load per thread and row instead costs 32 LSU transactions per
// instruction (rows are ld * 2 bytes apart): 17 % of the pair dgrad at cfg A / 4096 (measured with the mask switched off).
__device__ __forceinline__ void warp_sign_mask32(const __nv_bfloat16* base, int ld, int rows_valid, int lane,
                                                 uint32_t& even, uint32_t& odd) {
  uint32_t v[16];
  const int sub = lane >> 4, cl = lane & 15;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int r = 2 * i + sub;
    v[i] = (base != nullptr && r < rows_valid)
               ? __ldg(reinterpret_cast<const uint32_t*>(base + static_cast<size_t>(r) * ld) + cl) : 0x3F803F80u;
  }
  even = odd = 0xFFFFFFFFu;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const uint32_t b0 = __ballot_sync(0xffffffffu, bf16_bits_positive(v[i] & 0xFFFFu));
    const uint32_t b1 = __ballot_sync(0xffffffffu, bf16_bits_positive(v[i] >> 16));
    if ((lane >> 1) == i) {
      even = (lane & 1) ? (b0 >> 16) : (b0 & 0xFFFFu);
      odd = (lane & 1) ? (b1 >> 16) : (b1 & 0xFFFFu);
    }
  }
}

template <int kSplit, bool kDgrad = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
inn_hidden_pair_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                       const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                       const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                       const PairParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int BN = p.BN;
  const int w_half_bytes = (BN / 2) * BK * 2;
  const int stage_bytes = kSplit * (A_TILE_BYTES + w_half_bytes);
  const int stages = p.stages;
  const int nacc = p.nacc;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.num_m_pairs * p.num_n_tiles;

  uint8_t* epi_smem = smem + p.epi_off;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_smem + kSplit * 8 * (p.wide ? 4096 : 2048));
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    tma_prefetch_desc(&tmW0);
    tma_prefetch_desc(&tmO0);
    if (kSplit == 2) {
      tma_prefetch_desc(&tmA1);
      tma_prefetch_desc(&tmW1);
      tma_prefetch_desc(&tmO1);
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
  if (warp == 2) tmem_alloc_pair(tmem_slot, p.tmem_cols);
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
      const int m_pair = (p.rev ? num_tiles - 1 - tile : tile) / p.num_n_tiles;
      const int n_tile = (p.rev ? num_tiles - 1 - tile : tile) - m_pair * p.num_n_tiles;
      const int group = n_tile / p.tiles_per_group;
      const int m_tile = 2 * m_pair + static_cast<int>(rank);
      const int a_row0 = (p.a_m_wrap > 0 ? m_tile % p.a_m_wrap : m_tile) * BM;
      const int w_row0 = n_tile * BN + static_cast<int>(rank) * (BN / 2);
      // kDgrad: this CTA's half of the tile's W columns, inside the group's column range
      const int w_c0 = (n_tile - group * p.tiles_per_group) * BN + static_cast<int>(rank) * (BN / 2);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 21);
        if (rank == 0) mbar_arrive_expect_tx_el(&full_bar[stage], 2 * stage_bytes);
        uint8_t* sbase = smem + stage * stage_bytes;
        const int a_k = group * p.a_group_stride + kb * BK;
#pragma unroll
        for (int sp = 0; sp < kSplit; ++sp) {
          tma_load_2d_pair_el(sbase + sp * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, leader_full[stage], a_k, a_row0);
          if (kDgrad) {
            // contraction index = W row: {64 rows} x {BN / 2 columns} as MN-major chunks of 64 columns (8 KB each)
            const int w_rk = p.w_row0 + group * p.w_group_rows + kb * BK;
            for (int c = 0; c < (BN >> 7); ++c)
              tma_load_2d_pair_el(sbase + kSplit * A_TILE_BYTES + sp * w_half_bytes + c * (64 * 128),
                                  sp == 0 ? &tmW0 : &tmW1, leader_full[stage], w_c0 + c * 64, w_rk);
          } else {
            tma_load_2d_pair_el(sbase + kSplit * A_TILE_BYTES + sp * w_half_bytes, sp == 0 ? &tmW0 : &tmW1,
                                leader_full[stage], kb * BK, w_row0);
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ============================ MMA issuer (leader CTA only) ============================
    const uint32_t idesc = umma_idesc_bf16(2 * BM, BN, 0, kDgrad ? 1 : 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 22);      // both CTAs' epilogues have drained this accumulator
      tc_fence_after();
      const uint32_t d_base = tmem_base + acc * BN;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&full_bar[stage], phase, 23);
        tc_fence_after();
        const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
        const uint32_t w0 = sbase + kSplit * A_TILE_BYTES;
        const uint64_t da_hi = umma_desc_k_sw128(sbase);
        const uint64_t dw_hi = kDgrad ? umma_desc_mn_sw128(w0, 64 * 128) : umma_desc_k_sw128(w0);
        const uint64_t da_lo = umma_desc_k_sw128(sbase + (kSplit - 1) * A_TILE_BYTES);
        const uint64_t dw_lo = kDgrad ? umma_desc_mn_sw128(w0 + (kSplit - 1) * w_half_bytes, 64 * 128)
                                      : umma_desc_k_sw128(w0 + (kSplit - 1) * w_half_bytes);
#pragma unroll
        for (int kk = 0; kk < BK / UMMA_K; ++kk) {
          const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 2) >> 4);
          // MN-major W: one k-row is 128 B, UMMA_K rows further = UMMA_K * 128 B
          const uint64_t woff = kDgrad ? static_cast<uint64_t>((kk * UMMA_K * 128) >> 4) : koff;
          const uint32_t first = (kb == 0 && kk == 0) ? 0u : 1u;
          if (kSplit == 2) {                                 // hi*hi + hi*lo + lo*hi (drops lo*lo, ~2^-16 relative)
            umma_bf16_ss_pair_el(d_base, da_lo + koff, dw_hi + woff, idesc, first);
            umma_bf16_ss_pair_el(d_base, da_hi + koff, dw_lo + woff, idesc, 1u);
            umma_bf16_ss_pair_el(d_base, da_hi + koff, dw_hi + woff, idesc, 1u);
          } else {
            umma_bf16_ss_pair_el(d_base, da_hi + koff, dw_hi + woff, idesc, first);
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
    const int ch = (warp - 4) >> 2;
    const int we = warp - 4;
    const bool wide = p.wide != 0;
    const int wbytes = wide ? 4096 : 2048;
    const int n_slabs = BN >> 6;
    uint8_t* wb_hi = epi_smem + (we * kSplit) * wbytes + lane * (wide ? 128 : 64);
    uint8_t* wb_lo = wb_hi + wbytes;
    const uint32_t leader_tempty[2] = {mapa_u32(&tempty_bar[0], 0), mapa_u32(&tempty_bar[1], 0)};
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int m_pair = (p.rev ? num_tiles - 1 - tile : tile) / p.num_n_tiles;
      const int n_tile = (p.rev ? num_tiles - 1 - tile : tile) - m_pair * p.num_n_tiles;
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      const int row0 = (2 * m_pair + static_cast<int>(rank)) * BM + q * 32;     // first row of this warp's slab
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      const float* bias_t = kDgrad ? nullptr : p.bias + n_tile * BN;
      const int ocol0 = (kDgrad ? p.out_col0 : 0) + n_tile * BN;              // first output column of this tile
      // kDgrad: LeakyReLU'(pre) from the sign of the saved activation at the output position (modules.py:174); the
      // 32 x 32 sign block of a chunk is fetched one chunk ahead (and the first one before the accumulator is ready)
      const __nv_bfloat16* mask_blk = (kDgrad && p.mask_act != nullptr && row0 < p.M)
          ? p.mask_act + static_cast<size_t>(row0) * p.ld_mask + p.mask_col0 + ocol0 : nullptr;
      const int mask_rows = p.M - row0;
      // all of this warp's sign blocks of the tile (<= 4: BN <= 256) are fetched here, while the tile's MMAs still run
      uint32_t me[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mo[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
      if (kDgrad) {
        int ci = 0;
        for (int s = wide ? ch : 0; s < n_slabs; s += wide ? 2 : 1)
          for (int h = 0; h < (wide ? 2 : 1); ++h, ++ci)
            if (ci < 4)
              warp_sign_mask32(mask_blk != nullptr ? mask_blk + s * 64 + (wide ? h * 32 : ch * 32) : nullptr, p.ld_mask,
                               mask_rows, lane, me[ci], mo[ci]);
      }
      int chunk_i = 0;
      mbar_wait(&tfull_bar[acc], acc_phase, 24);
      tc_fence_after();
      for (int s = wide ? ch : 0; s < n_slabs; s += wide ? 2 : 1) {
        for (int h = 0; h < (wide ? 2 : 1); ++h) {
          const int c_in = wide ? h * 32 : ch * 32;
          float v[32];
#pragma unroll
          for (int c = 0; c < 2; ++c) tmem_ld16(t_row + s * 64 + c_in + c * 16, v + c * 16);
          const uint32_t mk_e = me[chunk_i & 3], mk_o = mo[chunk_i & 3];
          ++chunk_i;
          tmem_ld_wait();
          if (wide ? (h == 1 && s + 2 >= n_slabs) : (s + 1 >= n_slabs)) {     // this warp is done with the accumulator
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_tempty[acc]);
          }
          if (h == 0) {                                      // the store that last read this buffer has drained
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
          }
          const float4* b4 = kDgrad ? nullptr : reinterpret_cast<const float4*>(bias_t + s * 64 + c_in);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float bb[8];
            if (!kDgrad) {
              const float4 b0 = __ldg(b4 + 2 * j), b1 = __ldg(b4 + 2 * j + 1);
              bb[0] = b0.x; bb[1] = b0.y; bb[2] = b0.z; bb[3] = b0.w; bb[4] = b1.x; bb[5] = b1.y; bb[6] = b1.z; bb[7] = b1.w;
            }
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float a, b;
              if (kDgrad) {
                a = v[j * 8 + 2 * e] * (((mk_e >> (j * 4 + e)) & 1u) ? 1.f : p.slope);
                b = v[j * 8 + 2 * e + 1] * (((mk_o >> (j * 4 + e)) & 1u) ? 1.f : p.slope);
              } else {
                a = v[j * 8 + 2 * e] + bb[2 * e];
                b = v[j * 8 + 2 * e + 1] + bb[2 * e + 1];
                a = fmaxf(a, a * p.slope);                   // LeakyReLU, 0 <= slope <= 1
                b = fmaxf(b, b * p.slope);
              }
              if (kSplit == 1) {
                hi[e] = cvt_bf16x2(a, b);
              } else {
                __nv_bfloat16 h0, l0, h1, l1;
                split_bf16(a, h0, l0);
                split_bf16(b, h1, l1);
                hi[e] = pack_bf16x2(h0, h1);
                lo[e] = pack_bf16x2(l0, l1);
              }
            }
            const int chunk = (wide ? ((h * 4 + j) ^ (lane & 7)) : (j ^ ((lane >> 1) & 3))) << 4;
            *reinterpret_cast<uint4*>(wb_hi + chunk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            if (kSplit == 2) *reinterpret_cast<uint4*>(wb_lo + chunk) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0 && row0 < p.M) {
          const int col = ocol0 + s * 64 + (wide ? 0 : ch * 32);
          const int loff = lane * (wide ? 128 : 64);
          tma_store_2d(&tmO0, wb_hi - loff, col, row0);
          if (kSplit == 2) tma_store_2d(&tmO1, wb_lo - loff, col, row0);
          bulk_commit();
        }
      }
    }
    if (lane == 0) bulk_wait<0>();                   // stores landed before exit
  }

  tc_fence_before();
  cluster_sync_all();                                // neither CTA leaves while the peer may still signal / read it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, p.tmem_cols);
  }
}

}  // namespace inn
