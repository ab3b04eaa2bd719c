// Weight-gradient GEMM of the INN subnetworks on tcgen05 (sm_100a), contraction over the batch.
//
//   dW_g[n, k] = sum_b dY_g[b, n] * X_g[b, k]          db_g[n] = sum_b dY_g[b, n]
//
// for the G subnetworks of a lockstep chain in one launch.  The reference has no code of its own for this --
// it is what torch.autograd runs for nn.Linear (models/modules/INN/modules.py:171-188, addmm backward).
//
// Both operands are read exactly as the forward / dgrad kernels left them: row-major [B, features] bf16
// (hi [+ lo]).  The contraction index (batch) is the ROW index of both, i.e. both operands are "MN-major" for
// the tensor core: a TMA box {64 features, 64 batch rows} with the 128-B swizzle is one canonical MN-major
// SW128 chunk (8 batch rows x 128 B = one 1024-B swizzle atom, chunks of 64 features LBO apart).
//   A = dY^T : M = 128 output features (2 chunks), K = 64 batch rows per stage
//   B = X^T  : N = BN input features (BN/64 chunks)
// The bias gradient rides on the tensor core too: tiles with n_tile == 0 issue one extra N=16 MMA per k-step
// against an all-ones B tile, so column BN.. of the accumulator holds sum_b dY[b, n].
//
// kSplit = 2 (fp32-parity mode): lo*hi + hi*lo + hi*hi as in the forward kernel.
//
// kPair (large batches): a tcgen05 CTA pair (cta_group::2) on a 256 x BN tile.  A 128 x 128 tile pulls 2 x 16 KB of
// operands per 64 batch rows for 4 MMAs -- at cfg A / 4096 that is 604 MB through the L2 -> SM path per hidden x hidden
// launch, twice what the tensor pipe's 27 us allow (ncu r01 v12: 48 us, tensor pipe 41 %).  In a pair each CTA stages its
// own 128 output features of dY^T and HALF of the X^T columns (the MMA reads the other half from the peer's shared
// memory): 24 KB per CTA and k-block for twice the tile, 2.7 x fewer bytes per FLOP.  Each CTA's epilogue drains its own
// 128 accumulator rows exactly as the single-CTA kernel does.
// Warp roles as in gemm_kernel.cuh: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// warps 4-11 epilogue (TMEM -> regs -> padded smem tile -> coalesced fp32 stores / atomics into the
// parameter-gradient tensors, which keep the reference's [out, in] layout and true sizes).
#pragma once
#include "gemm_kernel.cuh"

namespace inn {

constexpr int WG_CHUNK_BYTES = 64 * 128;          // one {64 features x 64 batch rows} bf16 box
constexpr int WG_ONES_BYTES = 2 * 1024;           // 16 k-rows x 128 B of bf16 ones (covers any N=16 MN-major read)
constexpr int WG_MAX_GROUPS = 4;

struct WgradParams {
  int B;                  // batch rows (contraction length)
  int num_kb;             // ceil(B / 64)
  int G;                  // subnetworks in this launch
  int m_tiles;            // per group: ceil(rows_p / 128)
  int n_tiles;            // per group: ceil(kin_p / BN)
  int kslices;            // split of the batch loop across CTAs (atomics when > 1)
  int BN;                 // 64, 128 or 256
  int stages;
  int tmem_cols, nacc;
  int epi_off;            // byte offset of the epilogue staging tile
  int epi_alias;          // 1: the staging tile reuses the pipeline stages (one tile per CTA: they are idle once the last
                          //    MMA has retired), which is what lets a 128 x 256 tile keep 4 stages in flight
  int dy_col0, dy_group_stride;   // columns of dY: dy_col0 + g * stride + m_tile * 128
  int x_col0, x_group_stride;     // columns of X
  int rows_valid;         // true number of output features per group (H or C/2)
  int kin_p;              // padded input width (columns of X per group)
  int seg0_valid;         // X columns [0, seg0_valid)                 -> gradient columns [0, seg0_valid)
  int seg1_start;         // X columns [seg1_start, +seg1_valid)       -> gradient columns [seg0_valid, ...)
  int seg1_valid;
  int ld_gw;              // row stride of the gradient tensors (= true fan-in)
  float* gw[WG_MAX_GROUPS];
  float* gb[WG_MAX_GROUPS];
};

__host__ inline int wgrad_stage_bytes(int ksplit, int BN) { return ksplit * (2 + BN / 64) * WG_CHUNK_BYTES; }
__host__ __device__ inline int wgrad_epi_bytes(int BN) { return BM * (BN + 1) * 4; }
__host__ inline int wgrad_smem_bytes(int ksplit, int BN, int stages, int epi_alias = 0) {
  return 1024 + stages * wgrad_stage_bytes(ksplit, BN) + WG_ONES_BYTES +
         (epi_alias ? 0 : ((wgrad_epi_bytes(BN) + 15) & ~15)) + SMEM_BARRIER_BYTES;
}

template <int kSplit, bool kPair = false>
__global__ void __launch_bounds__(WG_THREADS, 1)
inn_wgrad_kernel(const __grid_constant__ CUtensorMap tmY0, const __grid_constant__ CUtensorMap tmY1,
                 const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                 const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform role index (see gemm_kernel.cuh)
  const int lane = threadIdx.x & 31;
  const int BN = p.BN;
  const int n_chunks = kPair ? (BN >> 7) : (BN >> 6);          // X^T chunks this CTA stages (a pair: half of the tile's)
  const int stage_bytes = kSplit * (2 + n_chunks) * WG_CHUNK_BYTES;
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;        // 0 = leader
  const int cta_id = kPair ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int cta_stride = kPair ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int stages = p.stages;
  const int acc_cols = BN + 16;                 // + the bias columns
  const int nacc = p.nacc;

  uint8_t* ones_smem = smem + stages * stage_bytes;
  float* epi_tile = p.epi_alias ? reinterpret_cast<float*>(smem) : reinterpret_cast<float*>(ones_smem + WG_ONES_BYTES);
  const int epi_bn = (kPair && BN > 128) ? 128 : BN;   // width of the staging tile (two passes for 256-wide pair tiles)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ones_smem + WG_ONES_BYTES +
                                                   (p.epi_alias ? 0 : ((wgrad_epi_bytes(epi_bn) + 15) & ~15)));
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmY0);
    tma_prefetch_desc(&tmX0);
    if (kSplit == 2) {
      tma_prefetch_desc(&tmY1);
      tma_prefetch_desc(&tmX1);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], kPair ? 16 : WG_EPI_THREADS);     // pair: one arrival per epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if (kPair) tmem_alloc_pair(tmem_slot, p.tmem_cols);
    else tmem_alloc(tmem_slot, p.tmem_cols);
  }
  if (warp == 3) {                               // all-ones B tile for the bias column sums
    uint32_t* o = reinterpret_cast<uint32_t*>(ones_smem);
    for (int i = lane; i < WG_ONES_BYTES / 4; i += 32) o[i] = 0x3F803F80u;   // two bf16 1.0
    fence_proxy_async();
  }
  tc_fence_before();
  if (kPair) cluster_sync_all();                 // both CTAs' barriers, TMEM and ones tile exist before anyone signals
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) grid_dep_launch();

  // pair: p.m_tiles counts 256-row tiles (the host passes ceil(rows / 256)); CTA `rank` owns rows [rank * 128, +128) of it
  const int tiles_per_group = p.m_tiles * p.n_tiles;
  const int num_tiles = p.kslices * p.G * tiles_per_group;
  const int kb_per_slice = (p.num_kb + p.kslices - 1) / p.kslices;

  if (warp == 0) {
    // ============================ TMA producer ============================
    grid_dep_wait();
    int stage = 0;
    uint32_t phase = 0;
    uint32_t leader_full[MAX_STAGES];
#pragma unroll
    for (int s = 0; s < MAX_STAGES; ++s) leader_full[s] = kPair ? mapa_u32(&full_bar[s], 0) : 0u;
    for (int tile = cta_id; tile < num_tiles; tile += cta_stride) {
      const int slice = tile / (p.G * tiles_per_group);
      int rem = tile - slice * p.G * tiles_per_group;
      const int g = rem / tiles_per_group;
      rem -= g * tiles_per_group;
      const int m_tile = rem / p.n_tiles;
      const int n_tile = rem - m_tile * p.n_tiles;
      const int kb0 = slice * kb_per_slice;
      const int kb1 = min(p.num_kb, kb0 + kb_per_slice);
      const int y_col = p.dy_col0 + g * p.dy_group_stride + (kPair ? (2 * m_tile + static_cast<int>(rank)) : m_tile) * BM;
      const int x_col = p.x_col0 + g * p.x_group_stride + n_tile * BN + (kPair ? static_cast<int>(rank) * (BN >> 1) : 0);
      for (int kb = kb0; kb < kb1; ++kb) {
        {
          mbar_wait(&empty_bar[stage], phase ^ 1, 11);
          uint8_t* sbase = smem + stage * stage_bytes;
          if (kPair) {
            if (rank == 0) mbar_arrive_expect_tx_el(&full_bar[stage], 2 * stage_bytes);   // both CTAs' bytes
#pragma unroll
            for (int sp = 0; sp < kSplit; ++sp) {
              uint8_t* a = sbase + sp * 2 * WG_CHUNK_BYTES;
              tma_load_2d_pair_el(a, sp == 0 ? &tmY0 : &tmY1, leader_full[stage], y_col, kb * 64);
              tma_load_2d_pair_el(a + WG_CHUNK_BYTES, sp == 0 ? &tmY0 : &tmY1, leader_full[stage], y_col + 64, kb * 64);
              uint8_t* b = sbase + kSplit * 2 * WG_CHUNK_BYTES + sp * n_chunks * WG_CHUNK_BYTES;
              for (int c = 0; c < n_chunks; ++c)
                tma_load_2d_pair_el(b + c * WG_CHUNK_BYTES, sp == 0 ? &tmX0 : &tmX1, leader_full[stage], x_col + c * 64, kb * 64);
            }
          } else {
          mbar_arrive_expect_tx_el(&full_bar[stage], stage_bytes);
#pragma unroll
          for (int sp = 0; sp < kSplit; ++sp) {
            uint8_t* a = sbase + sp * 2 * WG_CHUNK_BYTES;
            tma_load_2d_el(a, sp == 0 ? &tmY0 : &tmY1, &full_bar[stage], y_col, kb * 64);
            tma_load_2d_el(a + WG_CHUNK_BYTES, sp == 0 ? &tmY0 : &tmY1, &full_bar[stage], y_col + 64, kb * 64);
            uint8_t* b = sbase + kSplit * 2 * WG_CHUNK_BYTES + sp * n_chunks * WG_CHUNK_BYTES;
            for (int c = 0; c < n_chunks; ++c)
              tma_load_2d_el(b + c * WG_CHUNK_BYTES, sp == 0 ? &tmX0 : &tmX1, &full_bar[stage], x_col + c * 64, kb * 64);
          }
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ============================ MMA issuer (pair: leader CTA only) ==============================
    const uint32_t idesc = umma_idesc_bf16(kPair ? 2 * BM : BM, BN, 1, 1);
    const uint32_t idesc_b = umma_idesc_bf16(kPair ? 2 * BM : BM, 16, 1, 1);
    auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t id, uint32_t accum) {
      if (kPair) umma_bf16_ss_pair_el(d, da, db, id, accum);
      else umma_bf16_ss_el(d, da, db, id, accum);
    };
    const uint64_t d_ones = umma_desc_mn_sw128(smem_u32(ones_smem), 1024);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = cta_id; tile < num_tiles; tile += cta_stride, ++it) {
      const int slice = tile / (p.G * tiles_per_group);
      const int rem = tile - slice * p.G * tiles_per_group;
      const int n_tile = (rem % tiles_per_group) % p.n_tiles;
      const bool want_bias = n_tile == 0;
      const int kb0 = slice * kb_per_slice;
      const int kb1 = min(p.num_kb, kb0 + kb_per_slice);
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 12);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * acc_cols;
      for (int kb = kb0; kb < kb1; ++kb) {
        {
          mbar_wait(&full_bar[stage], phase, 13);
          tc_fence_after();
          const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
          const uint32_t a_hi = sbase;
          const uint32_t a_lo = sbase + (kSplit - 1) * 2 * WG_CHUNK_BYTES;
          const uint32_t b_hi = sbase + kSplit * 2 * WG_CHUNK_BYTES;
          const uint32_t b_lo = b_hi + (kSplit - 1) * n_chunks * WG_CHUNK_BYTES;
          const uint64_t da_hi = umma_desc_mn_sw128(a_hi, WG_CHUNK_BYTES);
          const uint64_t da_lo = umma_desc_mn_sw128(a_lo, WG_CHUNK_BYTES);
          const uint64_t db_hi = umma_desc_mn_sw128(b_hi, WG_CHUNK_BYTES);
          const uint64_t db_lo = umma_desc_mn_sw128(b_lo, WG_CHUNK_BYTES);
#pragma unroll
          for (int kk = 0; kk < 64 / UMMA_K; ++kk) {
            const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 128) >> 4);   // 16 batch rows x 128 B
            const uint32_t first = (kb == kb0 && kk == 0) ? 0u : 1u;
            if (kSplit == 2) {
              mma(d_tmem, da_lo + koff, db_hi + koff, idesc, first);
              mma(d_tmem, da_hi + koff, db_lo + koff, idesc, 1u);
              mma(d_tmem, da_hi + koff, db_hi + koff, idesc, 1u);
              if (want_bias) {
                mma(d_tmem + BN, da_lo + koff, d_ones, idesc_b, first);
                mma(d_tmem + BN, da_hi + koff, d_ones, idesc_b, 1u);
              }
            } else {
              mma(d_tmem, da_hi + koff, db_hi + koff, idesc, first);
              if (want_bias) mma(d_tmem + BN, da_hi + koff, d_ones, idesc_b, first);
            }
          }
          if (kPair) {
            umma_commit_pair_el(&empty_bar[stage]);              // frees the slot in BOTH CTAs when the MMAs retire
            if (kb == kb1 - 1) umma_commit_pair_el(&tfull_bar[acc]);
          } else {
            umma_commit_el(&empty_bar[stage]);
            if (kb == kb1 - 1) umma_commit_el(&tfull_bar[acc]);
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ============================ epilogue ================================
    grid_dep_wait();
    const int q = warp & 3;                                       // TMEM lane quarter (== warp % 4)
    const int ch = (warp - 4) >> 2;                               // column half this warp moves out of TMEM
    const int we = warp - 4;                                      // 0..7: row phase of the write-out
    const int r_loc = q * 32 + lane;
    // pair tiles 256 columns wide leave one accumulator set in TMEM and no room for a 128 x 256 staging tile: they are
    // drained and written out as two 128-column passes through the same 64 KB tile
    const int EBN = (kPair && BN > 128) ? 128 : BN;              // columns per epilogue pass
    const int n_pass = BN / EBN;
    const int ldt = EBN;
    const bool atomic = p.kslices > 1;
    const uint32_t leader_tempty[2] = {kPair ? mapa_u32(&tempty_bar[0], 0) : 0u, kPair ? mapa_u32(&tempty_bar[1], 0) : 0u};
    int it = 0;
    for (int tile = cta_id; tile < num_tiles; tile += cta_stride, ++it) {
      const int slice = tile / (p.G * tiles_per_group);
      int rem = tile - slice * p.G * tiles_per_group;
      const int g = rem / tiles_per_group;
      rem -= g * tiles_per_group;
      const int m_tile = kPair ? 2 * (rem / p.n_tiles) + static_cast<int>(rank) : rem / p.n_tiles;   // this CTA's 128 rows
      const int n_tile = rem - (rem / p.n_tiles) * p.n_tiles;
      const int kb0 = slice * kb_per_slice;
      const int kb1 = min(p.num_kb, kb0 + kb_per_slice);
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * acc_cols;
      if (kb0 >= kb1) continue;                                  // empty slice (never scheduled by the host)
      mbar_wait(&tfull_bar[acc], acc_phase, 14);
      tc_fence_after();
      // staging tile: row stride BN floats, the 16-byte chunk index XORed with (row & 7) -- the drain (lanes = rows) and
      // the write-out (lanes = chunks of one row) are both conflict-free with 128-bit shared accesses
      float4* my4 = reinterpret_cast<float4*>(epi_tile + r_loc * ldt);
      const int row = m_tile * BM + r_loc;                       // output feature inside the group
      for (int pass = 0; pass < n_pass; ++pass) {
      if (pass > 0) named_bar_sync(1, WG_EPI_THREADS);            // the previous pass's write-out has read the tile
      for (int c0 = ch * (EBN >> 1); c0 < (ch + 1) * (EBN >> 1); c0 += 16) {
        float v[16];
        tmem_ld16(t_row + pass * EBN + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          my4[((c0 >> 2) + j) ^ (r_loc & 7)] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      if (pass == n_pass - 1) {
      if (n_tile == 0 && ch == 0) {
        float v[16];
        tmem_ld16(t_row + BN, v);
        tmem_ld_wait();
        if (row < p.rows_valid && p.gb[g] != nullptr) {
          if (atomic) atomicAdd(p.gb[g] + row, v[0]);
          else p.gb[g][row] = v[0];
        }
      }
      tc_fence_before();
      if (kPair) {
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(leader_tempty[acc]);
      } else {
        mbar_arrive(&tempty_bar[acc]);
      }
      }
      named_bar_sync(1, WG_EPI_THREADS);
      // coalesced write-out: consecutive threads -> consecutive gradient columns of one row
      float* gw = p.gw[g];
      const int col_base = n_tile * BN + pass * EBN;
      // one warp per row, lanes along the columns: the column -> gradient-column map is row-invariant (computed once
      // per tile), and two rows' smem reads are issued before their stores so the loop is not one dependent
      // ld.shared -> st.global chain per element (at batch 64 the write-out IS the kernel: ncu r01 v11)
      const int ncb = EBN >> 5;                                     // 32-column groups per row: 2, 4 or 8
      int gc[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int xcol = col_base + lane + 32 * k;
        gc[k] = -1;
        if (k < ncb) {
          if (xcol < p.seg0_valid) gc[k] = xcol;
          else if (xcol >= p.seg1_start && xcol < p.seg1_start + p.seg1_valid) gc[k] = xcol - p.seg1_start + p.seg0_valid;
        }
      }
      const int rows_here = min(BM, p.rows_valid - m_tile * BM);
      const bool vec = !atomic && ((p.seg0_valid | p.seg1_start | p.seg1_valid | p.ld_gw) & 3) == 0 &&
                       (reinterpret_cast<uintptr_t>(gw) & 15) == 0;
      if (vec) {
        // one warp per row, one 16-byte chunk per lane, four rows in flight
        const int nch = EBN >> 2;                                   // chunks per row: 16, 32 or 64
        int gc4[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int chunk = lane + 32 * k;
          const int xcol = col_base + 4 * chunk;
          gc4[k] = -1;
          if (chunk < nch) {
            if (xcol < p.seg0_valid) gc4[k] = xcol;
            else if (xcol >= p.seg1_start && xcol < p.seg1_start + p.seg1_valid) gc4[k] = xcol - p.seg1_start + p.seg0_valid;
          }
        }
        for (int r0 = we; r0 < rows_here; r0 += 32) {
          float4 v[4][2];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int r = r0 + 8 * i;
            const float4* s4 = reinterpret_cast<const float4*>(epi_tile + r * ldt);
#pragma unroll
            for (int k = 0; k < 2; ++k)
              if (r < rows_here && gc4[k] >= 0) v[i][k] = s4[(lane + 32 * k) ^ (r & 7)];
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int r = r0 + 8 * i;
            float* d = gw + static_cast<size_t>(m_tile * BM + r) * p.ld_gw;
#pragma unroll
            for (int k = 0; k < 2; ++k)
              if (r < rows_here && gc4[k] >= 0) *reinterpret_cast<float4*>(d + gc4[k]) = v[i][k];
          }
        }
      } else {
      for (int r = we; r < rows_here; r += 16) {
        const int r2 = r + 8;
        const bool two = r2 < rows_here;
        float va[8], vb[8];
        const float* sa = epi_tile + r * ldt;
        const float* sb = epi_tile + r2 * ldt;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int c = lane + 32 * k;
          va[k] = (k < ncb) ? sa[((((c >> 2) ^ (r & 7)) << 2) | (c & 3))] : 0.f;
          vb[k] = (k < ncb && two) ? sb[((((c >> 2) ^ (r2 & 7)) << 2) | (c & 3))] : 0.f;
        }
        float* da = gw + static_cast<size_t>(m_tile * BM + r) * p.ld_gw;
        float* db = gw + static_cast<size_t>(m_tile * BM + r2) * p.ld_gw;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < ncb && gc[k] >= 0) {
            if (atomic) {
              atomicAdd(da + gc[k], va[k]);
              if (two) atomicAdd(db + gc[k], vb[k]);
            } else {
              da[gc[k]] = va[k];
              if (two) db[gc[k]] = vb[k];
            }
          }
        }
      }
      }
      }                                                             // pass
      named_bar_sync(1, WG_EPI_THREADS);                            // staging tile is reused by the next tile
    }
  }

  tc_fence_before();
  if (kPair) cluster_sync_all();                 // neither CTA leaves while the peer may still signal / read it
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace inn
