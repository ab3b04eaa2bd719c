// Persistent, warp-specialised tcgen05 GEMM for the INN subnetworks (sm_100a).
//
//   acc[sub] (TMEM, fp32) = A_sub[128 x K] (bf16, K-major, TMA/SW128)  x  W_sub[BN x K]^T (bf16, K-major)
//
// One kernel template serves every Linear of BasicFullyConnectedNet (reference
// models/modules/INN/modules.py:165-191) and fuses what follows it:
//   EPI_HIDDEN    bias + LeakyReLU(0.01) -> bf16 activations (optionally hi/lo split), staged through
//                 swizzled smem slabs and written with TMA stores                          [modules.py:173-180]
//   EPI_TAIL_FWD  last Linear of s and t together (two accumulators), tanh, exp, affine update of the kept
//                 half (tile staged in smem, coalesced in/out), per-row log-det sum        [flow_blocks.py:138-142]
//   EPI_TAIL_REV  same GEMMs, (x - t) * exp(-s)                                            [flow_blocks.py:151]
// kSplit = 2 is the fp32-parity mode: activations and weights are carried as bf16 hi + bf16 lo and each
// k-step issues lo*hi + hi*lo + hi*hi (3 MMAs, fp32 accumulate), dropping only the lo*lo term (~2^-16 rel).
//
// Warp roles (384 threads): warp 0 TMA producer, warp 1 MMA issuer (one lane), warp 2 TMEM allocator,
// warps 4-11 epilogue (TMEM lane quarter = warp % 4, column half = (warp - 4) / 4). Rings: smem full/empty (TMA <-> MMA), TMEM full/empty
// (MMA <-> epilogue, double-buffered when the accumulators fit twice in the allocated columns).
#pragma once
#include "ptx.cuh"

namespace inn {

constexpr int BM = 128;                 // UMMA M (rows of the batch per tile)
constexpr int BK = 64;                  // bf16 elements per k-block = one 128-B swizzle row
constexpr int UMMA_K = 16;
constexpr int A_TILE_BYTES = BM * BK * 2;
constexpr int SLAB_BYTES = BM * 128;    // one 128 x 64 bf16 output slab (TMA store box)
constexpr int MAX_STAGES = 8;
constexpr int GEMM_THREADS = 384;       // warps 0-3: TMA / MMA / TMEM alloc / idle; warps 4-11: epilogue
constexpr int EPI_THREADS = 256;        // two warps per TMEM lane quarter, each takes half of the columns
constexpr int WG_THREADS = 384;         // wgrad kernel (wgrad_kernel.cuh): two epilogue warps per TMEM lane quarter as well
constexpr int WG_EPI_THREADS = 256;
constexpr int SMEM_BARRIER_BYTES = 256;

enum EpiMode { EPI_HIDDEN = 0, EPI_TAIL_FWD = 1, EPI_TAIL_REV = 2, EPI_DGRAD = 3, EPI_DU = 4 };

struct GemmParams {
  int M;                  // valid rows
  int num_m_tiles;
  int num_n_tiles;
  int num_k_blocks;       // K / 64 (K zero-padded)
  int BN;                 // UMMA N of one accumulator: 16, 32, 64, 128 or 256
  int bn_log2;
  int stages;
  int tmem_cols;          // power of two >= nacc * NSUB * BN
  int nacc;               // accumulator stages in TMEM (1 or 2)
  int epi_off;            // byte offset of the epilogue smem region (1024-aligned) = stages * stage_bytes
  int w_prefetch;         // 1: fetch the first stages' W tiles before griddepcontrol.wait
  int slab_bufs;          // EPI_HIDDEN: 0 = direct global stores, 1 / 2 / 4 = TMA-store slab buffers per (hi|lo)
  int slab_warp;          // warp-private output slabs, no CTA barrier inside a tile: 1 = tmO box 64 x 32 (128-B swizzle,
                          // slab_bufs >= 2), 2 = tmO box 32 x 32 (64-B swizzle, slab_bufs == 1)
  int tiles_per_group;    // n-tiles per subnetwork group (hidden layers); >= num_n_tiles otherwise
  int a_group_stride;     // A column offset (elements) per group
  int a_off[2];           // A column offset per accumulator
  int w_row0[2];          // first W row per accumulator
  const float* bias;      // hidden: [N]; tail: s bias
  const float* bias2;     // tail: t bias
  // EPI_HIDDEN
  __nv_bfloat16* out_hi;
  __nv_bfloat16* out_lo;
  int ld_out;
  float slope;
  // EPI_TAIL_*
  float* state;           // [M, ld_state] fp32, updated in place at columns keep_off + [0, n_valid)
  int ld_state;
  int keep_off;
  int n_valid;            // true C/2
  int vec4;               // C/2, keep_off and the row strides are multiples of 4: 16-byte state / 8-byte bf16 accesses
  float* logdet;          // [M], forward only
  int logdet_atomic;      // >1 n-tile per row -> atomicAdd
  __nv_bfloat16* u_hi;    // next half-step's MLP input (bf16 copy of the new half), nullable
  __nv_bfloat16* u_lo;
  int ld_u;
  float* s_save;          // training: per half-step tanh output s and the pre-update kept half (nullable)
  float* keep_save;
  int ld_save;
  // Folded block boundary (inference programs): the updated half goes straight into the NEXT block's layout --
  // Shuffle scatter + ActNorm (+ InvLeakyRelu) of the next block in forward, their inverses in reverse
  // (flow_blocks.py:255-274, 322-326, 380-387; modules.py:292-316) -- so no boundary kernel runs between blocks.
  float* f_out;           // nullable = not folded; [M, f_ld] next state (or the caller's output)
  int f_ld;
  const int* f_dst;       // [C] destination column of physical source column (keep_off + c)
  const float* f_loc;     // f_mode 1: indexed by destination column; f_mode 2: by source column
  const float* f_scale;
  int f_mode;             // 0: scatter only, 1: v = scale*(v+loc) [, *alpha if < 0], 2: [v /= alpha if < 0,] v = v/scale - loc
  int f_leaky;
  float f_alpha;
  __nv_bfloat16* f_u_hi;  // next half-step's MLP input: destination columns [f_u_begin, f_u_begin + n_valid)
  __nv_bfloat16* f_u_lo;
  int f_u_begin;
  const float* ld_extra;  // nullable: scalar added once per row (the next block's ActNorm log-det)
  int direct;                  // tail epilogue without the smem transposition (one-wave launches)
  int ks;                      // tail_cluster_kernel.cuh: K slices per subnetwork (cluster = 2 * ks CTAs per tile)
  int a_m_wrap;                // > 0: A holds only a_m_wrap row tiles and m-tile t reads tile t % a_m_wrap -- the 'cond'
                               // couplings of a multi-draw sampling call, whose subnetworks see only the conditioning
                               // (flow_blocks.py:224, 237) and therefore run once per conditioning row, not per draw
  int dbg;                     // debug bit mask (IPOKE_INN_DBG), 0 in production
  unsigned long long* trace;   // nullable debug timeline (ipoke_inn_debug_trace): 16 slots for this launch
  // backward (EPI_DGRAD / EPI_DU): dX = dY * W with W read MN-major straight from the forward layout
  int w_mn;               // 1: the W tile is {64 rows of the contraction index} x {BN columns}, MN-major for the MMA
  int w_group_rows;       // w_mn: first W row of group g = w_row0[0] + g * w_group_rows
  int out_col0;           // first output column (EPI_HIDDEN / EPI_DGRAD: out col = out_col0 + n_tile * BN)
  const __nv_bfloat16* mask_act;   // EPI_DGRAD: saved post-LeakyReLU activation (hi part) at the output position
  int ld_mask;
  int mask_col0;
  // EPI_DU: fp32 accumulation of the MLP-input gradient into the state gradient and the conditioning gradient
  float* du_state; int du_ld_state; int du_state_off; int du_x_cols;
  float* du_cond;  int du_ld_cond;  int du_cond_col0; int du_E;
};

__host__ __device__ constexpr int gemm_nsub(int mode) { return (mode == EPI_TAIL_FWD || mode == EPI_TAIL_REV) ? 2 : 1; }

__host__ inline int gemm_stage_bytes(int ksplit, int mode, int BN) {
  return gemm_nsub(mode) * ksplit * (A_TILE_BYTES + BN * BK * 2);
}
// epilogue smem: [slabs (EPI_HIDDEN with TMA store)] [fp32 area: bias (2 x BN) | tail: keep tile + 2 biases +
// 3 per-column arrays of the folded boundary]
// (`direct` tails keep their rows in registers: no 128 x BN staging tile, which leaves room for one more pipeline stage)
__host__ inline int gemm_epi_bytes(int ksplit, int mode, int BN, int slab_bufs, int direct = 0) {
  if (mode == EPI_HIDDEN || mode == EPI_DGRAD) return ksplit * slab_bufs * SLAB_BYTES + 2 * BN * 4;
  if (mode == EPI_DU) return 2 * BN * 4;
  return ((direct ? 0 : BM * (BN + 1)) + 5 * BN) * 4;
}
__host__ inline int gemm_smem_bytes(int ksplit, int mode, int BN, int stages, int slab_bufs, int direct = 0) {
  return 1024 + stages * gemm_stage_bytes(ksplit, mode, BN) +
         ((gemm_epi_bytes(ksplit, mode, BN, slab_bufs, direct) + 15) & ~15) + SMEM_BARRIER_BYTES;
}

// kWarpSlab: instantiate the warp-private-slab epilogue (p.slab_warp; multi-wave hidden / dgrad launches).  Small-batch
// launches use the kWarpSlab = false instantiation: carrying the unused path cost them 130 ns per launch (r01 A/B).
template <int kSplit, int kMode, bool kWarpSlab = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
inn_gemm_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                const GemmParams p) {
  constexpr int NSUB = gemm_nsub(kMode);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  // warp index through a shuffle and one-lane roles through elect.sync: the compiler then knows the role code is
  // warp-uniform and keeps TMA / MMA operands in uniform registers (a plain `lane == 0` branch made it wrap every
  // UTCHMMA / UTMALDG in an ELECT + R2UR loop, ~90 ns per MMA: r01 trace)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int BN = p.BN;
  // debug timeline: first and last CTA only
  unsigned long long* const tr =
      (p.trace != nullptr && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? p.trace + (blockIdx.x == 0 ? 0 : 16) : nullptr;
  if (tr != nullptr && threadIdx.x == 0) tr[0] = globaltimer_ns();
  const int w_tile_bytes = BN * BK * 2;
  const int stage_bytes = NSUB * kSplit * (A_TILE_BYTES + w_tile_bytes);
  const int stages = p.stages;
  constexpr int kacc = 1;                    // K-interleaved accumulators: measured no gain (r01) -- compiled out
  const int sub_cols = kacc * BN;            // TMEM columns of one accumulator set (s or t)
  const int acc_cols = NSUB * sub_cols;
  const int nacc = p.nacc;
  constexpr int mc = 1;                      // TMA-multicast clusters along n: measured slower at one wave (r01) -- compiled out
  const uint32_t cta_rank = mc > 1 ? cluster_ctarank() : 0u;
  const uint16_t mc_mask = static_cast<uint16_t>((1u << mc) - 1u);

  uint8_t* epi_smem = smem + p.epi_off;
  constexpr bool kSlabMode = (kMode == EPI_HIDDEN || kMode == EPI_DGRAD);
  constexpr bool kTailMode = (kMode == EPI_TAIL_FWD || kMode == EPI_TAIL_REV);
  const int slab_total = kSlabMode ? kSplit * p.slab_bufs * SLAB_BYTES : 0;
  float* epi_f32 = reinterpret_cast<float*>(epi_smem + slab_total);
  const int keep_elems = (kTailMode && !p.direct) ? BM * (BN + 1) : 0;   // tail staging tile [128][BN + 1]
  const int epi_bytes = slab_total + (kTailMode ? (keep_elems + 5 * BN) * 4 : 2 * BN * 4);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_smem + ((epi_bytes + 15) & ~15));
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
    if (kSlabMode && p.slab_bufs > 0) {
      tma_prefetch_desc(&tmO0);
      if (kSplit == 2) tma_prefetch_desc(&tmO1);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], mc);          // every CTA sharing the multicast slot must release it
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], EPI_THREADS);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  if (mc > 1) cluster_sync_all();            // peers' barriers are initialised before anyone multicasts
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tr != nullptr && threadIdx.x == 0) tr[1] = globaltimer_ns();

  // PDL: let the next kernel start its prologue (and prefetch its weights) now; our own reads of the previous
  // kernel's outputs happen only after grid_dep_wait() in the roles below.
  if (threadIdx.x == 0) grid_dep_launch();

  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0) {
    // ============================ TMA producer ============================
    int stage = 0;
    uint32_t phase = 0;
    const int a_rows = BM / mc;                               // rows of the A box this CTA loads (and multicasts)
    // weights do not depend on the previous kernel: arm the first stages and fetch their W tiles before the wait
    int pre = 0;
    if (p.w_prefetch && static_cast<int>(blockIdx.x) < num_tiles) {
      const int n_tile0 = blockIdx.x % p.num_n_tiles;
      pre = p.num_k_blocks < stages ? p.num_k_blocks : stages;
      for (int ki = 0; ki < pre; ++ki) {
        const int kb = ki;
        mbar_arrive_expect_tx_el(&full_bar[ki], stage_bytes);
        uint8_t* sbase = smem + ki * stage_bytes;
#pragma unroll
        for (int sub = 0; sub < NSUB; ++sub)
#pragma unroll
          for (int sp = 0; sp < kSplit; ++sp)
            tma_load_2d_el(sbase + NSUB * kSplit * A_TILE_BYTES + (sub * kSplit + sp) * w_tile_bytes,
                        sp == 0 ? &tmW0 : &tmW1, &full_bar[ki], kb * BK, p.w_row0[sub] + n_tile0 * BN);
      }
    }
    // tile coordinates of the first tile (integer divisions) before the dependency wait, off the critical path
    int m_tile = static_cast<int>(blockIdx.x) / p.num_n_tiles;
    int n_tile = static_cast<int>(blockIdx.x) - m_tile * p.num_n_tiles;
    int group = n_tile / p.tiles_per_group;
    int a_row0 = m_tile * BM;
    if (p.a_m_wrap > 0) a_row0 = (m_tile % p.a_m_wrap) * BM;
    grid_dep_wait();
    if (tr != nullptr && lane == 0) tr[2] = globaltimer_ns();
    bool first_tile = true;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      if (!first_tile) {
        m_tile = tile / p.num_n_tiles;
        n_tile = tile - m_tile * p.num_n_tiles;
        group = n_tile / p.tiles_per_group;
        a_row0 = m_tile * BM;
        if (p.a_m_wrap > 0) a_row0 = (m_tile % p.a_m_wrap) * BM;
      }
      for (int ki = 0; ki < p.num_k_blocks; ++ki) {
        const int kb = ki;
        {
          uint8_t* sbase = smem + stage * stage_bytes;
          const bool w_done = first_tile && ki < pre;
          if (!w_done) {
            mbar_wait(&empty_bar[stage], phase ^ 1, 1);
            mbar_arrive_expect_tx_el(&full_bar[stage], stage_bytes);
          }
#pragma unroll
          for (int sub = 0; sub < NSUB; ++sub) {
            const int a_k = p.a_off[sub] + group * p.a_group_stride + kb * BK;
            const int w_r = p.w_row0[sub] + n_tile * BN;
#pragma unroll
            for (int sp = 0; sp < kSplit; ++sp) {
              const int t = sub * kSplit + sp;
              if (mc > 1)
                tma_load_2d_mc_el(sbase + t * A_TILE_BYTES + cta_rank * a_rows * 128, sp == 0 ? &tmA0 : &tmA1,
                               &full_bar[stage], a_k, a_row0 + cta_rank * a_rows, mc_mask);
              else
                tma_load_2d_el(sbase + t * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, &full_bar[stage], a_k, a_row0);
              if (p.w_mn) {
                // contraction index = W row: {64 rows} x {BN columns} as BN/64 MN-major chunks
                const int w_c = (n_tile - group * p.tiles_per_group) * BN;
                const int w_rk = p.w_row0[sub] + group * p.w_group_rows + kb * BK;
                for (int c = 0; c < (BN >> 6); ++c)
                  tma_load_2d_el(sbase + NSUB * kSplit * A_TILE_BYTES + t * w_tile_bytes + c * (64 * 128),
                              sp == 0 ? &tmW0 : &tmW1, &full_bar[stage], w_c + c * 64, w_rk);
              } else if (!w_done)
                tma_load_2d_el(sbase + NSUB * kSplit * A_TILE_BYTES + t * w_tile_bytes, sp == 0 ? &tmW0 : &tmW1,
                            &full_bar[stage], kb * BK, w_r);
            }
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
      first_tile = false;
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ==============================
    const uint32_t idesc = umma_idesc_bf16(BM, BN, 0, p.w_mn);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 2);
      tc_fence_after();
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        {
          mbar_wait(&full_bar[stage], phase, 3);
          tc_fence_after();
          if (tr != nullptr && lane == 0 && it == 0 && kb == 0) tr[3] = globaltimer_ns();
          const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
#pragma unroll
          for (int sub = 0; sub < NSUB; ++sub) {
            const uint32_t d_base = tmem_base + acc * acc_cols + sub * sub_cols;
            const uint32_t a0 = sbase + (sub * kSplit) * A_TILE_BYTES;
            const uint32_t w0 = sbase + NSUB * kSplit * A_TILE_BYTES + (sub * kSplit) * w_tile_bytes;
            const uint64_t da_hi = umma_desc_k_sw128(a0);
            const uint32_t w1 = w0 + (kSplit - 1) * w_tile_bytes;
            const uint64_t dw_hi = p.w_mn ? umma_desc_mn_sw128(w0, 64 * 128) : umma_desc_k_sw128(w0);
            const uint64_t da_lo = umma_desc_k_sw128(a0 + (kSplit - 1) * A_TILE_BYTES);
            const uint64_t dw_lo = p.w_mn ? umma_desc_mn_sw128(w1, 64 * 128) : umma_desc_k_sw128(w1);
            // per UMMA_K = 16: K-major operands advance 32 B inside the swizzle row, MN-major ones 16 rows x 128 B
            const uint64_t wstep = p.w_mn ? static_cast<uint64_t>((UMMA_K * 128) >> 4) : static_cast<uint64_t>((UMMA_K * 2) >> 4);
#pragma unroll
            for (int kk = 0; kk < BK / UMMA_K; ++kk) {
              const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 2) >> 4);
              const uint64_t woff = wstep * kk;
              // round-robin over the kacc accumulators; the first MMA into each one overwrites
              const int mi = (kb * (BK / UMMA_K) + kk) * (kSplit == 2 ? 3 : 1);
              if (kSplit == 2) {
                umma_bf16_ss_el(d_base + ((mi) & (kacc - 1)) * BN, da_lo + koff, dw_hi + woff, idesc, mi >= kacc ? 1u : 0u);
                umma_bf16_ss_el(d_base + ((mi + 1) & (kacc - 1)) * BN, da_hi + koff, dw_lo + woff, idesc, mi + 1 >= kacc ? 1u : 0u);
                umma_bf16_ss_el(d_base + ((mi + 2) & (kacc - 1)) * BN, da_hi + koff, dw_hi + woff, idesc, mi + 2 >= kacc ? 1u : 0u);
              } else {
                umma_bf16_ss_el(d_base + (mi & (kacc - 1)) * BN, da_hi + koff, dw_hi + woff, idesc, mi >= kacc ? 1u : 0u);
              }
            }
          }
          if (mc > 1) umma_commit_mc_el(&empty_bar[stage], mc_mask);   // slot is shared: release it in every peer
          else umma_commit_el(&empty_bar[stage]);                 // frees the smem slot when the MMAs retire
          if (kb == p.num_k_blocks - 1) {
            umma_commit_el(&tfull_bar[acc]);   // accumulators ready
            if (tr != nullptr && lane == 0) tr[4] = globaltimer_ns();
          }
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    grid_dep_wait();
#include "gemm_epilogue.inl"
  }

  tc_fence_before();
  if (mc > 1) cluster_sync_all();            // no CTA exits while peers may still multicast into it
  else __syncthreads();
  if (tr != nullptr && threadIdx.x == 0) tr[7] = globaltimer_ns();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace inn
