// Persistent, warp-specialised tcgen05 GEMM for the INN subnetworks (sm_100a).
//
//   acc[sub] (TMEM, fp32) = A_sub[128 x K] (bf16, K-major, TMA/SW128)  x  W_sub[BN x K]^T (bf16, K-major)
//
// One kernel template serves every Linear of BasicFullyConnectedNet (reference
// models/modules/INN/modules.py:165-191) and fuses what follows it:
//   EPI_HIDDEN    bias + LeakyReLU(0.01) -> bf16 activations (optionally hi/lo split)   [modules.py:173-180]
//   EPI_TAIL_FWD  last Linear of s and t together (two accumulators), tanh, exp, affine update of the kept
//                 half in place, per-row log-det sum                                      [flow_blocks.py:138-142]
//   EPI_TAIL_REV  same GEMMs, (x - t) * exp(-s)                                           [flow_blocks.py:151]
// kSplit = 2 is the fp32-parity mode: activations and weights are carried as bf16 hi + bf16 lo and each
// k-step issues hi*hi + lo*hi + hi*lo (3 MMAs, fp32 accumulate), dropping only the lo*lo term (~2^-16 rel).
//
// Warp roles (256 threads, 1 CTA / SM): warp 0 TMA producer, warp 1 MMA issuer (one lane), warp 2 TMEM
// allocator, warps 4-7 epilogue (TMEM lane quarter = warp % 4). Rings: smem full/empty (TMA <-> MMA),
// TMEM full/empty (MMA <-> epilogue, double-buffered when the accumulators fit twice in 512 columns).
#pragma once
#include "ptx.cuh"

namespace inn {

constexpr int BM = 128;                 // UMMA M (rows of the batch per tile)
constexpr int BK = 64;                  // bf16 elements per k-block = one 128-B swizzle row
constexpr int UMMA_K = 16;
constexpr int A_TILE_BYTES = BM * BK * 2;
constexpr int MAX_STAGES = 8;
constexpr int GEMM_THREADS = 256;
constexpr int TMEM_COLS = 512;
constexpr int SMEM_BARRIER_BYTES = 256;

enum EpiMode { EPI_HIDDEN = 0, EPI_TAIL_FWD = 1, EPI_TAIL_REV = 2, EPI_DGRAD = 3 };

struct GemmParams {
  int M;                  // valid rows
  int num_m_tiles;
  int num_n_tiles;
  int num_k_blocks;       // K / 64 (K zero-padded)
  int BN;                 // UMMA N of one accumulator: multiple of 16, <= 256
  int stages;
  int tiles_per_group;    // n-tiles per subnetwork group (hidden layers); >= num_n_tiles otherwise
  int a_group_stride;     // A column offset (elements) per group
  int a_off[2];           // A column offset per accumulator
  int w_row0[2];          // first W row per accumulator
  const float* bias;      // hidden: [N]; tail: s bias
  const float* bias2;     // tail: t bias
  // EPI_HIDDEN / EPI_DGRAD
  __nv_bfloat16* out_hi;
  __nv_bfloat16* out_lo;
  int ld_out;
  float slope;
  const __nv_bfloat16* mask_src;  // EPI_DGRAD: forward activation whose sign gates the gradient (nullable)
  int ld_mask;
  // EPI_TAIL_*
  float* state;           // [M, ld_state] fp32, updated in place at columns keep_off + [0, n_valid)
  int ld_state;
  int keep_off;
  int n_valid;            // true C/2
  float* logdet;          // [M], forward only
  int logdet_atomic;      // >1 n-tile per row -> atomicAdd
  __nv_bfloat16* u_hi;    // next half-step's MLP input (bf16 copy of the new half), nullable
  __nv_bfloat16* u_lo;
  int ld_u;
  float* s_save;          // training: tanh output s and exp(s) are recomputed in backward from these (nullable)
  float* t_save;
  int ld_save;
};

__host__ __device__ constexpr int gemm_nsub(int mode) { return (mode == EPI_HIDDEN || mode == EPI_DGRAD) ? 1 : 2; }

__host__ inline int gemm_stage_bytes(int ksplit, int mode, int BN) {
  return gemm_nsub(mode) * ksplit * (A_TILE_BYTES + BN * BK * 2);
}
__host__ inline int gemm_smem_bytes(int ksplit, int mode, int BN, int stages) {
  return 1024 + stages * gemm_stage_bytes(ksplit, mode, BN) + SMEM_BARRIER_BYTES;
}

template <int kSplit, int kMode>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
inn_gemm_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmW1,
                const GemmParams p) {
  constexpr int NSUB = gemm_nsub(kMode);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int BN = p.BN;
  const int w_tile_bytes = BN * BK * 2;
  const int stage_bytes = NSUB * kSplit * (A_TILE_BYTES + w_tile_bytes);
  const int stages = p.stages;
  const int acc_cols = NSUB * BN;
  const int nacc = (2 * acc_cols <= TMEM_COLS) ? 2 : 1;

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
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 128);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0) {
    // ============================ TMA producer ============================
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_tile = tile / p.num_n_tiles;
      const int n_tile = tile - m_tile * p.num_n_tiles;
      const int group = n_tile / p.tiles_per_group;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        if (lane == 0) {
          mbar_wait(&empty_bar[stage], phase ^ 1, 1);
          mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
          uint8_t* sbase = smem + stage * stage_bytes;
#pragma unroll
          for (int sub = 0; sub < NSUB; ++sub) {
            const int a_k = p.a_off[sub] + group * p.a_group_stride + kb * BK;
            const int w_r = p.w_row0[sub] + n_tile * BN;
#pragma unroll
            for (int sp = 0; sp < kSplit; ++sp) {
              const int t = sub * kSplit + sp;
              tma_load_2d(sbase + t * A_TILE_BYTES, sp == 0 ? &tmA0 : &tmA1, &full_bar[stage], a_k, m_tile * BM);
              tma_load_2d(sbase + NSUB * kSplit * A_TILE_BYTES + t * w_tile_bytes, sp == 0 ? &tmW0 : &tmW1,
                          &full_bar[stage], kb * BK, w_r);
            }
          }
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ==============================
    const uint32_t idesc = umma_idesc_bf16(BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it % nacc;
      const uint32_t acc_phase = (it / nacc) & 1;
      if (lane == 0) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 2);
        tc_fence_after();
      }
      __syncwarp();
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        if (lane == 0) {
          mbar_wait(&full_bar[stage], phase, 3);
          tc_fence_after();
          const uint32_t sbase = smem_u32(smem + stage * stage_bytes);
#pragma unroll
          for (int sub = 0; sub < NSUB; ++sub) {
            const uint32_t d_tmem = tmem_base + acc * acc_cols + sub * BN;
            const uint32_t a0 = sbase + (sub * kSplit) * A_TILE_BYTES;
            const uint32_t w0 = sbase + NSUB * kSplit * A_TILE_BYTES + (sub * kSplit) * w_tile_bytes;
            const uint64_t da_hi = umma_desc_k_sw128(a0);
            const uint64_t dw_hi = umma_desc_k_sw128(w0);
            const uint64_t da_lo = umma_desc_k_sw128(a0 + (kSplit - 1) * A_TILE_BYTES);
            const uint64_t dw_lo = umma_desc_k_sw128(w0 + (kSplit - 1) * w_tile_bytes);
#pragma unroll
            for (int kk = 0; kk < BK / UMMA_K; ++kk) {
              const uint64_t koff = static_cast<uint64_t>((kk * UMMA_K * 2) >> 4);
              const uint32_t first = (kb == 0 && kk == 0) ? 0u : 1u;
              if (kSplit == 2) {
                // small terms first, then the dominant hi*hi product
                umma_bf16_ss(d_tmem, da_lo + koff, dw_hi + koff, idesc, first);
                umma_bf16_ss(d_tmem, da_hi + koff, dw_lo + koff, idesc, 1u);
                umma_bf16_ss(d_tmem, da_hi + koff, dw_hi + koff, idesc, 1u);
              } else {
                umma_bf16_ss(d_tmem, da_hi + koff, dw_hi + koff, idesc, first);
              }
            }
          }
          umma_commit(&empty_bar[stage]);                      // frees the smem slot when the MMAs retire
          if (kb == p.num_k_blocks - 1) umma_commit(&tfull_bar[acc]);   // accumulators ready
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ============================ epilogue ================================
    const int q = warp - 4;                       // TMEM lane quarter (== warp % 4)
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m_tile = tile / p.num_n_tiles;
      const int n_tile = tile - m_tile * p.num_n_tiles;
      const int acc = it % nacc;
      const uint32_t acc_phase = (it / nacc) & 1;
      mbar_wait(&tfull_bar[acc], acc_phase, 4);
      tc_fence_after();
      const int row = m_tile * BM + q * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * acc_cols;

      if (kMode == EPI_HIDDEN || kMode == EPI_DGRAD) {
        for (int c0 = 0; c0 < BN; c0 += 16) {
          float v[16];
          tmem_ld16(t_row + c0, v);
          tmem_ld_wait();
          const int col = n_tile * BN + c0;
          if (kMode == EPI_HIDDEN) {
            const float4* b4 = reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 b = __ldg(b4 + j);
              v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = v[j] >= 0.f ? v[j] : v[j] * p.slope;
          } else if (p.mask_src != nullptr && row_ok) {
            // dgrad: multiply by LeakyReLU'(pre-activation); sign(pre) == sign(saved activation)
            const uint4* m4 = reinterpret_cast<const uint4*>(p.mask_src + static_cast<size_t>(row) * p.ld_mask + col);
            uint4 mm[2] = {__ldg(m4), __ldg(m4 + 1)};
            const uint16_t* mh = reinterpret_cast<const uint16_t*>(mm);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = (mh[j] & 0x8000u) ? v[j] * p.slope : v[j];
          }
          if (row_ok) {
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              __nv_bfloat16 h0, l0, h1, l1;
              split_bf16(v[2 * j], h0, l0);
              split_bf16(v[2 * j + 1], h1, l1);
              hi[j] = pack_bf16x2(h0, h1);
              lo[j] = pack_bf16x2(l0, l1);
            }
            uint4* o = reinterpret_cast<uint4*>(p.out_hi + static_cast<size_t>(row) * p.ld_out + col);
            o[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            o[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
            if (kSplit == 2) {
              uint4* ol = reinterpret_cast<uint4*>(p.out_lo + static_cast<size_t>(row) * p.ld_out + col);
              ol[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
              ol[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
            }
          }
        }
      } else {
        float ld_sum = 0.f;
        for (int c0 = 0; c0 < BN; c0 += 16) {
          float vs[16], vt[16];
          tmem_ld16(t_row + c0, vs);
          tmem_ld16(t_row + BN + c0, vt);
          tmem_ld_wait();
          const int col0 = n_tile * BN + c0;       // column inside the transformed half
          if (row_ok) {
            float* st = p.state + static_cast<size_t>(row) * p.ld_state + p.keep_off;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int col = col0 + j;
              if (col < p.n_valid) {
                const float s = tanhf(vs[j] + __ldg(p.bias + col));
                const float t = vt[j] + __ldg(p.bias2 + col);
                const float keep = st[col];
                float nw;
                if (kMode == EPI_TAIL_FWD) {
                  nw = keep * expf(s) + t;
                  ld_sum += s;
                } else {
                  nw = (keep - t) * expf(-s);
                }
                st[col] = nw;
                if (p.s_save != nullptr) {
                  p.s_save[static_cast<size_t>(row) * p.ld_save + col] = s;
                  p.t_save[static_cast<size_t>(row) * p.ld_save + col] = keep;   // pre-update kept half
                }
                if (p.u_hi != nullptr) {
                  __nv_bfloat16 h, l;
                  split_bf16(nw, h, l);
                  p.u_hi[static_cast<size_t>(row) * p.ld_u + col] = h;
                  if (kSplit == 2) p.u_lo[static_cast<size_t>(row) * p.ld_u + col] = l;
                }
              }
            }
          }
        }
        if (kMode == EPI_TAIL_FWD && row_ok) {
          if (p.logdet_atomic) atomicAdd(p.logdet + row, ld_sum);
          else p.logdet[row] += ld_sum;
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

}  // namespace inn
