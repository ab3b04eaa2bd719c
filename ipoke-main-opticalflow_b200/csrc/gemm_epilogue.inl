// Epilogue warps (4..7) of inn_gemm_kernel -- textually included inside the kernel body.
// Two threads own one accumulator row (TMEM lane): warps 4-7 take the even 16/32-column groups, warps 8-11 the
// odd ones; 256 threads cover the 128-row tile.
    // ============================ epilogue ================================
    const int q = warp & 3;                       // TMEM lane quarter (== warp % 4, a hardware restriction)
    const int ch = (warp - 4) >> 2;               // column half handled by this warp
    const int tid_e = threadIdx.x - 128;          // 0..255
    const int r_loc = q * 32 + lane;              // row inside the tile
    int it = 0;
    int slab_ctr = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m_tile = tile / p.num_n_tiles;
      const int n_tile = tile - m_tile * p.num_n_tiles;
      const int acc = (nacc == 2) ? (it & 1) : 0;
      const uint32_t acc_phase = ((nacc == 2) ? (it >> 1) : it) & 1;
      const int row = m_tile * BM + r_loc;
      const bool row_ok = row < p.M;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * acc_cols;

      if (kSlabMode) {
        // ---- bias slice -> smem while the MMAs of this tile run
        float* bias_s = epi_f32 + (it & 1) * BN;
        const int ocol0 = p.out_col0 + n_tile * BN;              // first output column of this tile
        if (kMode == EPI_HIDDEN) {
          for (int c = tid_e; c < BN; c += EPI_THREADS) bias_s[c] = __ldg(p.bias + n_tile * BN + c);
          named_bar_sync(1, EPI_THREADS);
        }
        // EPI_DGRAD: LeakyReLU'(pre) from the sign of the saved activation at the output position (modules.py:174)
        const __nv_bfloat16* mask_row =
            (kMode == EPI_DGRAD && row_ok) ? p.mask_act + static_cast<size_t>(row) * p.ld_mask + p.mask_col0 + ocol0 : nullptr;
        mbar_wait(&tfull_bar[acc], acc_phase, 4);
        tc_fence_after();
        if (tr != nullptr && tid_e == 0) tr[5] = globaltimer_ns();
        if (kWarpSlab && p.slab_warp) {
          // ---- warp-private slabs: every epilogue warp owns the 32 rows of its TMEM lane quarter: TMEM -> regs -> its own
          // swizzled smem buffer -> one TMA store per warp.  No CTA-wide barrier inside the tile: the warps drift apart
          // and overlap each other's TMEM-load and store-drain latencies (ncu, r01 v6: the CTA-wide version with one
          // slab buffer serialises drain -> barrier -> load -> math -> barrier -> store, 1.3 us per slab).
          //   slab_warp 1 (slab_bufs >= 2): 32 x 128 B buffers, 64-column boxes, warps 4-7 take the even slabs, 8-11 the odd
          //   slab_warp 2 (slab_bufs == 1): 32 x  64 B buffers, 32-column boxes (64-B swizzle), warp `ch` takes its half
          const int we = warp - 4;
          const bool wide = p.slab_warp == 1;
          const int nbw = wide ? (p.slab_bufs >> 1) : 1;         // buffers per warp
          const int wbytes = wide ? 4096 : 2048;
          const int n_slabs = BN >> 6;
          bool arrived = false;
          for (int s = wide ? ch : 0; s < n_slabs; s += wide ? 2 : 1, ++slab_ctr) {
            uint8_t* wb_hi = epi_smem + ((we * nbw + (slab_ctr & (nbw - 1))) * kSplit) * wbytes + lane * (wide ? 128 : 64);
            uint8_t* wb_lo = wb_hi + wbytes;
            for (int h = 0; h < (wide ? 2 : 1); ++h) {
              const int c_in = wide ? h * 32 : ch * 32;          // first column of this half inside the slab
              float v[32];
#pragma unroll
              for (int c = 0; c < 2; ++c) tmem_ld16(t_row + s * 64 + c_in + c * 16, v + c * 16);
              tmem_ld_wait();
              for (int a = 1; a < kacc; ++a) {
                float w[32];
#pragma unroll
                for (int c = 0; c < 2; ++c) tmem_ld16(t_row + a * BN + s * 64 + c_in + c * 16, w + c * 16);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 32; ++c) v[c] += w[c];
              }
              if (wide ? (h == 1 && s + 2 >= n_slabs) : (s + 1 >= n_slabs)) {   // this warp has read its part of the accumulator
                tc_fence_before();
                mbar_arrive(&tempty_bar[acc]);
                arrived = true;
              }
              if (h == 0) {                                      // the store that last read this buffer has drained
                if (lane == 0) {
                  if (nbw == 2) bulk_wait_read<1>();
                  else bulk_wait_read<0>();
                }
                __syncwarp();
              }
              const float* bs = bias_s + s * 64 + c_in;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint32_t hi[4], lo[4];
                if (kMode == EPI_HIDDEN) {
                  const float4 b0 = *reinterpret_cast<const float4*>(bs + j * 8);
                  const float4 b1 = *reinterpret_cast<const float4*>(bs + j * 8 + 4);
                  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    float a = v[j * 8 + 2 * e] + bb[2 * e];
                    float b = v[j * 8 + 2 * e + 1] + bb[2 * e + 1];
                    a = fmaxf(a, a * p.slope);                   // LeakyReLU for 0 <= slope <= 1 (checked at plan creation)
                    b = fmaxf(b, b * p.slope);
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
                } else {
                  uint4 mk = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
                  if (mask_row != nullptr) mk = __ldg(reinterpret_cast<const uint4*>(mask_row + s * 64 + c_in) + j);
                  const uint32_t mw[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    const float a = v[j * 8 + 2 * e] * (bf16_bits_positive(mw[e] & 0xFFFFu) ? 1.f : p.slope);
                    const float b = v[j * 8 + 2 * e + 1] * (bf16_bits_positive(mw[e] >> 16) ? 1.f : p.slope);
                    __nv_bfloat16 h0, l0, h1, l1;
                    split_bf16(a, h0, l0);
                    split_bf16(b, h1, l1);
                    hi[e] = pack_bf16x2(h0, h1);
                    lo[e] = pack_bf16x2(l0, l1);
                  }
                }
                // 128-B swizzle: 16-B chunk index XOR (row % 8); 64-B swizzle: XOR ((row / 2) % 4)
                const int chunk = (wide ? ((h * 4 + j) ^ (lane & 7)) : (j ^ ((lane >> 1) & 3))) << 4;
                *reinterpret_cast<uint4*>(wb_hi + chunk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                if (kSplit == 2) *reinterpret_cast<uint4*>(wb_lo + chunk) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
              }
            }
            fence_proxy_async();                                 // generic-proxy smem writes -> visible to TMA
            __syncwarp();
            if (lane == 0) {
              const int col = ocol0 + s * 64 + (wide ? 0 : ch * 32);
              const int loff = lane * (wide ? 128 : 64);
              tma_store_2d(&tmO0, wb_hi - loff, col, m_tile * BM + q * 32);
              if (kSplit == 2) tma_store_2d(&tmO1, wb_lo - loff, col, m_tile * BM + q * 32);
              bulk_commit();
            }
          }
          if (!arrived) {                                        // BN = 64: the odd-slab warps read nothing
            tc_fence_before();
            mbar_arrive(&tempty_bar[acc]);
          }
        } else if (p.slab_bufs > 0) {
          // ---- 64-column slabs: TMEM -> regs -> swizzled smem -> TMA store
          const int n_slabs = BN >> 6;
          for (int s = 0; s < n_slabs; ++s, ++slab_ctr) {
            const int buf = slab_ctr & (p.slab_bufs - 1);        // slab_bufs is 1, 2 or 4
            if (tid_e == 0) {
              if (p.slab_bufs == 4) bulk_wait_read<3>();
              else if (p.slab_bufs == 2) bulk_wait_read<1>();
              else bulk_wait_read<0>();
            }
            named_bar_sync(1, EPI_THREADS);                      // slab buffer `buf` is free again
            float v[32];                                         // this thread's half of the 64-column slab row
#pragma unroll
            for (int c = 0; c < 2; ++c) tmem_ld16(t_row + s * 64 + ch * 32 + c * 16, v + c * 16);
            tmem_ld_wait();
            for (int a = 1; a < kacc; ++a) {                     // K-interleaved partial accumulators
              float w[32];
#pragma unroll
              for (int c = 0; c < 2; ++c) tmem_ld16(t_row + a * BN + s * 64 + ch * 32 + c * 16, w + c * 16);
              tmem_ld_wait();
#pragma unroll
              for (int c = 0; c < 32; ++c) v[c] += w[c];
            }
            if (s == n_slabs - 1) {                              // accumulator fully read: hand TMEM back
              tc_fence_before();
              mbar_arrive(&tempty_bar[acc]);
            }
            const float* bs = bias_s + s * 64 + ch * 32;
            uint8_t* slab_hi = epi_smem + (buf * kSplit) * SLAB_BYTES + r_loc * 128;
            uint8_t* slab_lo = slab_hi + SLAB_BYTES;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint32_t hi[4], lo[4];
              uint4 mk = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);   // bf16 ones: mask = 1
              if (kMode == EPI_DGRAD && mask_row != nullptr)
                mk = __ldg(reinterpret_cast<const uint4*>(mask_row + s * 64 + ch * 32) + j);
              const uint32_t mw[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float a, b;
                if (kMode == EPI_HIDDEN) {
                  a = v[j * 8 + 2 * e] + bs[j * 8 + 2 * e];
                  b = v[j * 8 + 2 * e + 1] + bs[j * 8 + 2 * e + 1];
                  a = a >= 0.f ? a : a * p.slope;
                  b = b >= 0.f ? b : b * p.slope;
                } else {
                  a = v[j * 8 + 2 * e] * (bf16_bits_positive(mw[e] & 0xFFFFu) ? 1.f : p.slope);
                  b = v[j * 8 + 2 * e + 1] * (bf16_bits_positive(mw[e] >> 16) ? 1.f : p.slope);
                }
                __nv_bfloat16 h0, l0, h1, l1;
                split_bf16(a, h0, l0);
                split_bf16(b, h1, l1);
                hi[e] = pack_bf16x2(h0, h1);
                lo[e] = pack_bf16x2(l0, l1);
              }
              const int chunk = ((ch * 4 + j) ^ (r_loc & 7)) << 4;   // 128-B swizzle: 16-B chunk index XOR (row % 8)
              *reinterpret_cast<uint4*>(slab_hi + chunk) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
              if (kSplit == 2) *reinterpret_cast<uint4*>(slab_lo + chunk) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            fence_proxy_async();                                 // generic-proxy smem writes -> visible to TMA
            named_bar_sync(1, EPI_THREADS);
            if (tid_e == 0) {
              const int col = ocol0 + s * 64;
              tma_store_2d(&tmO0, epi_smem + (buf * kSplit) * SLAB_BYTES, col, m_tile * BM);
              if (kSplit == 2) tma_store_2d(&tmO1, epi_smem + (buf * kSplit + 1) * SLAB_BYTES, col, m_tile * BM);
              bulk_commit();
            }
          }
        } else {
          // ---- narrow tiles (BN < 64, small batches): direct 32-B global stores per thread
          for (int c0 = ch * 16; c0 < BN; c0 += 32) {
            float v[16];
            tmem_ld16(t_row + c0, v);
            tmem_ld_wait();
            for (int a = 1; a < kacc; ++a) {
              float w[16];
              tmem_ld16(t_row + a * BN + c0, w);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += w[j];
            }
            const int col = ocol0 + c0;
            if (kMode == EPI_HIDDEN) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                v[j] += bias_s[c0 + j];
                v[j] = v[j] >= 0.f ? v[j] : v[j] * p.slope;
              }
            } else if (mask_row != nullptr) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                v[j] *= bf16_bits_positive(__bfloat16_as_ushort(mask_row[c0 + j])) ? 1.f : p.slope;
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
          tc_fence_before();
          mbar_arrive(&tempty_bar[acc]);
        }
      } else if (kMode == EPI_DU) {
        // ---- gradient of the MLP input: columns [0, x_cols) add into the state gradient of the conditioner
        //      half, columns [cond_col0, cond_col0 + E) into the conditioning gradient (flow_blocks.py:137, 224)
        mbar_wait(&tfull_bar[acc], acc_phase, 4);
        tc_fence_after();
        if (tr != nullptr && tid_e == 0) tr[5] = globaltimer_ns();
        for (int c0 = ch * 16; c0 < BN; c0 += 32) {
          float v[16];
          tmem_ld16(t_row + c0, v);
          tmem_ld_wait();
          // whole 16-column group inside one destination and 16-byte aligned: 4 float4 read-modify-writes in flight
          // (the scalar += chain below is one dependent load-add-store per element: 57 us per launch at cfg A / 4096)
          float* dst16 = nullptr;
          if (row_ok) {
            const int cb = n_tile * BN + c0;
            if (cb + 16 <= p.du_x_cols)
              dst16 = p.du_state + static_cast<size_t>(row) * p.du_ld_state + p.du_state_off + cb;
            else if (cb >= p.du_cond_col0 && cb + 16 <= p.du_cond_col0 + p.du_E)
              dst16 = p.du_cond + static_cast<size_t>(row) * p.du_ld_cond + (cb - p.du_cond_col0);
            if ((reinterpret_cast<uintptr_t>(dst16) & 15u) != 0) dst16 = nullptr;
          }
          if (dst16 != nullptr) {
            float4* d4 = reinterpret_cast<float4*>(dst16);
            float4 a4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) a4[j] = d4[j];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              d4[j] = make_float4(a4[j].x + v[4 * j], a4[j].y + v[4 * j + 1], a4[j].z + v[4 * j + 2], a4[j].w + v[4 * j + 3]);
          } else if (row_ok) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int c = n_tile * BN + c0 + j;
              if (c < p.du_x_cols) {
                p.du_state[static_cast<size_t>(row) * p.du_ld_state + p.du_state_off + c] += v[j];
              } else if (c >= p.du_cond_col0 && c < p.du_cond_col0 + p.du_E) {
                p.du_cond[static_cast<size_t>(row) * p.du_ld_cond + (c - p.du_cond_col0)] += v[j];
              }
            }
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty_bar[acc]);
      } else {
        // ---- coupling tail: stage the kept half of this tile and both biases in smem (coalesced), overlapping
        //      the MMAs; then one row per thread out of smem; then coalesced write-back.
        float* keep_s = epi_f32;                                  // [128][BN + 1]
        float* bias_s = epi_f32 + keep_elems;
        float* bias_t = bias_s + BN;
        float* col_scale = bias_t + BN;                           // folded boundary: per-column constants of this tile
        float* col_loc = col_scale + BN;
        int* col_dst = reinterpret_cast<int*>(col_loc + BN);
        const int ldk = BN + 1;
        const int col_base = n_tile * BN;
        const int bn4_log2 = p.bn_log2 - 2;                       // float4 groups per tile row
        const int groups = BM << bn4_log2;                        // multiple of EPI_THREADS (BN >= 16)
        if (p.direct) {
          // ---- one-wave launches (small batches): no smem transposition, no second pass. Each thread owns a row and
          //      its 16-column groups: kept values straight from global into registers (whole 64-byte runs, the first
          //      group prefetched before the accumulator wait), math, then stores / folded scatter from registers.
          for (int c = tid_e; c < BN; c += EPI_THREADS) {
            bias_s[c] = __ldg(p.bias + col_base + c);
            bias_t[c] = __ldg(p.bias2 + col_base + c);
            if (p.f_out != nullptr && col_base + c < p.n_valid) {
              const int src = p.keep_off + col_base + c;
              const int dst = p.f_dst != nullptr ? __ldg(p.f_dst + src) : src;
              col_dst[c] = dst;
              col_scale[c] = 1.f;
              col_loc[c] = 0.f;
              if (p.f_mode != 0) {
                const int pi = p.f_mode == 1 ? dst : src;
                const float sc = __ldg(p.f_scale + pi);
                col_scale[c] = p.f_mode == 2 ? 1.0f / sc : sc;
                col_loc[c] = __ldg(p.f_loc + pi);
              }
            }
          }
          float* krow = p.state + static_cast<size_t>(row_ok ? row : 0) * p.ld_state + p.keep_off + col_base;
          float kp[16];
          auto load_keep = [&](int c0) {
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              const int c = c0 + 4 * j4;
              float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
              if (row_ok && col_base + c < p.n_valid) {
                if (p.vec4) {
                  v = *reinterpret_cast<const float4*>(krow + c);
                } else {
                  v.x = krow[c];
                  if (col_base + c + 1 < p.n_valid) v.y = krow[c + 1];
                  if (col_base + c + 2 < p.n_valid) v.z = krow[c + 2];
                  if (col_base + c + 3 < p.n_valid) v.w = krow[c + 3];
                }
              }
              kp[4 * j4] = v.x; kp[4 * j4 + 1] = v.y; kp[4 * j4 + 2] = v.z; kp[4 * j4 + 3] = v.w;
            }
          };
          if (ch * 16 < BN) load_keep(ch * 16);
          named_bar_sync(1, EPI_THREADS);                           // biases / column constants are in smem
          mbar_wait(&tfull_bar[acc], acc_phase, 4);
          tc_fence_after();
          if (tr != nullptr && tid_e == 0) tr[5] = globaltimer_ns();
          float ld_sum = 0.f;
          const float neg = p.f_leaky ? (p.f_mode == 2 ? 1.0f / p.f_alpha : p.f_alpha) : 1.0f;
          const bool sv = p.s_save != nullptr && row_ok;
          for (int c0 = ch * 16; c0 < BN; c0 += 32) {
            float vs[16], vt[16];
            tmem_ld16(t_row + c0, vs);
            tmem_ld16(t_row + sub_cols + c0, vt);
            tmem_ld_wait();
            for (int a = 1; a < kacc; ++a) {
              float ws_[16], wt_[16];
              tmem_ld16(t_row + a * BN + c0, ws_);
              tmem_ld16(t_row + sub_cols + a * BN + c0, wt_);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 16; ++j) { vs[j] += ws_[j]; vt[j] += wt_[j]; }
            }
            if (c0 + 32 >= BN) {                                    // last group: hand the accumulator back early
              tc_fence_before();
              mbar_arrive(&tempty_bar[acc]);
            }
            const int gcol = col_base + c0;
            const int nv = p.n_valid - gcol < 16 ? p.n_valid - gcol : 16;   // <= 0 for padded groups
            if (sv) {
              float* ks = p.keep_save + static_cast<size_t>(row) * p.ld_save + gcol;
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nv) ks[j] = kp[j];
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float s = kSplit == 2 ? tanhf(vs[j] + bias_s[c0 + j]) : tanh_approx(vs[j] + bias_s[c0 + j]);
              const float t = vt[j] + bias_t[c0 + j];
              if (kMode == EPI_TAIL_FWD) {
                kp[j] = kp[j] * (kSplit == 2 ? expf(s) : ex2_approx(s * 1.4426950408889634f)) + t;
                ld_sum += s;                                        // padded columns: zero weights and biases -> s = 0
              } else {
                kp[j] = (kp[j] - t) * (kSplit == 2 ? expf(-s) : ex2_approx(s * -1.4426950408889634f));
              }
              vs[j] = s;
            }
            if (row_ok && nv > 0) {
              if (sv) {
                float* ss = p.s_save + static_cast<size_t>(row) * p.ld_save + gcol;
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  if (j < nv) ss[j] = vs[j];
              }
              if (p.u_hi != nullptr) {                              // bf16 copy for this block's second half-step
                __nv_bfloat16* uh = p.u_hi + static_cast<size_t>(row) * p.ld_u + gcol;
                __nv_bfloat16* ul = kSplit == 2 ? p.u_lo + static_cast<size_t>(row) * p.ld_u + gcol : nullptr;
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4) {
                  __nv_bfloat16 h[4], l[4];
#pragma unroll
                  for (int e = 0; e < 4; ++e) split_bf16(kp[4 * j4 + e], h[e], l[e]);
                  if (p.vec4) {
                    if (4 * j4 < nv) {
                      *reinterpret_cast<uint2*>(uh + 4 * j4) = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
                      if (kSplit == 2)
                        *reinterpret_cast<uint2*>(ul + 4 * j4) = make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
                    }
                  } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                      if (4 * j4 + e < nv) {
                        uh[4 * j4 + e] = h[e];
                        if (kSplit == 2) ul[4 * j4 + e] = l[e];
                      }
                  }
                }
              }
              if (p.f_out == nullptr) {
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4) {
                  if (p.vec4) {
                    if (4 * j4 < nv)
                      *reinterpret_cast<float4*>(krow + c0 + 4 * j4) =
                          make_float4(kp[4 * j4], kp[4 * j4 + 1], kp[4 * j4 + 2], kp[4 * j4 + 3]);
                  } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                      if (4 * j4 + e < nv) krow[c0 + 4 * j4 + e] = kp[4 * j4 + e];
                  }
                }
              } else {
                float* orow = p.f_out + static_cast<size_t>(row) * p.f_ld;
                const size_t uoff = static_cast<size_t>(row) * p.ld_u;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  if (j < nv) {
                    const int dst = col_dst[c0 + j];
                    float v = kp[j];
                    if (p.f_mode == 1) {
                      v = col_scale[c0 + j] * (v + col_loc[c0 + j]);
                      v = v * (v >= 0.f ? 1.0f : neg);
                    } else if (p.f_mode == 2) {
                      v = v * (v >= 0.f ? 1.0f : neg);
                      v = v * col_scale[c0 + j] - col_loc[c0 + j];
                    }
                    orow[dst] = v;
                    const int ucol = dst - p.f_u_begin;
                    if (p.f_u_hi != nullptr && ucol >= 0 && ucol < p.n_valid) {
                      __nv_bfloat16 h, l;
                      split_bf16(v, h, l);
                      p.f_u_hi[uoff + ucol] = h;
                      if (kSplit == 2) p.f_u_lo[uoff + ucol] = l;
                    }
                  }
                }
              }
            }
            if (c0 + 32 < BN) load_keep(c0 + 32);
          }
          if (ch * 16 >= BN) {                                      // warps without a column group still release TMEM
            tc_fence_before();
            mbar_arrive(&tempty_bar[acc]);
          }
          if (kMode == EPI_TAIL_FWD && row_ok) {
            if (p.ld_extra != nullptr && n_tile == 0 && ch == 0) ld_sum += __ldg(p.ld_extra);
            atomicAdd(p.logdet + row, ld_sum);
          }
          named_bar_sync(1, EPI_THREADS);                           // smem constants are rewritten by the next tile
          continue;
        }
        // 8 independent 16-byte loads in flight per thread (a scalar loop here is latency-bound: ncu r01)
        for (int g0 = tid_e; g0 < groups; g0 += 8 * EPI_THREADS) {
          float4 kv[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int idx = g0 + u * EPI_THREADS;
            kv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < groups) {
              const int r = idx >> bn4_log2;
              const int gcol = col_base + ((idx & ((1 << bn4_log2) - 1)) << 2);
              const int grow = m_tile * BM + r;
              if (grow < p.M && gcol < p.n_valid) {
                const float* src = p.state + static_cast<size_t>(grow) * p.ld_state + p.keep_off + gcol;
                if (p.vec4) {
                  kv[u] = *reinterpret_cast<const float4*>(src);
                } else {
                  kv[u].x = src[0];
                  if (gcol + 1 < p.n_valid) kv[u].y = src[1];
                  if (gcol + 2 < p.n_valid) kv[u].z = src[2];
                  if (gcol + 3 < p.n_valid) kv[u].w = src[3];
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int idx = g0 + u * EPI_THREADS;
            if (idx < groups) {
              float* d = keep_s + (idx >> bn4_log2) * ldk + ((idx & ((1 << bn4_log2) - 1)) << 2);
              d[0] = kv[u].x; d[1] = kv[u].y; d[2] = kv[u].z; d[3] = kv[u].w;
            }
          }
        }
        for (int c = tid_e; c < BN; c += EPI_THREADS) {
          bias_s[c] = __ldg(p.bias + col_base + c);
          bias_t[c] = __ldg(p.bias2 + col_base + c);
          if (p.f_out != nullptr && col_base + c < p.n_valid) {
            const int src = p.keep_off + col_base + c;
            const int dst = p.f_dst != nullptr ? __ldg(p.f_dst + src) : src;
            col_dst[c] = dst;
            if (p.f_mode != 0) {
              const int pi = p.f_mode == 1 ? dst : src;           // forward: next block's ActNorm; reverse: this block's
              col_scale[c] = __ldg(p.f_scale + pi);
              col_loc[c] = __ldg(p.f_loc + pi);
            }
          }
        }
        if (tr != nullptr && tid_e == 0) tr[8] = globaltimer_ns();      // keep tile staged (this thread)
        named_bar_sync(1, EPI_THREADS);
        if (tr != nullptr && tid_e == 0) tr[9] = globaltimer_ns();
        mbar_wait(&tfull_bar[acc], acc_phase, 4);
        tc_fence_after();
        if (tr != nullptr && tid_e == 0) tr[5] = globaltimer_ns();
        float ld_sum = 0.f;
        float* my_keep = keep_s + r_loc * ldk;
        const bool sv = p.s_save != nullptr && row_ok;            // training forward keeps s and the old kept half
        for (int c0 = ch * 16; c0 < BN; c0 += 32) {
          float vs[16], vt[16];
          tmem_ld16(t_row + c0, vs);
          tmem_ld16(t_row + sub_cols + c0, vt);
          tmem_ld_wait();
          for (int a = 1; a < kacc; ++a) {
            float ws_[16], wt_[16];
            tmem_ld16(t_row + a * BN + c0, ws_);
            tmem_ld16(t_row + sub_cols + a * BN + c0, wt_);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) { vs[j] += ws_[j]; vt[j] += wt_[j]; }
          }
          if (sv) {
            float* ks = p.keep_save + static_cast<size_t>(row) * p.ld_save + col_base + c0;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (col_base + c0 + j < p.n_valid) ks[j] = my_keep[c0 + j];
          }
          // padded columns (>= n_valid) have zero weights and biases: s = tanh(0) = 0 adds nothing to the log-det
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const float4 bs4 = *reinterpret_cast<const float4*>(bias_s + c0 + 4 * j4);
            const float4 bt4 = *reinterpret_cast<const float4*>(bias_t + c0 + 4 * j4);
            const float bsv[4] = {bs4.x, bs4.y, bs4.z, bs4.w};
            const float btv[4] = {bt4.x, bt4.y, bt4.z, bt4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int j = 4 * j4 + e;
              const int c = c0 + j;
              // bf16 mode: MUFU tanh / ex2 (abs error <= 5e-4 / 2e-7 relative, below the bf16 GEMM error);
              // fp32-parity mode keeps libm-accurate tanhf / expf
              const float s = kSplit == 2 ? tanhf(vs[j] + bsv[e]) : tanh_approx(vs[j] + bsv[e]);
              const float t = vt[j] + btv[e];
              const float keep = my_keep[c];
              float nw;
              if (kMode == EPI_TAIL_FWD) {
                nw = keep * (kSplit == 2 ? expf(s) : ex2_approx(s * 1.4426950408889634f)) + t;
                ld_sum += s;
              } else {
                nw = (keep - t) * (kSplit == 2 ? expf(-s) : ex2_approx(s * -1.4426950408889634f));
              }
              my_keep[c] = nw;
              vs[j] = s;
            }
          }
          if (sv) {
            float* ss = p.s_save + static_cast<size_t>(row) * p.ld_save + col_base + c0;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (col_base + c0 + j < p.n_valid) ss[j] = vs[j];
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty_bar[acc]);
        if (tr != nullptr && tid_e == 0) tr[10] = globaltimer_ns();     // per-row math done
        if (kMode == EPI_TAIL_FWD && row_ok) {
          if (p.ld_extra != nullptr && n_tile == 0 && ch == 0) ld_sum += __ldg(p.ld_extra);
          atomicAdd(p.logdet + row, ld_sum);       // two column halves (and possibly several n-tiles) per row
        }
        named_bar_sync(1, EPI_THREADS);
        if (tr != nullptr && tid_e == 0) tr[11] = globaltimer_ns();
        if (p.f_out != nullptr) {
          // ---- folded boundary: lane = column (conflict-free smem reads, per-column constants live in registers;
          //      BN = 16 packs two rows into one warp pass), the 8 warps walk the rows; every element goes to its
          //      place in the next block's state
          const int we = warp - 4;
          const int lpr = BN < 32 ? BN : 32;                        // lanes per row
          const int rpp = 32 / lpr;                                 // rows per warp pass
          const int cl0 = lane & (lpr - 1);
          const int kmax = (BN + 31) >> 5;
          float cs[4], cl[4];
          int cd[4], cu[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int c = cl0 + 32 * k;
            cd[k] = -1;
            cu[k] = -1;
            cs[k] = 1.f;
            cl[k] = 0.f;
            if (k < kmax && col_base + c < p.n_valid) {
              cd[k] = col_dst[c];
              cs[k] = p.f_mode == 2 ? 1.0f / col_scale[c] : col_scale[c];
              cl[k] = col_loc[c];
              const int ucol = cd[k] - p.f_u_begin;
              if (p.f_u_hi != nullptr && ucol >= 0 && ucol < p.n_valid) cu[k] = ucol;
            }
          }
          const float neg = p.f_leaky ? (p.f_mode == 2 ? 1.0f / p.f_alpha : p.f_alpha) : 1.0f;
          const int rows_here = p.M - m_tile * BM < BM ? p.M - m_tile * BM : BM;
          const bool has_u = p.u_hi != nullptr;
          if (tr != nullptr && tid_e == 0) tr[13] = globaltimer_ns();   // column constants in registers
          // 4 rows per pass: all smem reads first (they cannot be hoisted over the global stores by the compiler),
          // then 4 independent chains of math + stores
          for (int r0 = we * rpp + (lane / lpr); r0 < rows_here; r0 += 32 * rpp) {
            float nwv[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int r = r0 + u * 8 * rpp;
#pragma unroll
              for (int k = 0; k < 4; ++k)
                nwv[u][k] = (k < kmax && cd[k] >= 0 && r < rows_here) ? keep_s[r * ldk + cl0 + 32 * k] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int r = r0 + u * 8 * rpp;
              if (r < rows_here) {
                const size_t grow = static_cast<size_t>(m_tile * BM + r);
                float* orow = p.f_out + grow * p.f_ld;
                const size_t uoff = grow * p.ld_u;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (k < kmax && cd[k] >= 0) {
                    const float nw = nwv[u][k];
                    if (has_u) {                                      // this block's second half-step reads the raw value
                      __nv_bfloat16 h, l;
                      split_bf16(nw, h, l);
                      p.u_hi[uoff + col_base + cl0 + 32 * k] = h;
                      if (kSplit == 2) p.u_lo[uoff + col_base + cl0 + 32 * k] = l;
                    }
                    float v = nw;
                    if (p.f_mode == 1) {
                      v = cs[k] * (v + cl[k]);
                      v = v * (v >= 0.f ? 1.0f : neg);
                    } else if (p.f_mode == 2) {
                      v = v * (v >= 0.f ? 1.0f : neg);
                      v = v * cs[k] - cl[k];
                    }
                    orow[cd[k]] = v;
                    if (cu[k] >= 0) {
                      __nv_bfloat16 h, l;
                      split_bf16(v, h, l);
                      p.f_u_hi[uoff + cu[k]] = h;
                      if (kSplit == 2) p.f_u_lo[uoff + cu[k]] = l;
                    }
                  }
                }
              }
            }
          }
        } else
        for (int idx = tid_e; idx < groups; idx += EPI_THREADS) {
          const int r = idx >> bn4_log2;
          const int c = (idx & ((1 << bn4_log2) - 1)) << 2;
          const int grow = m_tile * BM + r;
          const int gcol = col_base + c;
          if (grow < p.M && gcol < p.n_valid) {
            const float* ks = keep_s + r * ldk + c;
            const float nw[4] = {ks[0], ks[1], ks[2], ks[3]};
            const int nv = p.n_valid - gcol < 4 ? p.n_valid - gcol : 4;
            if (p.u_hi != nullptr) {
              __nv_bfloat16 h[4], l[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) split_bf16(nw[e], h[e], l[e]);
              __nv_bfloat16* uh = p.u_hi + static_cast<size_t>(grow) * p.ld_u + gcol;
              if (p.vec4) {
                *reinterpret_cast<uint2*>(uh) = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
                if (kSplit == 2)
                  *reinterpret_cast<uint2*>(p.u_lo + static_cast<size_t>(grow) * p.ld_u + gcol) =
                      make_uint2(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
              } else {
                for (int e = 0; e < nv; ++e) {
                  uh[e] = h[e];
                  if (kSplit == 2) p.u_lo[static_cast<size_t>(grow) * p.ld_u + gcol + e] = l[e];
                }
              }
            }
            float* d = p.state + static_cast<size_t>(grow) * p.ld_state + p.keep_off + gcol;
            if (p.vec4) *reinterpret_cast<float4*>(d) = make_float4(nw[0], nw[1], nw[2], nw[3]);
            else for (int e = 0; e < nv; ++e) d[e] = nw[e];
          }
        }
        if (tr != nullptr && tid_e == 0) tr[12] = globaltimer_ns();     // write-back issued (this thread)
        named_bar_sync(1, EPI_THREADS);                           // keep_s is reused by the next tile
      }
    }
    if (kSlabMode && p.slab_bufs > 0 && ((kWarpSlab && p.slab_warp) ? lane == 0 : tid_e == 0)) bulk_wait<0>();   // stores landed before exit
    if (tr != nullptr && tid_e == 0) tr[6] = globaltimer_ns();
