// libipoke_inn.so -- host side of the C ABI declared in include/ipoke_inn.h.
// Builds, per (batch, workspace, direction), a linear program of kernel launches:
//   boundary (gather + ActNorm + leaky)  ->  per half-step: (depth+1) x inn_gemm<EPI_HIDDEN>, inn_gemm<EPI_TAIL_*>
// and replays it on the caller's stream. No device allocation, no host synchronisation.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/ipoke_inn.h"
#include "elementwise.cuh"
#include "gemm_kernel.cuh"

using namespace inn;

namespace {

constexpr int kMaxDynSmem = 232448;  // 227 KB
constexpr int MODE_UNCOND = 0, MODE_NORMAL = 1, MODE_COND = 2;

inline long long round_up(long long v, long long a) { return (v + a - 1) / a * a; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct LayerW {
  long long hi = -1, lo = -1;   // byte offsets of the bf16 matrices
  long long bias = -1;          // byte offset of the fp32 bias (rows entries)
  int rows = 0, Kp = 0;
};
struct Chain {                   // (depth+1) hidden layers run for G subnetworks in lockstep
  int G = 2;
  int Kin_p = 0;                 // padded width of the first layer's input view
  int a_col_off = 0;             // first column of that view inside the MLP input buffer U
  std::vector<LayerW> L;
};
struct Tail {                    // last Linear of s_i and t_i, rows [0,Np) = s, [Np,2Np) = t
  LayerW w;
  long long bias_t = -1;
  int chain = 0, g_s = 0, g_t = 1;
};
struct Block {
  int mode = MODE_UNCOND;
  Chain chains[2];
  int n_chains = 2;
  Tail tails[2];
  long long loc = -1, scale = -1, an_ld = -1, map_fwd = -1, map_rev = -1;   // byte offsets
};

struct GemmOp {
  CUtensorMap tmA[2], tmW[2], tmO[2];
  GemmParams p;
  int mode, grid, smem, cluster;
};
struct Op {
  int kind;  // 0 boundary, 1 gemm
  BoundaryParams b;
  int b_grid;
  int gemm;  // index into gemm ops
  // which user pointers this boundary op takes (patched per call)
  int user_in, user_cond, user_out_final, block_index;
};
struct ProgramKey {
  const void* packed; void* ws; int B, kb, ke, dir, want_blocks;
  bool operator==(const ProgramKey& o) const {
    return packed == o.packed && ws == o.ws && B == o.B && kb == o.kb && ke == o.ke && dir == o.dir &&
           want_blocks == o.want_blocks;
  }
};
struct GraphEntry {
  const void* ptrs[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaGraphExec_t exec = nullptr;
  unsigned long long last_use = 0;
};
struct Program {
  ProgramKey key;
  std::vector<Op> ops;
  std::vector<GemmOp> gemms;
  int launches = 0;
  std::vector<GraphEntry> graphs;     // instantiated CUDA graphs keyed by the caller's pointers
  int graph_misses = 0, graph_hits = 0;
  ~Program() {
    for (auto& g : graphs)
      if (g.exec) cudaGraphExecDestroy(g.exec);
  }
};

struct WsLayout {
  long long S[2], U_hi, U_lo, ACT_hi[2], ACT_lo[2], logdet, xin, condin, zout, ldout, total;
};

}  // namespace

struct ipoke_inn_plan {
  ipoke_inn_desc d;
  int device = 0;
  int ksplit = 2;
  int C2 = 0, C2p = 0, Ep = 0, Hp = 0, Np = 0, ldU = 0, Gmax = 2, ldACT = 0;
  std::vector<Block> blocks;
  long long packed_bytes = 0;
  // device-side job tables live inside the packed buffer
  long long jobs_linear = -1, jobs_block = -1;
  // lazy device state
  bool dev_ready = false;
  int num_sms = 148;
  EncodeTiledFn encode = nullptr;
  std::mutex mu;
  std::vector<std::unique_ptr<Program>> programs;
  int last_launches = 0;
  // launch options: 1 = replay the program as a CUDA graph (captured on an internal stream), PDL attribute on GEMMs
  int use_graph = 1, use_pdl = 1, use_mcast = 0;
  int w_prefetch = 1;
  int latency_smem_cap_kb = 0;   // > 0: cap smem of one-wave launches so two kernels can be co-resident under PDL
  cudaStream_t cap_stream = nullptr;
  unsigned long long tick = 0;
  // optional per-launch CUDA-event timing (bench roofline): events of the last profiled run
  bool profile = false;
  std::vector<cudaEvent_t> prof_ev;
  std::vector<int> prof_kind;   // 0 boundary, 1 hidden GEMM, 2 tail GEMM
};

namespace {

long long bump(long long& cur, long long bytes) {
  cur = round_up(cur, 1024);
  const long long o = cur;
  cur += bytes;
  return o;
}

void layout_layer(ipoke_inn_plan* P, long long& cur, LayerW& L, int rows, int Kp) {
  L.rows = rows;
  L.Kp = Kp;
  L.hi = bump(cur, 2ll * rows * Kp);
  if (P->ksplit == 2) L.lo = bump(cur, 2ll * rows * Kp);
  L.bias = bump(cur, 4ll * rows);
}

WsLayout ws_layout(const ipoke_inn_plan* P, int B) {
  WsLayout w;
  long long cur = 0;
  const long long C = P->d.C;
  w.S[0] = bump(cur, 4ll * B * C);
  w.S[1] = bump(cur, 4ll * B * C);
  w.U_hi = bump(cur, 2ll * B * P->ldU);
  w.U_lo = bump(cur, 2ll * B * P->ldU);
  for (int i = 0; i < 2; ++i) {
    w.ACT_hi[i] = bump(cur, 2ll * B * P->ldACT);
    w.ACT_lo[i] = bump(cur, 2ll * B * P->ldACT);
  }
  w.logdet = bump(cur, 4ll * B);
  w.xin = bump(cur, 4ll * B * C);
  w.condin = bump(cur, 4ll * B * std::max(1, (int)P->d.E));
  w.zout = bump(cur, 4ll * B * C);
  w.ldout = bump(cur, 4ll * B);
  w.total = round_up(cur, 1024);
  return w;
}

int ensure_device(ipoke_inn_plan* P) {
  if (P->dev_ready) return 0;
  cudaError_t e = cudaSetDevice(P->device);
  if (e != cudaSuccess) return IPOKE_INN_E_NO_DEVICE;
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, P->device);
  if (e != cudaSuccess) return IPOKE_INN_E_NO_DEVICE;
  if (prop.major != 10) return IPOKE_INN_E_NO_DEVICE;
  P->num_sms = prop.multiProcessorCount;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || fn == nullptr || q != cudaDriverEntryPointSuccess) return IPOKE_INN_E_NO_DEVICE;
  P->encode = reinterpret_cast<EncodeTiledFn>(fn);
#define SET_SMEM(K)                                                                                    \
  e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);               \
  if (e != cudaSuccess) return (int)e;
  SET_SMEM((inn_gemm_kernel<1, EPI_HIDDEN>))
  SET_SMEM((inn_gemm_kernel<2, EPI_HIDDEN>))
  SET_SMEM((inn_gemm_kernel<1, EPI_TAIL_FWD>))
  SET_SMEM((inn_gemm_kernel<2, EPI_TAIL_FWD>))
  SET_SMEM((inn_gemm_kernel<1, EPI_TAIL_REV>))
  SET_SMEM((inn_gemm_kernel<2, EPI_TAIL_REV>))
#undef SET_SMEM
  P->dev_ready = true;
  return 0;
}

int make_map(const ipoke_inn_plan* P, CUtensorMap* m, const void* base, long long inner, long long outer,
             long long row_stride_bytes, int box_outer) {
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)row_stride_bytes};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = P->encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : IPOKE_INN_E_INTERNAL;
}

struct TileCfg { int bn = 0, stages = 0, slab_bufs = 0, mc = 1; };

// smem plan for one tile width: pipeline stages + epilogue staging must fit in 227 KB
TileCfg tile_cfg(int ksplit, int mode, int bn) {
  TileCfg t;
  t.bn = bn;
  const int avail = kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES;
  const int sb = gemm_stage_bytes(ksplit, mode, bn);
  auto stages_with = [&](int slabs) {
    const int epi = (gemm_epi_bytes(ksplit, mode, bn, slabs) + 15) & ~15;
    return std::min(MAX_STAGES, (avail - epi) / sb);
  };
  if (mode == EPI_HIDDEN && bn >= 64) {
    const int s2 = stages_with(2), s1 = stages_with(1);
    if (s2 >= 4 || s2 == s1) { t.slab_bufs = 2; t.stages = s2; }
    else { t.slab_bufs = 1; t.stages = s1; }
  } else {
    t.slab_bufs = 0;
    t.stages = stages_with(0);
  }
  return t;
}

// Largest UMMA N that divides `rows_per_group`, keeps >= 2 pipeline stages and still yields about one CTA
// per SM; otherwise the smallest (most tiles) -- small batches are latency/weight-streaming bound.
TileCfg pick_bn(const ipoke_inn_plan* P, int mode, int rows_per_group, int tiles_other) {
  static const int cand[5] = {256, 128, 64, 32, 16};
  TileCfg best;
  for (int i = 0; i < 5; ++i) {
    const int bn = cand[i];
    if (rows_per_group % bn != 0) continue;
    const TileCfg t = tile_cfg(P->ksplit, mode, bn);
    if (t.stages < 2) continue;
    best = t;
    if ((long long)tiles_other * (rows_per_group / bn) >= (long long)(P->num_sms * 4) / 5) return t;
  }
  return best;
}

// Latency mode (one wave of tiles): CTAs of consecutive n-tiles share their A tile -> cluster + TMA multicast.
int pick_mc(const ipoke_inn_plan* P, int tiles, int n_tiles, int tiles_per_group) {
  if (!P->use_mcast || tiles > P->num_sms) return 1;
  static const int cand[3] = {8, 4, 2};
  for (int i = 0; i < 3; ++i)
    if (n_tiles % cand[i] == 0 && tiles_per_group % cand[i] == 0) return cand[i];
  return 1;
}

struct Builder {
  ipoke_inn_plan* P;
  Program* prog;
  const char* packed;
  char* ws;
  WsLayout wl;
  int B, m_tiles;
  int err = 0;

  template <typename T>
  T* pk(long long off) const { return reinterpret_cast<T*>(const_cast<char*>(packed) + off); }
  template <typename T>
  T* w(long long off) const { return reinterpret_cast<T*>(ws + off); }

  Op& add_boundary() {
    prog->ops.emplace_back();
    Op& o = prog->ops.back();
    std::memset(&o, 0, sizeof(Op));
    o.kind = 0;
    o.gemm = -1;
    o.block_index = -1;
    BoundaryParams& b = o.b;
    b.B = B;
    b.C = P->d.C;
    b.alpha = P->d.alpha;
    b.ld_in = b.ld_out = b.ld_raw = P->d.C;
    const long long total = (long long)B * std::max(P->d.C, P->ldU);
    o.b_grid = (int)std::min<long long>((total + 255) / 256, (long long)P->num_sms * 8);
    prog->launches++;
    return o;
  }
  void set_u(BoundaryParams& b, int u_begin) {
    b.u_hi = w<__nv_bfloat16>(wl.U_hi);
    b.u_lo = P->ksplit == 2 ? w<__nv_bfloat16>(wl.U_lo) : nullptr;
    b.ld_u = P->ldU;
    b.u_begin = u_begin;
    b.u_count = P->C2;
    b.u_pad = P->C2p;
  }

  GemmOp& add_gemm(int mode) {
    prog->gemms.emplace_back();
    prog->ops.emplace_back();
    Op& o = prog->ops.back();
    std::memset(&o, 0, sizeof(Op));
    o.kind = 1;
    o.gemm = (int)prog->gemms.size() - 1;
    GemmOp& g = prog->gemms.back();
    std::memset(&g, 0, sizeof(GemmOp));
    g.mode = mode;
    prog->launches++;
    return g;
  }
  void finish_gemm(GemmOp& g, const TileCfg& t) {
    g.p.M = B;
    g.p.num_m_tiles = m_tiles;
    g.p.BN = t.bn;
    g.p.bn_log2 = 0;
    while ((1 << g.p.bn_log2) < t.bn) g.p.bn_log2++;
    g.p.stages = std::min(t.stages, std::max(2, g.p.num_k_blocks));
    g.p.slab_bufs = t.slab_bufs;
    g.p.mcast = t.mc;
    g.p.w_prefetch = P->w_prefetch;
    const int tiles = g.p.num_m_tiles * g.p.num_n_tiles;
    if (tiles <= P->num_sms && P->latency_smem_cap_kb > 0) {
      // optional: keep the footprint small so the next kernel's CTA (PDL) can be co-resident
      while (g.p.stages > 2 &&
             gemm_smem_bytes(P->ksplit, g.mode, t.bn, g.p.stages, t.slab_bufs) > P->latency_smem_cap_kb * 1024)
        g.p.stages--;
    }
    const int acc_cols = gemm_nsub(g.mode) * t.bn;
    g.p.nacc = (2 * acc_cols <= 512) ? 2 : 1;
    int cols = 32;
    while (cols < g.p.nacc * acc_cols) cols <<= 1;
    g.p.tmem_cols = cols;
    g.p.epi_off = g.p.stages * gemm_stage_bytes(P->ksplit, g.mode, t.bn);
    g.smem = gemm_smem_bytes(P->ksplit, g.mode, t.bn, g.p.stages, t.slab_bufs);
    g.grid = t.mc > 1 ? tiles : std::min(tiles, P->num_sms);
    g.cluster = t.mc;
    if (P->ksplit == 1) {
      g.tmA[1] = g.tmA[0];
      g.tmW[1] = g.tmW[0];
      g.tmO[1] = g.tmO[0];
    }
  }

  // hidden layers of one chain; returns the ACT buffer index holding the last hidden activation
  int chain(const Chain& c) {
    const int rows = c.G * P->Hp;
    TileCfg tc = pick_bn(P, EPI_HIDDEN, P->Hp, m_tiles * c.G);
    const int bn = tc.bn;
    tc.mc = pick_mc(P, m_tiles * (rows / bn), rows / bn, P->Hp / bn);
    const int a_box = BM / tc.mc;
    for (size_t l = 0; l < c.L.size(); ++l) {
      const LayerW& L = c.L[l];
      GemmOp& g = add_gemm(EPI_HIDDEN);
      g.p.num_n_tiles = rows / bn;
      g.p.tiles_per_group = P->Hp / bn;
      g.p.num_k_blocks = L.Kp / BK;
      g.p.a_group_stride = l == 0 ? 0 : P->Hp;
      g.p.bias = pk<float>(L.bias);
      g.p.out_hi = w<__nv_bfloat16>(wl.ACT_hi[l & 1]);
      g.p.out_lo = w<__nv_bfloat16>(wl.ACT_lo[l & 1]);
      g.p.ld_out = P->ldACT;
      g.p.slope = P->d.mlp_slope;
      for (int sp = 0; sp < P->ksplit; ++sp) {
        if (l == 0) {
          const long long base = (sp == 0 ? wl.U_hi : wl.U_lo) + 2ll * c.a_col_off;
          err |= make_map(P, &g.tmA[sp], ws + base, c.Kin_p, B, 2ll * P->ldU, a_box);
        } else {
          const long long base = sp == 0 ? wl.ACT_hi[(l - 1) & 1] : wl.ACT_lo[(l - 1) & 1];
          err |= make_map(P, &g.tmA[sp], ws + base, rows, B, 2ll * P->ldACT, a_box);
        }
        err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? L.hi : L.lo), L.Kp, L.rows, 2ll * L.Kp, bn);
        // output slabs (TMA store): same 64 x 128 box as the next layer's A loads
        err |= make_map(P, &g.tmO[sp], ws + (sp == 0 ? wl.ACT_hi[l & 1] : wl.ACT_lo[l & 1]), rows, B, 2ll * P->ldACT, BM);
      }
      finish_gemm(g, tc);
    }
    return (int)(c.L.size() - 1) & 1;
  }

  void tail(const Block& blk, int i, int act_buf, int G, int state_buf, bool reverse, bool write_u) {
    const Tail& t = blk.tails[i];
    GemmOp& g = add_gemm(reverse ? EPI_TAIL_REV : EPI_TAIL_FWD);
    TileCfg tc = pick_bn(P, g.mode, P->Np, m_tiles);
    const int bn = tc.bn;
    g.p.num_n_tiles = P->Np / bn;
    tc.mc = pick_mc(P, m_tiles * (P->Np / bn), P->Np / bn, P->Np / bn);
    g.p.tiles_per_group = 1 << 30;
    g.p.num_k_blocks = P->Hp / BK;
    g.p.a_group_stride = 0;
    g.p.a_off[0] = t.g_s * P->Hp;
    g.p.a_off[1] = t.g_t * P->Hp;
    g.p.w_row0[0] = 0;
    g.p.w_row0[1] = P->Np;
    g.p.bias = pk<float>(t.w.bias);
    g.p.bias2 = pk<float>(t.bias_t);
    g.p.state = w<float>(wl.S[state_buf]);
    g.p.ld_state = P->d.C;
    g.p.keep_off = (i == 0) ? P->C2 : 0;
    g.p.n_valid = P->C2;
    g.p.logdet = w<float>(wl.logdet);
    g.p.logdet_atomic = g.p.num_n_tiles > 1;
    if (write_u) {
      g.p.u_hi = w<__nv_bfloat16>(wl.U_hi);
      g.p.u_lo = w<__nv_bfloat16>(wl.U_lo);
      g.p.ld_u = P->ldU;
    }
    for (int sp = 0; sp < P->ksplit; ++sp) {
      const long long base = sp == 0 ? wl.ACT_hi[act_buf] : wl.ACT_lo[act_buf];
      err |= make_map(P, &g.tmA[sp], ws + base, G * P->Hp, B, 2ll * P->ldACT, BM / tc.mc);
      err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? t.w.hi : t.w.lo), t.w.Kp, t.w.rows, 2ll * t.w.Kp, bn);
      g.tmO[sp] = g.tmA[sp];      // unused by the tail epilogue
    }
    finish_gemm(g, tc);
  }

  void coupling(const Block& blk, int state_buf, bool reverse) {
    if (blk.mode == MODE_COND) {
      const int ab = chain(blk.chains[0]);
      tail(blk, reverse ? 1 : 0, ab, 4, state_buf, reverse, false);
      tail(blk, reverse ? 0 : 1, ab, 4, state_buf, reverse, false);
    } else {
      const int first = reverse ? 1 : 0;
      int ab = chain(blk.chains[first]);
      tail(blk, first, ab, 2, state_buf, reverse, true);
      ab = chain(blk.chains[1 - first]);
      tail(blk, 1 - first, ab, 2, state_buf, reverse, false);
    }
  }

  void build_forward(int kb, int ke, bool want_blocks) {
    const ipoke_inn_desc& d = P->d;
    int cur = 0;
    for (int k = kb; k < ke; ++k) {
      const Block& blk = P->blocks[k];
      Op& o = add_boundary();
      BoundaryParams& b = o.b;
      const int nxt = (k == kb) ? 0 : (cur ^ 1);
      if (k == kb) {
        o.user_in = 1;
        o.user_cond = d.E > 0;
        b.logdet_init = 1;
      } else {
        b.in = w<float>(wl.S[cur]);
        b.src = pk<int>(P->blocks[k - 1].map_fwd);
        if (want_blocks) o.block_index = k - 1 - kb;
      }
      b.out = w<float>(wl.S[nxt]);
      b.post_loc = pk<float>(blk.loc);
      b.post_scale = pk<float>(blk.scale);
      b.post_leaky = d.conditional;
      set_u(b, 0);
      if (d.E > 0) {
        b.E = d.E;
        b.ld_cond = d.E;
        b.u_cond_off = P->C2p;
        b.e_pad = P->Ep;
      }
      b.logdet = w<float>(wl.logdet);
      b.an_logdet = pk<float>(blk.an_ld);
      cur = nxt;
      coupling(blk, cur, false);
    }
    Op& o = add_boundary();
    BoundaryParams& b = o.b;
    b.in = w<float>(wl.S[cur]);
    b.src = pk<int>(P->blocks[ke - 1].map_fwd);
    b.logdet = w<float>(wl.logdet);
    o.user_out_final = 1;
    if (want_blocks) o.block_index = ke - 1 - kb;
  }

  void build_reverse(int kb, int ke) {
    const ipoke_inn_desc& d = P->d;
    int cur = 0;
    for (int k = ke - 1; k >= kb; --k) {
      const Block& blk = P->blocks[k];
      Op& o = add_boundary();
      BoundaryParams& b = o.b;
      const int nxt = (k == ke - 1) ? 0 : (cur ^ 1);
      if (k == ke - 1) {
        o.user_in = 1;
        o.user_cond = d.E > 0;
      } else {
        b.in = w<float>(wl.S[cur]);
        b.pre_loc = pk<float>(P->blocks[k + 1].loc);
        b.pre_scale = pk<float>(P->blocks[k + 1].scale);
        b.pre_leaky = d.conditional;
      }
      b.src = pk<int>(blk.map_rev);
      b.out = w<float>(wl.S[nxt]);
      set_u(b, P->C2);
      if (d.E > 0) {
        b.E = d.E;
        b.ld_cond = d.E;
        b.u_cond_off = P->C2p;
        b.e_pad = P->Ep;
      }
      cur = nxt;
      coupling(blk, cur, true);
    }
    Op& o = add_boundary();
    BoundaryParams& b = o.b;
    b.in = w<float>(wl.S[cur]);
    b.pre_loc = pk<float>(P->blocks[kb].loc);
    b.pre_scale = pk<float>(P->blocks[kb].scale);
    b.pre_leaky = d.conditional;
    o.user_out_final = 1;
  }
};

Program* get_program(ipoke_inn_plan* P, const ProgramKey& key, int* err) {
  for (auto& pr : P->programs)
    if (pr->key == key) return pr.get();
  std::unique_ptr<Program> prog(new Program());
  prog->key = key;
  Builder bld;
  bld.P = P;
  bld.prog = prog.get();
  bld.packed = static_cast<const char*>(key.packed);
  bld.ws = static_cast<char*>(key.ws);
  bld.wl = ws_layout(P, key.B);
  bld.B = key.B;
  bld.m_tiles = (key.B + BM - 1) / BM;
  // reserve so references stay valid while building
  const int per_block = 1 + 2 * (P->d.depth + 2);
  prog->ops.reserve((size_t)(key.ke - key.kb) * per_block + 2);
  prog->gemms.reserve((size_t)(key.ke - key.kb) * per_block);
  if (key.dir == 0) bld.build_forward(key.kb, key.ke, key.want_blocks != 0);
  else bld.build_reverse(key.kb, key.ke);
  if (bld.err) {
    *err = IPOKE_INN_E_INTERNAL;
    return nullptr;
  }
  if (P->programs.size() >= 16) P->programs.erase(P->programs.begin());
  P->programs.push_back(std::move(prog));
  return P->programs.back().get();
}

template <int kSplit, int kMode>
cudaError_t launch_gemm_t(const GemmOp& g, cudaStream_t st, int pdl) {
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)g.grid);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = (size_t)g.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (g.cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)g.cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, inn_gemm_kernel<kSplit, kMode>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], g.tmO[0],
                            g.tmO[1], g.p);
}

cudaError_t launch_gemm(const ipoke_inn_plan* P, const GemmOp& g, cudaStream_t st) {
  const int pdl = P->use_pdl;
  if (P->ksplit == 2) {
    switch (g.mode) {
      case EPI_HIDDEN: return launch_gemm_t<2, EPI_HIDDEN>(g, st, pdl);
      case EPI_TAIL_FWD: return launch_gemm_t<2, EPI_TAIL_FWD>(g, st, pdl);
      default: return launch_gemm_t<2, EPI_TAIL_REV>(g, st, pdl);
    }
  }
  switch (g.mode) {
    case EPI_HIDDEN: return launch_gemm_t<1, EPI_HIDDEN>(g, st, pdl);
    case EPI_TAIL_FWD: return launch_gemm_t<1, EPI_TAIL_FWD>(g, st, pdl);
    default: return launch_gemm_t<1, EPI_TAIL_REV>(g, st, pdl);
  }
}

int issue_program(ipoke_inn_plan* P, Program* prog, const float* in, const float* cond, float* out, float* logdet_out,
                  float* block_outs, float* block_lds, cudaStream_t st) {
  const long long BC = (long long)prog->key.B * P->d.C;
  size_t op_i = 0;
  if (P->profile) {
    while (P->prof_ev.size() < 2 * prog->ops.size()) {
      cudaEvent_t ev;
      cudaError_t e = cudaEventCreate(&ev);
      if (e != cudaSuccess) return (int)e;
      P->prof_ev.push_back(ev);
    }
    P->prof_kind.assign(prog->ops.size(), 0);
  }
  struct ProfScope {
    ipoke_inn_plan* P; size_t i; cudaStream_t st;
    ProfScope(ipoke_inn_plan* P_, size_t i_, cudaStream_t st_, int kind) : P(P_), i(i_), st(st_) {
      if (P->profile) { P->prof_kind[i] = kind; cudaEventRecord(P->prof_ev[2 * i], st); }
    }
    ~ProfScope() { if (P->profile) cudaEventRecord(P->prof_ev[2 * i + 1], st); }
  };
  for (Op& o : prog->ops) {
    ProfScope scope(P, op_i++, st, o.kind == 0 ? 0 : (prog->gemms[o.gemm].mode == EPI_HIDDEN ? 1 : 2));
    if (o.kind == 0) {
      BoundaryParams b = o.b;
      if (o.user_in) b.in = in;
      if (o.user_cond) b.cond = cond;
      if (o.block_index >= 0) {
        if (block_outs) b.raw_out = block_outs + o.block_index * BC;
        if (block_lds) b.block_logdet = block_lds + (long long)o.block_index * prog->key.B;
      }
      if (o.user_out_final) {
        if (prog->key.dir == 0) {
          // z (and last_outs[-1]); cumulative logdet to the caller (and last_logdets[-1])
          b.raw_out2 = b.raw_out;
          b.raw_out = out;
          b.block_logdet2 = b.block_logdet;
          b.block_logdet = logdet_out;
        } else {
          b.raw_out = out;
        }
      }
      cudaLaunchConfig_t cfg;
      std::memset(&cfg, 0, sizeof(cfg));
      cfg.gridDim = dim3((unsigned)o.b_grid);
      cfg.blockDim = dim3(256);
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = P->use_pdl ? 1 : 0;
      cudaError_t e = cudaLaunchKernelEx(&cfg, inn_boundary_kernel, b);
      if (e != cudaSuccess) return (int)e;
    } else {
      cudaError_t e = launch_gemm(P, prog->gemms[o.gemm], st);
      if (e != cudaSuccess) return (int)e;
    }
  }
  P->last_launches = prog->launches;
  return 0;
}

// Replay as a CUDA graph when the caller's pointers repeat (steady-state loops); plain launches otherwise.
int run_program(ipoke_inn_plan* P, Program* prog, const float* in, const float* cond, float* out, float* logdet_out,
                float* block_outs, float* block_lds, cudaStream_t st) {
  if (!P->use_graph || P->profile || prog->graph_misses > 8 + 4 * prog->graph_hits)
    return issue_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st);
  const void* ptrs[6] = {in, cond, out, logdet_out, block_outs, block_lds};
  ++P->tick;
  for (auto& g : prog->graphs) {
    if (std::memcmp(g.ptrs, ptrs, sizeof(ptrs)) == 0) {
      g.last_use = P->tick;
      ++prog->graph_hits;
      cudaError_t e = cudaGraphLaunch(g.exec, st);
      P->last_launches = prog->launches;
      return e == cudaSuccess ? 0 : (int)e;
    }
  }
  ++prog->graph_misses;
  if (!P->cap_stream) {
    cudaError_t e = cudaStreamCreateWithFlags(&P->cap_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e = cudaStreamBeginCapture(P->cap_stream, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) return (int)e;
  int rc = issue_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, P->cap_stream);
  cudaGraph_t graph = nullptr;
  e = cudaStreamEndCapture(P->cap_stream, &graph);
  if (rc != 0 || e != cudaSuccess || graph == nullptr) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    P->use_graph = 0;          // capture unsupported here: fall back to plain launches for good
    return issue_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st);
  }
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess || exec == nullptr) {
    cudaGetLastError();
    P->use_graph = 0;
    return issue_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st);
  }
  if (prog->graphs.size() >= 4) {                       // evict the least recently used
    size_t lru = 0;
    for (size_t i = 1; i < prog->graphs.size(); ++i)
      if (prog->graphs[i].last_use < prog->graphs[lru].last_use) lru = i;
    cudaGraphExecDestroy(prog->graphs[lru].exec);
    prog->graphs.erase(prog->graphs.begin() + lru);
  }
  GraphEntry ge;
  std::memcpy(ge.ptrs, ptrs, sizeof(ptrs));
  ge.exec = exec;
  ge.last_use = P->tick;
  prog->graphs.push_back(ge);
  e = cudaGraphLaunch(exec, st);
  return e == cudaSuccess ? 0 : (int)e;
}

int check_common(const ipoke_inn_plan* P, const void* packed, const void* a, const void* b, int B, int kb, int ke,
                 const void* cond, const void* ws) {
  if (!P || !packed || !a || !b || !ws) return IPOKE_INN_E_INVALID;
  if (B <= 0 || kb < 0 || ke > P->d.n_flows || kb >= ke) return IPOKE_INN_E_INVALID;
  if (P->d.E > 0 && !cond) return IPOKE_INN_E_INVALID;
  return 0;
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

int ipoke_inn_version(void) { return IPOKE_INN_VERSION; }

const char* ipoke_inn_last_error_string(int code) {
  switch (code) {
    case IPOKE_INN_OK: return "ok";
    case IPOKE_INN_E_INVALID: return "ipoke_inn: invalid argument";
    case IPOKE_INN_E_UNSUPPORTED: return "ipoke_inn: unsupported shape (C must be even and >= 2; H, depth, n_flows >= 1; E == 0 iff unconditional)";
    case IPOKE_INN_E_NO_DEVICE: return "ipoke_inn: no sm_100 CUDA device / driver entry point unavailable";
    case IPOKE_INN_E_INTERNAL: return "ipoke_inn: internal error (tensor map encode failed)";
    default: break;
  }
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "ipoke_inn: unknown error";
}

int ipoke_inn_plan_create(const ipoke_inn_desc* desc, int device, ipoke_inn_plan** out) {
  if (!desc || !out) return IPOKE_INN_E_INVALID;
  const ipoke_inn_desc& d = *desc;
  if (d.C < 2 || (d.C & 1) || d.H < 1 || d.depth < 0 || d.n_flows < 1 || d.E < 0) return IPOKE_INN_E_UNSUPPORTED;
  if (d.conditional ? d.E < 1 : d.E != 0) return IPOKE_INN_E_UNSUPPORTED;
  if (d.precision != IPOKE_INN_PREC_FP32_SPLIT && d.precision != IPOKE_INN_PREC_BF16) return IPOKE_INN_E_INVALID;
  ipoke_inn_plan* P = new (std::nothrow) ipoke_inn_plan();
  if (!P) return IPOKE_INN_E_INTERNAL;
  P->d = d;
  P->device = device;
  if (const char* e = std::getenv("IPOKE_INN_GRAPH")) P->use_graph = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_PDL")) P->use_pdl = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_MCAST")) P->use_mcast = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_WPREFETCH")) P->w_prefetch = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_STAGECAP_KB")) P->latency_smem_cap_kb = std::atoi(e);
  P->ksplit = d.precision == IPOKE_INN_PREC_FP32_SPLIT ? 2 : 1;
  P->C2 = d.C / 2;
  P->C2p = (int)round_up(P->C2, BK);
  P->Ep = d.E > 0 ? (int)round_up(d.E, BK) : 0;
  P->Hp = (int)round_up(d.H, BK);
  P->Np = (int)round_up(P->C2, 16);
  P->ldU = P->C2p + P->Ep;
  P->Gmax = d.conditional ? 4 : 2;
  P->ldACT = P->Gmax * P->Hp;
  long long cur = 0;
  P->blocks.resize(d.n_flows);
  for (int k = 0; k < d.n_flows; ++k) {
    Block& b = P->blocks[k];
    b.mode = !d.conditional ? MODE_UNCOND : (k % 2 == 0 ? MODE_COND : MODE_NORMAL);   // flow_blocks.py:27
    b.n_chains = b.mode == MODE_COND ? 1 : 2;
    for (int c = 0; c < b.n_chains; ++c) {
      Chain& ch = b.chains[c];
      ch.G = b.mode == MODE_COND ? 4 : 2;
      ch.Kin_p = b.mode == MODE_COND ? P->Ep : (b.mode == MODE_NORMAL ? P->ldU : P->C2p);
      ch.a_col_off = b.mode == MODE_COND ? P->C2p : 0;
      ch.L.resize(d.depth + 1);
      for (int l = 0; l <= d.depth; ++l) layout_layer(P, cur, ch.L[l], ch.G * P->Hp, l == 0 ? ch.Kin_p : P->Hp);
    }
    for (int i = 0; i < 2; ++i) {
      Tail& t = b.tails[i];
      layout_layer(P, cur, t.w, 2 * P->Np, P->Hp);      // bias holds [s | t]
      t.bias_t = t.w.bias + 4ll * P->Np;
      t.chain = b.mode == MODE_COND ? 0 : i;
      t.g_s = b.mode == MODE_COND ? 2 * i : 0;
      t.g_t = t.g_s + 1;
    }
    b.loc = bump(cur, 4ll * d.C);
    b.scale = bump(cur, 4ll * d.C);
    b.an_ld = bump(cur, 16);
    b.map_fwd = bump(cur, 4ll * d.C);
    b.map_rev = bump(cur, 4ll * d.C);
  }
  const int n_lin = d.n_flows * 4 * (d.depth + 2);
  P->jobs_linear = bump(cur, (long long)sizeof(PackJob) * n_lin);
  P->jobs_block = bump(cur, (long long)sizeof(PackBlockJob) * d.n_flows);
  P->packed_bytes = round_up(cur, 1024);
  *out = P;
  return 0;
}

void ipoke_inn_plan_destroy(ipoke_inn_plan* plan) { delete plan; }

size_t ipoke_inn_packed_bytes(const ipoke_inn_plan* plan) { return plan ? (size_t)plan->packed_bytes : 0; }

size_t ipoke_inn_workspace_bytes(const ipoke_inn_plan* plan, int B, int mode) {
  (void)mode;
  if (!plan || B <= 0) return 0;
  return (size_t)ws_layout(plan, B).total;
}

int ipoke_inn_num_linears(const ipoke_inn_plan* plan) {
  return plan ? plan->d.n_flows * 4 * (plan->d.depth + 2) : 0;
}

int ipoke_inn_pack(const ipoke_inn_plan* cplan, const float* const* linear_w, const float* const* linear_b,
                   const float* const* an_loc, const float* const* an_scale, const int64_t* const* perm_fwd,
                   const int64_t* const* perm_bwd, void* packed, void* stream) {
  ipoke_inn_plan* P = const_cast<ipoke_inn_plan*>(cplan);
  if (!P || !linear_w || !linear_b || !an_loc || !an_scale || !perm_fwd || !perm_bwd || !packed) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  int rc = ensure_device(P);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const ipoke_inn_desc& d = P->d;
  const int nl = d.depth + 2;
  std::vector<PackJob> jobs;
  std::vector<PackBlockJob> bjobs;
  for (int k = 0; k < d.n_flows; ++k) {
    const Block& b = P->blocks[k];
    for (int i = 0; i < 2; ++i)
      for (int net = 0; net < 2; ++net)
        for (int l = 0; l < nl; ++l) {
          const int li = ((k * 2 + i) * 2 + net) * nl + l;
          if (!linear_w[li] || !linear_b[li]) return IPOKE_INN_E_INVALID;
          PackJob j;
          std::memset(&j, 0, sizeof(j));
          j.w = linear_w[li];
          j.b = linear_b[li];
          j.dstT_hi = j.dstT_lo = -1;
          j.col_split = 1 << 30;
          if (l <= d.depth) {
            const Chain& ch = b.chains[b.mode == MODE_COND ? 0 : i];
            const LayerW& L = ch.L[l];
            const int g = b.mode == MODE_COND ? 2 * i + net : net;
            j.N = d.H;
            j.K = l == 0 ? (b.mode == MODE_COND ? d.E : (b.mode == MODE_NORMAL ? P->C2 + d.E : P->C2)) : d.H;
            if (l == 0 && b.mode == MODE_NORMAL) {   // cat(x_apply, cond): cond columns start at C2p in U
              j.col_split = P->C2;
              j.col_shift = P->C2p - P->C2;
            }
            j.ld_dst = L.Kp;
            j.dst_hi = L.hi / 2;
            j.dst_lo = L.lo >= 0 ? L.lo / 2 : -1;
            j.row0 = g * P->Hp;
            j.bias_off = L.bias / 4;
          } else {
            const Tail& t = b.tails[i];
            j.N = P->C2;
            j.K = d.H;
            j.ld_dst = t.w.Kp;
            j.dst_hi = t.w.hi / 2;
            j.dst_lo = t.w.lo >= 0 ? t.w.lo / 2 : -1;
            j.row0 = net * P->Np;
            j.bias_off = t.w.bias / 4;
          }
          jobs.push_back(j);
        }
    PackBlockJob bj;
    bj.loc = an_loc[k];
    bj.scale = an_scale[k];
    bj.perm_fwd = reinterpret_cast<const long long*>(perm_fwd[k]);
    bj.perm_bwd = reinterpret_cast<const long long*>(perm_bwd[k]);
    if (!bj.loc || !bj.scale || !bj.perm_fwd || !bj.perm_bwd) return IPOKE_INN_E_INVALID;
    bj.loc_off = b.loc / 4;
    bj.scale_off = b.scale / 4;
    bj.an_logdet_off = b.an_ld / 4;
    bj.map_fwd_off = b.map_fwd / 4;
    bj.map_rev_off = b.map_rev / 4;
    bjobs.push_back(bj);
  }
  char* pk = static_cast<char*>(packed);
  cudaError_t e = cudaMemsetAsync(pk, 0, (size_t)P->jobs_linear, st);   // zero padding of every matrix / bias
  if (e != cudaSuccess) return (int)e;
  // job tables: pageable host memory -> the copy is staged before the call returns
  e = cudaMemcpyAsync(pk + P->jobs_linear, jobs.data(), sizeof(PackJob) * jobs.size(), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemcpyAsync(pk + P->jobs_block, bjobs.data(), sizeof(PackBlockJob) * bjobs.size(), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return (int)e;
  const int max_elems = std::max(d.H * std::max(d.H, P->C2 + d.E), P->C2 * d.H);
  dim3 grid((unsigned)std::min(64, (max_elems + 255) / 256), (unsigned)jobs.size());
  inn_pack_linear_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const PackJob*>(pk + P->jobs_linear), packed);
  e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  inn_pack_block_kernel<<<d.n_flows, 256, 0, st>>>(reinterpret_cast<const PackBlockJob*>(pk + P->jobs_block), packed, d.C);
  e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_forward(ipoke_inn_plan* P, const void* packed, const float* x, const float* cond, float* z,
                      float* logdet, float* block_outs, float* block_logdets, int B, int block_begin, int block_end,
                      void* workspace, void* stream) {
  int rc = check_common(P, packed, x, z, B, block_begin, block_end, cond, workspace);
  if (rc) return rc;
  if (!logdet || x == z) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, block_begin, block_end, 0, (block_outs || block_logdets) ? 1 : 0};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, x, cond, z, logdet, block_outs, block_logdets, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_reverse(ipoke_inn_plan* P, const void* packed, const float* z, const float* cond, float* x, int B,
                      int block_begin, int block_end, void* workspace, void* stream) {
  int rc = check_common(P, packed, z, x, B, block_begin, block_end, cond, workspace);
  if (rc) return rc;
  if (x == z) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, block_begin, block_end, 1, 0};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, z, cond, x, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_forward_host(ipoke_inn_plan* P, const void* packed, const float* x_host, const float* cond_host,
                           float* z_host, float* logdet_host, int B, void* workspace, void* stream) {
  if (!P || !x_host || !z_host || !logdet_host || !workspace || B <= 0) return IPOKE_INN_E_INVALID;
  if (P->d.E > 0 && !cond_host) return IPOKE_INN_E_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const WsLayout wl = ws_layout(P, B);
  char* ws = static_cast<char*>(workspace);
  const size_t bc = sizeof(float) * (size_t)B * P->d.C;
  cudaError_t e = cudaMemcpyAsync(ws + wl.xin, x_host, bc, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return (int)e;
  if (P->d.E > 0) {
    e = cudaMemcpyAsync(ws + wl.condin, cond_host, sizeof(float) * (size_t)B * P->d.E, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
  }
  int rc = ipoke_inn_forward(P, packed, reinterpret_cast<float*>(ws + wl.xin),
                             P->d.E > 0 ? reinterpret_cast<float*>(ws + wl.condin) : nullptr,
                             reinterpret_cast<float*>(ws + wl.zout), reinterpret_cast<float*>(ws + wl.ldout), nullptr,
                             nullptr, B, 0, P->d.n_flows, workspace, stream);
  if (rc) return rc;
  e = cudaMemcpyAsync(z_host, ws + wl.zout, bc, cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemcpyAsync(logdet_host, ws + wl.ldout, sizeof(float) * (size_t)B, cudaMemcpyDeviceToHost, st);
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_reverse_host(ipoke_inn_plan* P, const void* packed, const float* z_host, const float* cond_host,
                           float* x_host, int B, void* workspace, void* stream) {
  if (!P || !z_host || !x_host || !workspace || B <= 0) return IPOKE_INN_E_INVALID;
  if (P->d.E > 0 && !cond_host) return IPOKE_INN_E_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const WsLayout wl = ws_layout(P, B);
  char* ws = static_cast<char*>(workspace);
  const size_t bc = sizeof(float) * (size_t)B * P->d.C;
  cudaError_t e = cudaMemcpyAsync(ws + wl.xin, z_host, bc, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return (int)e;
  if (P->d.E > 0) {
    e = cudaMemcpyAsync(ws + wl.condin, cond_host, sizeof(float) * (size_t)B * P->d.E, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
  }
  int rc = ipoke_inn_reverse(P, packed, reinterpret_cast<float*>(ws + wl.xin),
                             P->d.E > 0 ? reinterpret_cast<float*>(ws + wl.condin) : nullptr,
                             reinterpret_cast<float*>(ws + wl.zout), B, 0, P->d.n_flows, workspace, stream);
  if (rc) return rc;
  e = cudaMemcpyAsync(x_host, ws + wl.zout, bc, cudaMemcpyDeviceToHost, st);
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_actnorm_stats(const float* x, int B, int C, double* moments, void* stream) {
  if (!x || !moments || B <= 0 || C <= 0) return IPOKE_INN_E_INVALID;
  dim3 grid((unsigned)((C + 255) / 256), (unsigned)std::min(B, 128));
  inn_actnorm_stats_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, B, C, moments);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_actnorm_finalize(const double* moments, double count, int C, float* loc, float* scale, void* stream) {
  if (!moments || !loc || !scale || C <= 0 || count < 2.0) return IPOKE_INN_E_INVALID;
  inn_actnorm_finalize_kernel<<<(C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(moments, count, C, loc, scale);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_flow_loss(const float* z, const float* logdet, int B, int C, float nll_weight, float logdet_weight,
                        float* out3, float* dz, float* dlogdet, void* stream) {
  if (!z || !logdet || !out3 || B <= 0 || C <= 0) return IPOKE_INN_E_INVALID;
  inn_flow_loss_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(z, logdet, B, C, nll_weight, logdet_weight,
                                                                          out3, dz, dlogdet);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_last_launch_count(const ipoke_inn_plan* plan) { return plan ? plan->last_launches : 0; }

int ipoke_inn_set_option(ipoke_inn_plan* P, const char* name, int value) {
  if (!P || !name) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  if (std::strcmp(name, "graph") == 0) P->use_graph = value != 0;
  else if (std::strcmp(name, "pdl") == 0) { P->use_pdl = value != 0; P->programs.clear(); }
  else if (std::strcmp(name, "mcast") == 0) { P->use_mcast = value != 0; P->programs.clear(); }
  else return IPOKE_INN_E_INVALID;
  return 0;
}

int ipoke_inn_profile_enable(ipoke_inn_plan* P, int on) {
  if (!P) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  P->profile = on != 0;
  P->prof_kind.clear();
  return 0;
}

int ipoke_inn_profile_read(ipoke_inn_plan* P, double* ms3, int* launches3) {
  if (!P || !ms3 || !launches3) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  for (int k = 0; k < 3; ++k) { ms3[k] = 0.0; launches3[k] = 0; }
  const size_t n = P->prof_kind.size();
  if (n == 0) return 0;
  cudaError_t e = cudaEventSynchronize(P->prof_ev[2 * n - 1]);
  if (e != cudaSuccess) return (int)e;
  for (size_t i = 0; i < n; ++i) {
    float ms = 0.f;
    e = cudaEventElapsedTime(&ms, P->prof_ev[2 * i], P->prof_ev[2 * i + 1]);
    if (e != cudaSuccess) return (int)e;
    ms3[P->prof_kind[i]] += ms;
    launches3[P->prof_kind[i]] += 1;
  }
  return 0;
}

int ipoke_inn_debug_layout(const ipoke_inn_plan* P, int B, long long* out, int n) {
  if (!P || !out || n < 20 || B <= 0) return IPOKE_INN_E_INVALID;
  const WsLayout w = ws_layout(P, B);
  const long long v[20] = {w.S[0], w.S[1], w.U_hi, w.U_lo, w.ACT_hi[0], w.ACT_lo[0], w.ACT_hi[1], w.ACT_lo[1],
                           w.logdet, w.total, P->ldU, P->ldACT, P->Hp, P->C2p, P->Ep, P->Np, P->ksplit, P->Gmax,
                           P->packed_bytes, P->num_sms};
  for (int i = 0; i < 20; ++i) out[i] = v[i];
  return 0;
}

}  // extern "C"
