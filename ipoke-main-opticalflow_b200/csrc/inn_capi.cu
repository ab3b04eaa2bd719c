// libipoke_inn.so -- host side of the C ABI declared in include/ipoke_inn.h.
// Builds, per (batch, workspace, direction), a linear program of kernel launches:
//   boundary (gather + ActNorm + leaky)  ->  per half-step: (depth+1) x inn_gemm<EPI_HIDDEN>, inn_gemm<EPI_TAIL_*>
// and replays it on the caller's stream. No device allocation, no host synchronisation.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/ipoke_inn.h"
#include "elementwise.cuh"
#include "gemm_kernel.cuh"
#include "optim.cuh"
#include "tail_cluster_kernel.cuh"
#include "hidden_pair_kernel.cuh"
#include "tail_pair_kernel.cuh"
#include "wgrad_kernel.cuh"

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
  long long loc = -1, scale = -1, an_ld = -1, map_fwd = -1, map_rev = -1, map_fwd_inv = -1, map_rev_inv = -1;   // byte offsets
};

struct GemmOp {
  CUtensorMap tmA[2], tmW[2], tmO[2];
  GemmParams p;
  int mode, grid, smem, cluster;
  int tail_cluster;   // 1: inn_tail_cluster_kernel (small batches), cluster = 2 * p.ks
  int pair;           // 1: inn_hidden_pair_kernel (multi-wave hidden layers as tcgen05 CTA pairs), params in pp
  PairParams pp;
  int tail_pair;      // 1: inn_tail_pair_kernel (multi-wave unfolded coupling tails as CTA pairs), config in tp
  TailPairCfg tp;
};
struct WgradOp {
  CUtensorMap tmY[2], tmX[2];
  WgradParams p;
  int grid, smem;
  int pair;                 // 1: CTA-pair instantiation (cluster of 2)
  int lin[WG_MAX_GROUPS];   // index of each group's nn.Linear in the caller's gradient pointer arrays
};
enum OpKind { OP_BOUNDARY = 0, OP_GEMM = 1, OP_BWD_INIT = 2, OP_TAIL_BWD = 3, OP_BOUNDARY_BWD = 4, OP_WGRAD = 5,
              OP_ZERO_GRAD = 6 };
struct Op {
  int kind;
  BoundaryParams b;
  int b_grid;
  int gemm;  // index into gemm ops
  // which user pointers this boundary op takes (patched per call)
  int user_in, user_cond, user_out_final, block_index;
  int user_logdet;    // folded programs: the log-det accumulates in the caller's buffer (boundary and tail ops)
  // backward ops
  TailBwdParams tb;
  BoundaryBwdParams bb;
  int wgrad;          // index into wgrad ops
  int h0_block;       // OP_BOUNDARY_BWD: the block whose input was saved (0 -> the caller's x, k -> block_outs[k-1])
  int to_user_dx;     // OP_BOUNDARY_BWD: dst is the caller's dx
  int du_to_user_dcond;   // OP_GEMM (EPI_DU): du_cond is patched with the conditioning-gradient accumulator
  int zero_dcond;     // OP_BWD_INIT
  int lin, lin_rows, lin_cols;   // OP_ZERO_GRAD
};
struct ProgramKey {
  const void* packed; void* ws; int B, kb, ke, dir, want_blocks;
  const void* saved;
  int Bc = 0;   // multi-draw calls: rows of the conditioning (B = draws * Bc); 0 = one draw per conditioning row
  bool operator==(const ProgramKey& o) const {
    return packed == o.packed && ws == o.ws && B == o.B && kb == o.kb && ke == o.ke && dir == o.dir &&
           want_blocks == o.want_blocks && saved == o.saved && Bc == o.Bc;
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
  std::vector<WgradOp> wgrads;
  int launches = 0;
  std::vector<GraphEntry> graphs;     // instantiated CUDA graphs keyed by the caller's pointers
  int graph_misses = 0, graph_hits = 0;
  struct BwdGraph {                   // backward programs: keyed by every pointer the launches are patched with
    std::vector<const void*> key;
    cudaGraphExec_t exec = nullptr;
    unsigned long long last_use = 0;
  };
  std::vector<BwdGraph> bwd_graphs;
  ~Program() {
    for (auto& g : graphs)
      if (g.exec) cudaGraphExecDestroy(g.exec);
    for (auto& g : bwd_graphs)
      if (g.exec) cudaGraphExecDestroy(g.exec);
  }
};

struct WsLayout {
  long long S[2], U_hi, U_lo, U2_hi, U2_lo, ACT_hi[2], ACT_lo[2], logdet, xin, condin, zout, ldout, total;
  // backward scratch
  long long dt_hi[2], dt_lo[2], dcond, an_acc, an_cnt, dld_sum;
};
// What the training forward keeps for the backward pass (per flow block).
struct SavedBlock {
  long long U_hi[2], U_lo[2];                       // MLP input of half-step i (cond blocks: i = 0 only)
  std::vector<long long> act_hi[2], act_lo[2];      // [chain][layer] post-LeakyReLU activations [B, G*Hp]
  long long s[2], keep[2];                          // per half-step: tanh output and the kept half before the update
};
struct SavedLayout {
  std::vector<SavedBlock> blk;
  long long total = 0;
};

}  // namespace

struct ipoke_inn_plan {
  ipoke_inn_desc d;
  int device = 0;
  int ksplit = 2;
  int C2 = 0, C2p = 0, Ep = 0, Hp = 0, Np = 0, Npk = 0, ldU = 0, Gmax = 2, ldACT = 0;
  std::vector<Block> blocks;
  long long packed_bytes = 0;
  // device-side job tables live inside the packed buffer
  long long jobs_linear = -1, jobs_block = -1;
  long long w_arena_bytes = 0, lo_arena_off = -1;       // hi arena = [0, w_arena_bytes); lo arena right after it
  std::vector<long long> block_w_begin, block_w_end;    // bf16 element range of every flow block inside the hi arena
  // lazy device state
  bool dev_ready = false;
  int num_sms = 148;
  EncodeTiledFn encode = nullptr;
  std::mutex mu;
  std::vector<std::unique_ptr<Program>> programs;
  int last_launches = 0;
  // launch options: 1 = replay the program as a CUDA graph (captured on an internal stream), PDL attribute on GEMMs
  int use_graph = 1, use_pdl = 1;
  int tail_bn_max = 128;
  int tail_direct = 1;
  int streams = 1;                // calls the caller runs concurrently on different streams: one-wave launches are
                                  // sized for num_sms / streams SMs so that the concurrent kernels are co-resident
  int bwd_graph = 1;              // replay the backward pass as a CUDA graph when the caller's pointers repeat
  int use_pair = 1;               // multi-wave hidden layers as CTA pairs (hidden_pair_kernel.cuh)
  int pair_dgrad = 1;             // ... and the multi-wave dgrad GEMMs of the backward pass
  int pair_wgrad = 1;             // large-batch weight-gradient GEMMs as CTA pairs on 256 x 256 (or 256 x 128) tiles
  int pair_wgrad_bn = 256;
  int serpentine = 1;             // pair kernels of consecutive launches walk the row tiles in alternating directions
  int pair_tail = 1;              // multi-wave unfolded coupling tails as CTA pairs over the full N (tail_pair_kernel.cuh)
  int pair_tail_bn = 128;         // widest column tile of the pair tail: 256 covers the full N of cfg B (operands once per 128
                                  // rows: 302 MB instead of 572 MB through the L2 -> SM path, ncu r02) but fills TMEM with one
                                  // accumulator set, and the exposed epilogue gives the saving back (74 us either way); at 128
                                  // TMEM double-buffers and the launch measures 69 us against 74 (profiles/r02_ab_tail_pair.txt)
  int warp_slab = 1;              // hidden / dgrad epilogue: warp-private output slabs (no CTA barrier inside a tile)
  int tail_cluster = 1;           // small batches: coupling tails as split-K clusters (tail_cluster_kernel.cuh)
  int dbg = 0;
  int hidden_bn_force = 0;       // debug: fixed tile width of the hidden-layer GEMMs
  int widek_cover = 30;          // SM coverage target (%) of the hidden GEMMs when Hp > 1024 and ...
  int widek_mtiles = 8;          // ... the batch has at most this many 128-row tiles (one request at a time)
  unsigned long long* trace_buf = nullptr;   // ipoke_inn_debug_trace
  int trace_slots = 0;
  int use_fold = 1;              // inference programs: block boundaries folded into the coupling tails
  int fast_boundary = 512;       // boundary kernel: warp-per-row path from this many rows (0 = never)
  int fold_max_rows = -1;        // ... for batches up to this many rows (larger: separate warp-per-row boundary kernels);
                                 // -1 = by width: the folded scatter costs one 32-byte sector per element, which wide states
                                 // (C >= 256) pay for from ~2048 rows on (cfg B: -3 % at 2048, -11 % at 4096, -17 % at 8192, -2 % at
                                 // 32768; cfg A, C = 128: folding stays ahead) -- profiles/r01_ab_unfold_v10.txt
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

// The packed buffer is three arenas: [bf16 hi matrices of every layer | their lo residuals (parity mode) | fp32 biases,
// ActNorm state, gather maps, job tables].  A layer's hi and lo matrix sit at the same offset of their arenas, and the hi
// arena is ordered block by block -- for shapes without padding it is, element for element, the concatenation of the
// nn.Linear weights, which is what lets an optimizer shard own a contiguous span of it (ipoke_inn_param_layout).
struct ArenaCur { long long w = 0, small = 0; };
void layout_layer(ipoke_inn_plan* P, ArenaCur& cur, LayerW& L, int rows, int Kp) {
  L.rows = rows;
  L.Kp = Kp;
  L.hi = bump(cur.w, 2ll * rows * Kp);            // relative to the hi arena; rebased in ipoke_inn_plan_create
  L.lo = P->ksplit == 2 ? L.hi : -1;              // same offset inside the lo arena
  L.bias = bump(cur.small, 4ll * rows);           // relative to the small arena
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
  for (int i = 0; i < 2; ++i) {
    w.dt_hi[i] = bump(cur, 2ll * B * 2 * P->Npk);
    w.dt_lo[i] = bump(cur, 2ll * B * 2 * P->Npk);
  }
  w.dcond = bump(cur, 4ll * B * std::max(1, (int)P->d.E));
  w.an_acc = bump(cur, 4ll * 2 * C * P->d.n_flows);
  w.an_cnt = bump(cur, 4ll * P->d.n_flows);
  w.dld_sum = bump(cur, 16);
  w.U2_hi = bump(cur, 2ll * B * P->ldU);       // second MLP-input buffer (folded programs alternate per block)
  w.U2_lo = bump(cur, 2ll * B * P->ldU);
  w.total = round_up(cur, 1024);
  return w;
}

SavedLayout saved_layout(const ipoke_inn_plan* P, int B) {
  SavedLayout L;
  long long cur = 0;
  L.blk.resize(P->d.n_flows);
  for (int k = 0; k < P->d.n_flows; ++k) {
    const Block& b = P->blocks[k];
    SavedBlock& sb = L.blk[k];
    for (int i = 0; i < 2; ++i) {
      sb.U_hi[i] = sb.U_lo[i] = -1;
      if (i == 0 || b.mode != MODE_COND) {
        sb.U_hi[i] = bump(cur, 2ll * B * P->ldU);
        sb.U_lo[i] = P->ksplit == 2 ? bump(cur, 2ll * B * P->ldU) : -1;
      }
      sb.s[i] = bump(cur, 4ll * B * P->C2);
      sb.keep[i] = bump(cur, 4ll * B * P->C2);
    }
    for (int c = 0; c < b.n_chains; ++c)
      for (int l = 0; l <= P->d.depth; ++l) {
        sb.act_hi[c].push_back(bump(cur, 2ll * B * b.chains[c].G * P->Hp));
        sb.act_lo[c].push_back(P->ksplit == 2 ? bump(cur, 2ll * B * b.chains[c].G * P->Hp) : -1);
      }
  }
  L.total = round_up(cur, 1024);
  return L;
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
  SET_SMEM((inn_hidden_pair_kernel<1, false>))
  SET_SMEM((inn_hidden_pair_kernel<2, false>))
  SET_SMEM((inn_hidden_pair_kernel<1, true>))
  SET_SMEM((inn_hidden_pair_kernel<2, true>))
  SET_SMEM((inn_gemm_kernel<1, EPI_HIDDEN>))
  SET_SMEM((inn_gemm_kernel<1, EPI_HIDDEN, true>))
  SET_SMEM((inn_gemm_kernel<2, EPI_HIDDEN, true>))
  SET_SMEM((inn_gemm_kernel<1, EPI_DGRAD, true>))
  SET_SMEM((inn_gemm_kernel<2, EPI_DGRAD, true>))
  SET_SMEM((inn_gemm_kernel<2, EPI_HIDDEN>))
  SET_SMEM((inn_gemm_kernel<1, EPI_TAIL_FWD>))
  SET_SMEM((inn_gemm_kernel<2, EPI_TAIL_FWD>))
  SET_SMEM((inn_gemm_kernel<1, EPI_TAIL_REV>))
  SET_SMEM((inn_gemm_kernel<2, EPI_TAIL_REV>))
  SET_SMEM((inn_gemm_kernel<1, EPI_DGRAD>))
  SET_SMEM((inn_gemm_kernel<2, EPI_DGRAD>))
  SET_SMEM((inn_gemm_kernel<1, EPI_DU>))
  SET_SMEM((inn_gemm_kernel<2, EPI_DU>))
  SET_SMEM((inn_tail_pair_kernel<1, false>))
  SET_SMEM((inn_tail_pair_kernel<2, false>))
  SET_SMEM((inn_tail_pair_kernel<1, true>))
  SET_SMEM((inn_tail_pair_kernel<2, true>))
  SET_SMEM((inn_tail_cluster_kernel<1, false>))
  SET_SMEM((inn_tail_cluster_kernel<2, false>))
  SET_SMEM((inn_tail_cluster_kernel<1, true>))
  SET_SMEM((inn_tail_cluster_kernel<2, true>))
  SET_SMEM((inn_wgrad_kernel<1, false>))
  SET_SMEM((inn_wgrad_kernel<2, false>))
  SET_SMEM((inn_wgrad_kernel<1, true>))
  SET_SMEM((inn_wgrad_kernel<2, true>))
#undef SET_SMEM
  P->dev_ready = true;
  return 0;
}

int make_map(const ipoke_inn_plan* P, CUtensorMap* m, const void* base, long long inner, long long outer,
             long long row_stride_bytes, int box_outer, int box_inner = BK) {
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)row_stride_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  // 64 bf16 per box row = one 128-B swizzle row; 32 per row (warp-private half slabs) = 64-B swizzle
  CUresult r = P->encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, box_inner == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : IPOKE_INN_E_INTERNAL;
}
// slab_warp mode of a hidden / dgrad GEMM with `slab_bufs` slab buffers (0 = CTA-wide slabs)
inline int slab_warp_mode(const ipoke_inn_plan* P, int slab_bufs) {
  if (!P->warp_slab || slab_bufs < 1) return 0;
  return slab_bufs >= 2 ? 1 : 2;
}

struct TileCfg { int bn = 0, stages = 0, slab_bufs = 0, direct = 0; };

// smem plan for one tile width: pipeline stages + epilogue staging must fit in 227 KB
TileCfg tile_cfg(int ksplit, int mode, int bn, int direct = 0) {
  TileCfg t;
  t.bn = bn;
  t.direct = direct;
  const int avail = kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES;
  const int sb = gemm_stage_bytes(ksplit, mode, bn);
  auto stages_with = [&](int slabs) {
    const int epi = (gemm_epi_bytes(ksplit, mode, bn, slabs, direct) + 15) & ~15;
    return std::min(MAX_STAGES, (avail - epi) / sb);
  };
  if ((mode == EPI_HIDDEN || mode == EPI_DGRAD) && bn >= 64) {
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
TileCfg pick_bn(const ipoke_inn_plan* P, int mode, int rows_per_group, int tiles_other, int bn_max = 256,
                int cover_pct = 80) {
  static const int cand[5] = {256, 128, 64, 32, 16};
  TileCfg best;
  for (int i = 0; i < 5; ++i) {
    const int bn = cand[i];
    if (bn > bn_max || rows_per_group % bn != 0) continue;
    const TileCfg t = tile_cfg(P->ksplit, mode, bn);
    if (t.stages < 2) continue;
    best = t;
    if ((long long)tiles_other * (rows_per_group / bn) * 100 >= (long long)(P->num_sms / P->streams) * cover_pct) return t;
  }
  return best;
}

struct Builder {
  ipoke_inn_plan* P;
  Program* prog;
  const char* packed;
  char* ws;
  WsLayout wl;
  int B, m_tiles;
  // multi-draw calls (B = draws * Bc rows share Bc conditioning rows): the subnetworks of 'cond' couplings read only the
  // conditioning (flow_blocks.py:224, 237), so their hidden layers run on Bc rows and every draw's tail reads them
  int Bc = 0;
  int gemm_rows = 0, gemm_m_tiles = 0;     // rows / row tiles of the GEMM being built (B or Bc)
  int err = 0;
  // training forward / backward: activations live in the caller's `saved` buffer instead of the ping-pong scratch
  char* saved = nullptr;
  SavedLayout sl;
  bool save = false;
  int cur_block = 0;
  // inference programs: Shuffle / ActNorm / InvLeakyRelu between blocks run inside the tail epilogues
  bool fold = false;
  int f_kb = 0, f_ke = 0;

  template <typename T>
  T* sv(long long off) const { return off < 0 ? nullptr : reinterpret_cast<T*>(saved + off); }
  // MLP input buffer of (block, half-step)
  char* u_base(int k, int i, int sp) const {
    if (!save) {
      if (fold && (k & 1)) return ws + (sp == 0 ? wl.U2_hi : wl.U2_lo);
      return ws + (sp == 0 ? wl.U_hi : wl.U_lo);
    }
    const SavedBlock& sb = sl.blk[k];
    const int j = sb.U_hi[i] >= 0 ? i : 0;
    return saved + (sp == 0 ? sb.U_hi[j] : sb.U_lo[j]);
  }
  // activation buffer written by layer l of chain c
  char* act_base(int k, int c, int l, int sp) const {
    if (!save) return ws + (sp == 0 ? wl.ACT_hi[l & 1] : wl.ACT_lo[l & 1]);
    return saved + (sp == 0 ? sl.blk[k].act_hi[c][l] : sl.blk[k].act_lo[c][l]);
  }

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
  void set_u(BoundaryParams& b, int u_begin, int k, int i) {
    b.u_hi = reinterpret_cast<__nv_bfloat16*>(u_base(k, i, 0));
    b.u_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(u_base(k, i, 1)) : nullptr;
    b.ld_u = P->ldU;
    b.u_begin = u_begin;
    b.u_count = P->C2;
    b.u_pad = P->C2p;
  }

  // Alternate the row-tile direction of consecutive pair launches when one layer's activations are between one and two
  // L2 capacities: a layer then starts on the rows its producer wrote last, which are still cached (cfg B / 32768: 134 MB
  // per layer, step 12.02 -> 11.80 ms; cfg A / 32768 with 400 MB and cfg A / 4096 with 50 MB do not move)
  bool serpentine_ok(long long act_cols) const {
    const long long bytes = (long long)B * act_cols * 2 * P->ksplit;
    return P->serpentine && Bc == 0 && (P->serpentine > 1 || (bytes >= (96ll << 20) && bytes <= (256ll << 20)));
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
    g.p.dbg = P->dbg;
    if (P->trace_buf != nullptr && (int)prog->gemms.size() <= P->trace_slots)
      g.p.trace = P->trace_buf + 32ull * (prog->gemms.size() - 1);
    prog->launches++;
    return g;
  }
  void finish_gemm(GemmOp& g, const TileCfg& t) {
    g.p.M = gemm_rows;
    g.p.num_m_tiles = gemm_m_tiles;
    g.p.BN = t.bn;
    g.p.bn_log2 = 0;
    while ((1 << g.p.bn_log2) < t.bn) g.p.bn_log2++;
    g.p.stages = std::min(t.stages, std::max(2, g.p.num_k_blocks));
    int slab_bufs = t.slab_bufs;
    g.p.slab_bufs = slab_bufs;
    g.p.slab_warp = (g.mode == EPI_HIDDEN || g.mode == EPI_DGRAD) ? slab_warp_mode(P, slab_bufs) : 0;
    g.p.w_prefetch = P->w_prefetch;
    const int tiles = g.p.num_m_tiles * g.p.num_n_tiles;
    if (tiles <= P->num_sms && P->latency_smem_cap_kb > 0) {
      // optional: keep the footprint small so the next kernel's CTA (PDL) can be co-resident
      while (g.p.stages > 2 &&
             gemm_smem_bytes(P->ksplit, g.mode, t.bn, g.p.stages, slab_bufs, t.direct) > P->latency_smem_cap_kb * 1024)
        g.p.stages--;
    }
    g.p.direct = t.direct;
    const int acc_cols = gemm_nsub(g.mode) * t.bn;
    g.p.nacc = (2 * acc_cols <= 512) ? 2 : 1;
    int cols = 32;
    while (cols < g.p.nacc * acc_cols) cols <<= 1;
    g.p.tmem_cols = cols;
    g.p.epi_off = g.p.stages * gemm_stage_bytes(P->ksplit, g.mode, t.bn);
    g.smem = gemm_smem_bytes(P->ksplit, g.mode, t.bn, g.p.stages, slab_bufs, t.direct);
    g.grid = std::min(tiles, P->num_sms);
    g.cluster = 1;
    if (P->ksplit == 1) {
      g.tmA[1] = g.tmA[0];
      g.tmW[1] = g.tmW[0];
      g.tmO[1] = g.tmO[0];
    }
  }

  // hidden layers of one chain; returns the ACT buffer index holding the last hidden activation
  int chain(const Chain& c, int k, int ci, int ui, bool cond_only = false) {
    const int rows = c.G * P->Hp;
    const int B = (cond_only && Bc > 0) ? Bc : this->B;
    const int m_tiles = (B + BM - 1) / BM;
    gemm_rows = B;
    gemm_m_tiles = m_tiles;
    // Every CTA pulls a whole 128 x K activation panel through its L2 port: for wide hidden layers (K = Hp > 1024,
    // > 256 KB per panel) at one or two row tiles, covering 80 % of the SMs with narrow tiles multiplies that traffic --
    // cfg A (Hp 1536) batch 64 / 256: 48 CTAs of 64- / 128-wide tiles beat the 96 - 192 CTAs of the 80 % rule by 25 % /
    // 23 % per step, batch 512 / 1024 gain 6 % / 4 % (profiles/r01_ab_hidden_bn_v13.txt).  cfg B (Hp 1024) keeps the
    // 80 % rule it was tuned on, and so do plans sized for concurrent requests (streams > 1: measured 3 % slower).
    const int cover = (P->Hp > 1024 && m_tiles <= P->widek_mtiles && P->streams == 1) ? P->widek_cover : 80;
    TileCfg tc = pick_bn(P, EPI_HIDDEN, P->Hp, m_tiles * c.G, 256, cover);
    if (P->hidden_bn_force > 0 && P->Hp % P->hidden_bn_force == 0) tc = tile_cfg(P->ksplit, EPI_HIDDEN, P->hidden_bn_force);
    const int bn = tc.bn;
    const int a_box = BM;
    // multi-wave launches: CTA pairs (cta_group::2) on 256 x bn tiles, each CTA staging half of the W tile
    const int m_pairs = (m_tiles + 1) / 2;
    const bool pair = P->use_pair && (bn == 128 || bn == 256) && P->trace_buf == nullptr &&
                      (P->use_pair > 1 ? (long long)m_tiles * (rows / bn) > P->num_sms      // 2: every multi-wave launch
                                       : 2ll * m_tiles * (rows / bn) >= 5ll * P->num_sms);  // default: >= 2.5 waves (measured:
                                                                                           // 1.7 waves at cfg B / 4096 lose 3 %)
    for (size_t l = 0; l < c.L.size(); ++l) {
      const LayerW& L = c.L[l];
      GemmOp& g = add_gemm(EPI_HIDDEN);
      if (pair) {
        const long long ld_act = save ? rows : P->ldACT;
        PairParams& q = g.pp;
        g.pair = 1;
        q.M = B;
        q.num_m_pairs = m_pairs;
        q.num_n_tiles = rows / bn;
        q.tiles_per_group = P->Hp / bn;
        q.num_k_blocks = L.Kp / BK;
        q.BN = bn;
        q.a_group_stride = l == 0 ? 0 : P->Hp;
        q.a_m_wrap = 0;
        q.rev = serpentine_ok(rows) ? ((int)prog->gemms.size() & 1) : 0;
        q.bias = pk<float>(L.bias);
        q.slope = P->d.mlp_slope;
        q.nacc = 2 * bn <= 512 ? 2 : 1;
        q.tmem_cols = 32;
        while (q.tmem_cols < q.nacc * bn) q.tmem_cols <<= 1;
        // 32-column store boxes (16 KB of epilogue smem per split) when they buy a pipeline stage, else 64-column ones
        const int avail = kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES;
        const int sb = pair_stage_bytes(P->ksplit, bn);
        const int st_wide = (avail - pair_epi_bytes(P->ksplit, 1)) / sb, st_half = (avail - pair_epi_bytes(P->ksplit, 0)) / sb;
        q.wide = st_wide >= st_half ? 1 : 0;
        q.stages = std::max(2, std::min(std::min(MAX_STAGES, q.wide ? st_wide : st_half), std::max(2, q.num_k_blocks)));
        q.epi_off = q.stages * sb;
        g.smem = pair_smem_bytes(P->ksplit, bn, q.stages, q.wide);
        g.cluster = 2;
        g.grid = 2 * (int)std::min<long long>((long long)m_pairs * q.num_n_tiles, P->num_sms / 2);
        g.p.num_k_blocks = q.num_k_blocks;      // read by the profiling / accounting code
        for (int sp = 0; sp < P->ksplit; ++sp) {
          if (l == 0) err |= make_map(P, &g.tmA[sp], u_base(k, ui, sp) + 2ll * c.a_col_off, c.Kin_p, B, 2ll * P->ldU, BM);
          else err |= make_map(P, &g.tmA[sp], act_base(k, ci, (int)l - 1, sp), rows, B, 2ll * ld_act, BM);
          err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? L.hi : L.lo), L.Kp, L.rows, 2ll * L.Kp, bn / 2);
          err |= make_map(P, &g.tmO[sp], act_base(k, ci, (int)l, sp), rows, B, 2ll * ld_act, 32, q.wide ? BK : 32);
        }
        if (P->ksplit == 1) {
          g.tmA[1] = g.tmA[0];
          g.tmW[1] = g.tmW[0];
          g.tmO[1] = g.tmO[0];
        }
        continue;
      }
      g.p.num_n_tiles = rows / bn;
      g.p.tiles_per_group = P->Hp / bn;
      g.p.num_k_blocks = L.Kp / BK;
      g.p.a_group_stride = l == 0 ? 0 : P->Hp;
      g.p.bias = pk<float>(L.bias);
      const long long ld_act = save ? rows : P->ldACT;     // saved activations are stored densely per chain
      g.p.out_hi = reinterpret_cast<__nv_bfloat16*>(act_base(k, ci, (int)l, 0));
      g.p.out_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(act_base(k, ci, (int)l, 1)) : nullptr;
      g.p.ld_out = (int)ld_act;
      g.p.slope = P->d.mlp_slope;
      for (int sp = 0; sp < P->ksplit; ++sp) {
        if (l == 0) {
          err |= make_map(P, &g.tmA[sp], u_base(k, ui, sp) + 2ll * c.a_col_off, c.Kin_p, B, 2ll * P->ldU, a_box);
        } else {
          err |= make_map(P, &g.tmA[sp], act_base(k, ci, (int)l - 1, sp), rows, B, 2ll * ld_act, a_box);
        }
        err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? L.hi : L.lo), L.Kp, L.rows, 2ll * L.Kp, bn);
        // output slabs (TMA store): same 64 x 128 box as the next layer's A loads
        {                                                  // warp-private slabs: 64 x 32 or 32 x 32 boxes
          const int swm = slab_warp_mode(P, tc.slab_bufs);
          err |= make_map(P, &g.tmO[sp], act_base(k, ci, (int)l, sp), rows, B, 2ll * ld_act, swm ? 32 : BM, swm == 2 ? 32 : BK);
        }
      }
      finish_gemm(g, tc);
    }
    return (int)(c.L.size() - 1);
  }

  // act_layer: index of the last hidden layer of chain `ci` (its output feeds the tail)
  void tail(const Block& blk, int k, int i, int ci, int act_layer, int G, int state_buf, bool reverse, bool write_u,
            bool last_in_block) {
    const Tail& t = blk.tails[i];
    gemm_rows = B;
    gemm_m_tiles = m_tiles;
    const bool wrap = blk.mode == MODE_COND && Bc > 0 && Bc != B;   // the hidden activations hold Bc rows
    GemmOp& g = add_gemm(reverse ? EPI_TAIL_REV : EPI_TAIL_FWD);
    g.p.a_m_wrap = wrap ? Bc / BM : 0;
    Op& op = prog->ops.back();
    // two accumulators (s, t) per stage: cap the tile at 128 columns so TMEM still double-buffers and the
    // (transcendental-heavy) epilogue of tile i overlaps the MMAs of tile i+1
    TileCfg tc = pick_bn(P, g.mode, P->Np, m_tiles, P->tail_bn_max);
    // small batches: 128 x 16 tiles, each shared by a cluster of {s, t} x ks K-slices (tail_cluster_kernel.cuh)
    int ks = 0;
    if (P->tail_cluster) {
      const int tiles16 = m_tiles * (P->Np / TC_BN);
      for (int c = 4; c >= 1; c >>= 1)
        if ((P->Hp / BK) % c == 0 && tiles16 * 2 * c <= std::max(P->num_sms / P->streams, std::min(P->num_sms, tiles16 * 2))) { ks = c; break; }
      if (P->tail_cluster > 1 && ks > P->tail_cluster / 2) ks = P->tail_cluster / 2;   // debug: cap the cluster size
    }
    if (ks > 0) {
      tc = TileCfg();
      tc.bn = TC_BN;
    } else if (P->tail_direct == 2 || (P->tail_direct == 1 && m_tiles * (P->Np / tc.bn) <= P->num_sms)) {
      // register-resident epilogue: always for one-wave launches; tail_direct = 2 also for multi-wave launches, where
      // dropping the 128 x BN staging tile buys one more pipeline stage (2 x 64 KB in flight do not cover the TMA latency)
      tc = tile_cfg(P->ksplit, g.mode, tc.bn, 1);
    }
    const int bn = tc.bn;
    g.p.num_n_tiles = P->Np / bn;
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
    g.p.vec4 = (P->C2 % 4 == 0 && P->d.C % 4 == 0 && P->ldU % 4 == 0) ? 1 : 0;
    g.p.logdet = w<float>(wl.logdet);
    g.p.logdet_atomic = g.p.num_n_tiles > 1;
    if (write_u) {
      g.p.u_hi = reinterpret_cast<__nv_bfloat16*>(u_base(k, 1, 0));
      g.p.u_lo = reinterpret_cast<__nv_bfloat16*>(u_base(k, 1, 1));
      g.p.ld_u = P->ldU;
    }
    if (fold) {
      const ipoke_inn_desc& d = P->d;
      const int nk = reverse ? k - 1 : k + 1;                      // the block that consumes this one's output
      const bool has_next = reverse ? nk >= f_kb : nk < f_ke;
      g.p.f_ld = d.C;
      g.p.f_alpha = d.alpha;
      g.p.f_leaky = d.conditional;
      g.p.ld_u = P->ldU;
      if (has_next) {
        g.p.f_out = w<float>(wl.S[state_buf ^ 1]);
        g.p.f_u_hi = reinterpret_cast<__nv_bfloat16*>(u_base(nk, 0, 0));
        g.p.f_u_lo = reinterpret_cast<__nv_bfloat16*>(u_base(nk, 0, 1));
        g.p.f_u_begin = reverse ? P->C2 : 0;
      } else {
        op.user_out_final = 1;                                     // f_out = the caller's output
      }
      if (!reverse) {
        g.p.f_dst = pk<int>(blk.map_fwd_inv);
        g.p.f_mode = has_next ? 1 : 0;
        if (has_next) {
          g.p.f_loc = pk<float>(P->blocks[nk].loc);
          g.p.f_scale = pk<float>(P->blocks[nk].scale);
          if (last_in_block) g.p.ld_extra = pk<float>(P->blocks[nk].an_ld);
        }
        op.user_logdet = 1;
      } else {
        g.p.f_dst = has_next ? pk<int>(P->blocks[nk].map_rev_inv) : nullptr;   // nullptr = identity
        g.p.f_mode = 2;
        g.p.f_loc = pk<float>(blk.loc);
        g.p.f_scale = pk<float>(blk.scale);
      }
    }
    if (save && !reverse) {
      g.p.s_save = sv<float>(sl.blk[k].s[i]);
      g.p.keep_save = sv<float>(sl.blk[k].keep[i]);
      g.p.ld_save = P->C2;
    }
    const long long ld_act = save ? (long long)G * P->Hp : P->ldACT;
    for (int sp = 0; sp < P->ksplit; ++sp) {
      err |= make_map(P, &g.tmA[sp], act_base(k, ci, act_layer, sp), G * P->Hp, wrap ? Bc : B, 2ll * ld_act, BM);
      err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? t.w.hi : t.w.lo), t.w.Kp, t.w.rows, 2ll * t.w.Kp, bn);
      g.tmO[sp] = g.tmA[sp];      // unused by the tail epilogue
    }
    if (ks > 0) {
      g.p.M = B;
      g.p.num_m_tiles = m_tiles;
      g.p.BN = TC_BN;
      g.p.ks = ks;
      g.p.w_prefetch = P->w_prefetch;
      const int nkb = g.p.num_k_blocks / ks;
      const int fit = (kMaxDynSmem - 1024 - TC_RECV_BYTES - SMEM_BARRIER_BYTES) / tail_cluster_stage_bytes(P->ksplit);
      g.p.stages = std::max(1, std::min(std::min(nkb, MAX_STAGES), fit));
      g.smem = tail_cluster_smem_bytes(P->ksplit, g.p.stages);
      g.cluster = 2 * ks;
      g.grid = m_tiles * g.p.num_n_tiles * g.cluster;
      g.tail_cluster = 1;
      if (P->ksplit == 1) {
        g.tmA[1] = g.tmA[0];
        g.tmW[1] = g.tmW[0];
      }
      return;
    }
    finish_gemm(g, tc);
    // multi-wave unfolded tails: a CTA pair per 256 rows x the widest column tile that fits, so that every activation
    // panel goes through the L2 -> SM path once per pair instead of once per 128 x bn tile (tail_pair_kernel.cuh)
    if (P->pair_tail && !fold && P->trace_buf == nullptr && g.p.vec4 &&
        (P->pair_tail > 1 ? m_tiles >= 2          // 2 (debug / tests): every multi-tile unfolded tail
                          : (long long)m_tiles * g.p.num_n_tiles * 2 >= 3ll * P->num_sms)) {
      static const int cand[5] = {256, 128, 64, 32, 16};
      for (int ci2 = 0; ci2 < 5; ++ci2) {
        const int pbn = cand[ci2];
        if (P->Np % pbn != 0 || pbn > P->pair_tail_bn) continue;
        const int sb = tail_pair_stage_bytes(P->ksplit, pbn);
        const int st = (kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES) / sb;
        if (st < 2) continue;
        TailPairCfg& q = g.tp;
        q.num_m_pairs = (m_tiles + 1) / 2;
        q.BN = pbn;
        q.stages = std::min(std::min(st, MAX_STAGES), std::max(2, g.p.num_k_blocks));
        q.nacc = 4 * pbn <= 512 ? 2 : 1;
        q.rev = serpentine_ok((long long)G * P->Hp) ? ((int)prog->gemms.size() & 1) : 0;
        q.tmem_cols = 32;
        while (q.tmem_cols < q.nacc * 2 * pbn) q.tmem_cols <<= 1;
        g.tail_pair = 1;
        g.p.BN = pbn;
        g.p.num_n_tiles = P->Np / pbn;
        g.smem = tail_pair_smem_bytes(P->ksplit, pbn, q.stages);
        g.cluster = 2;
        g.grid = 2 * (int)std::min<long long>((long long)q.num_m_pairs * g.p.num_n_tiles, P->num_sms / 2);
        for (int sp = 0; sp < P->ksplit; ++sp)
          err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? t.w.hi : t.w.lo), t.w.Kp, t.w.rows, 2ll * t.w.Kp, pbn / 2);
        if (P->ksplit == 1) g.tmW[1] = g.tmW[0];
        break;
      }
    }
  }

  void coupling(const Block& blk, int k, int state_buf, bool reverse) {
    if (blk.mode == MODE_COND) {
      const int al = chain(blk.chains[0], k, 0, 0, true);
      tail(blk, k, reverse ? 1 : 0, 0, al, 4, state_buf, reverse, false, false);
      tail(blk, k, reverse ? 0 : 1, 0, al, 4, state_buf, reverse, false, true);
    } else {
      // forward: half-step 0 reads U(k,0) (written by the boundary) and writes U(k,1); reverse runs 1 then 0 and,
      // without `save`, both U views are the same scratch buffer
      const int first = reverse ? 1 : 0;
      int al = chain(blk.chains[first], k, first, reverse ? 0 : first);
      tail(blk, k, first, first, al, 2, state_buf, reverse, true, false);
      al = chain(blk.chains[1 - first], k, 1 - first, reverse ? 1 : 1 - first);
      tail(blk, k, 1 - first, 1 - first, al, 2, state_buf, reverse, false, true);
    }
  }

  int fold_limit() const {
    if (P->fold_max_rows >= 0) return P->fold_max_rows;
    return P->d.C >= 256 ? 2047 : (1 << 30);
  }

  void build_forward(int kb, int ke, bool want_blocks) {
    const ipoke_inn_desc& d = P->d;
    int cur = 0;
    fold = P->use_fold && !save && !want_blocks && B <= fold_limit();
    f_kb = kb;
    f_ke = ke;
    for (int k = kb; k < ke; ++k) {
      const Block& blk = P->blocks[k];
      if (fold && k > kb) {            // the previous block's tails already wrote this block's state and MLP input
        cur ^= 1;
        coupling(blk, k, cur, false);
        continue;
      }
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
      set_u(b, 0, k, 0);
      if (save && blk.mode != MODE_COND) {       // half-step 1 reads its own saved buffer: same cond columns / padding
        b.u2_hi = reinterpret_cast<__nv_bfloat16*>(u_base(k, 1, 0));
        b.u2_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(u_base(k, 1, 1)) : nullptr;
      }
      if (fold) {                                // both alternating MLP-input buffers get the cond columns / padding
        b.u2_hi = reinterpret_cast<__nv_bfloat16*>(u_base(k + 1, 0, 0));
        b.u2_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(u_base(k + 1, 0, 1)) : nullptr;
        o.user_logdet = 1;
      }
      if (save) o.user_cond = d.E > 0;
      if (d.E > 0) {
        b.E = d.E;
        b.ld_cond = d.E;
        b.u_cond_off = P->C2p;
        b.e_pad = P->Ep;
        b.cond_rows = (Bc > 0 && Bc != B) ? Bc : 0;
      }
      b.logdet = w<float>(wl.logdet);
      b.an_logdet = pk<float>(blk.an_ld);
      cur = nxt;
      coupling(blk, k, cur, false);
    }
    if (fold) return;
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
    fold = P->use_fold && !save && B <= fold_limit();
    f_kb = kb;
    f_ke = ke;
    for (int k = ke - 1; k >= kb; --k) {
      const Block& blk = P->blocks[k];
      if (fold && k < ke - 1) {
        cur ^= 1;
        coupling(blk, k, cur, true);
        continue;
      }
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
      set_u(b, P->C2, k, 0);
      if (fold) {
        b.u2_hi = reinterpret_cast<__nv_bfloat16*>(u_base(k + 1, 0, 0));
        b.u2_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(u_base(k + 1, 0, 1)) : nullptr;
      }
      if (d.E > 0) {
        b.E = d.E;
        b.ld_cond = d.E;
        b.u_cond_off = P->C2p;
        b.e_pad = P->Ep;
        b.cond_rows = (Bc > 0 && Bc != B) ? Bc : 0;
      }
      cur = nxt;
      coupling(blk, k, cur, true);
    }
    if (fold) return;
    Op& o = add_boundary();
    BoundaryParams& b = o.b;
    b.in = w<float>(wl.S[cur]);
    b.pre_loc = pk<float>(P->blocks[kb].loc);
    b.pre_scale = pk<float>(P->blocks[kb].scale);
    b.pre_leaky = d.conditional;
    o.user_out_final = 1;
  }
};


// ------------------------------------------------------------------ backward program
struct BwdBuilder : Builder {
  int nl() const { return P->d.depth + 2; }
  int lin_index(int k, int i, int net, int l) const { return ((k * 2 + i) * 2 + net) * nl() + l; }

  Op& add_op(int kind) {
    prog->ops.emplace_back();
    Op& o = prog->ops.back();
    std::memset(&o, 0, sizeof(Op));
    o.kind = kind;
    o.gemm = o.wgrad = -1;
    o.block_index = o.h0_block = -1;
    prog->launches++;
    return o;
  }
  int ew_grid(long long total) const { return (int)std::min<long long>((total + 255) / 256, (long long)P->num_sms * 8); }
  char* d_base(int idx, int sp) const { return ws + (sp == 0 ? wl.ACT_hi[idx] : wl.ACT_lo[idx]); }
  char* dt_base(int idx, int sp) const { return ws + (sp == 0 ? wl.dt_hi[idx] : wl.dt_lo[idx]); }

  void tail_bwd(int k, int i, int state_buf, int dt_idx) {
    Op& o = add_op(OP_TAIL_BWD);
    TailBwdParams& t = o.tb;
    t.dstate = w<float>(wl.S[state_buf]);
    t.ld_state = P->d.C;
    t.keep_off = (i == 0) ? P->C2 : 0;
    t.s_save = sv<float>(sl.blk[k].s[i]);
    t.keep_save = sv<float>(sl.blk[k].keep[i]);
    t.ld_save = P->C2;
    t.dt_hi = reinterpret_cast<__nv_bfloat16*>(dt_base(dt_idx, 0));
    t.dt_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(dt_base(dt_idx, 1)) : nullptr;
    t.ld_dt = 2 * P->Npk;
    t.npk = P->Npk;
    t.B = B;
    t.C2 = P->C2;
    o.b_grid = ew_grid((long long)B * P->Npk);
  }

  TileCfg pick_bn_bwd(int mode, int cols_per_group, int tiles_other) {
    static const int cand[3] = {256, 128, 64};
    TileCfg best;
    for (int c = 0; c < 3; ++c) {
      const int bn = cand[c];
      if (cols_per_group % bn != 0) continue;
      const TileCfg t = tile_cfg(P->ksplit, mode, bn);
      if (t.stages < 2) continue;
      best = t;
      if ((long long)tiles_other * (cols_per_group / bn) >= (long long)(P->num_sms * 4) / 5) return t;
    }
    return best;
  }

  // dX = dY * W (W read MN-major from the forward layout), grouped like the forward chain
  //   a_sp        : A operand (dY) base per split, [B, a_cols] row stride ld_a (elements)
  //   groups      : number of independent groups; contraction length per group kcon (multiple of 64)
  GemmOp& dgrad(int mode, char* const a_sp[2], long long a_cols, long long ld_a, int a_group_stride, const LayerW& W,
                int w_row0, int w_group_rows, int groups, int kcon, int out_cols_per_group) {
    GemmOp& g = add_gemm(mode);
    TileCfg tc = pick_bn_bwd(mode, out_cols_per_group, m_tiles * groups);
    const int bn = tc.bn;
    g.p.num_n_tiles = groups * (out_cols_per_group / bn);
    g.p.tiles_per_group = out_cols_per_group / bn;
    g.p.num_k_blocks = kcon / BK;
    g.p.a_group_stride = a_group_stride;
    g.p.w_mn = 1;
    g.p.w_row0[0] = w_row0;
    g.p.w_group_rows = w_group_rows;
    g.p.slope = P->d.mlp_slope;
    for (int sp = 0; sp < P->ksplit; ++sp) {
      err |= make_map(P, &g.tmA[sp], a_sp[sp], a_cols, B, 2ll * ld_a, BM);
      err |= make_map(P, &g.tmW[sp], packed + (sp == 0 ? W.hi : W.lo), W.Kp, W.rows, 2ll * W.Kp, 64);
      g.tmO[sp] = g.tmA[sp];
    }
    finish_gemm(g, tc);
    g.p.w_prefetch = 0;
    return g;
  }
  void set_dgrad_out(GemmOp& g, int d_idx, int out_col0, long long out_cols, char* const mask_sp0, long long ld_mask) {
    g.p.out_hi = reinterpret_cast<__nv_bfloat16*>(d_base(d_idx, 0));
    g.p.out_lo = P->ksplit == 2 ? reinterpret_cast<__nv_bfloat16*>(d_base(d_idx, 1)) : nullptr;
    g.p.ld_out = P->ldACT;
    g.p.out_col0 = out_col0;
    g.p.mask_act = reinterpret_cast<const __nv_bfloat16*>(mask_sp0);
    g.p.ld_mask = (int)ld_mask;
    g.p.mask_col0 = 0;
    for (int sp = 0; sp < P->ksplit; ++sp)
      err |= make_map(P, &g.tmO[sp], d_base(d_idx, sp), out_cols, B, 2ll * P->ldACT, g.p.slab_warp ? 32 : BM,
                      g.p.slab_warp == 2 ? 32 : BK);
    // multi-wave launches: the same CTA-pair kernel as the forward hidden layers (cta_group::2 on 256 x bn tiles, each CTA
    // staging its own 128 rows of dY and half of the MN-major W tile), with the LeakyReLU' mask in the epilogue
    const int bn = g.p.BN;
    const long long tiles = (long long)m_tiles * g.p.num_n_tiles;
    const bool pair = P->use_pair && P->pair_dgrad && (bn == 128 || bn == 256) && P->trace_buf == nullptr &&
                      (P->use_pair > 1 ? tiles > P->num_sms : 2 * tiles >= 5ll * P->num_sms);
    if (!pair) return;
    PairParams& q = g.pp;
    std::memset(&q, 0, sizeof(q));
    g.pair = 2;                                  // 2 = the dgrad instantiation
    q.M = B;
    q.num_m_pairs = (m_tiles + 1) / 2;
    q.num_n_tiles = g.p.num_n_tiles;
    q.tiles_per_group = g.p.tiles_per_group;
    q.num_k_blocks = g.p.num_k_blocks;
    q.BN = bn;
    q.a_group_stride = g.p.a_group_stride;
    q.a_m_wrap = 0;
    q.rev = serpentine_ok(P->ldACT) ? ((int)prog->gemms.size() & 1) : 0;
    q.bias = nullptr;
    q.slope = P->d.mlp_slope;
    q.w_row0 = g.p.w_row0[0];
    q.w_group_rows = g.p.w_group_rows;
    q.out_col0 = out_col0;
    q.mask_act = std::getenv("IPOKE_INN_DBG_NOMASK") ? nullptr : g.p.mask_act;   // timing experiment only
    q.ld_mask = g.p.ld_mask;
    q.mask_col0 = g.p.mask_col0;
    q.nacc = 2 * bn <= 512 ? 2 : 1;
    q.tmem_cols = 32;
    while (q.tmem_cols < q.nacc * bn) q.tmem_cols <<= 1;
    const int avail = kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES;
    const int sb = pair_stage_bytes(P->ksplit, bn);
    const int st_wide = (avail - pair_epi_bytes(P->ksplit, 1)) / sb, st_half = (avail - pair_epi_bytes(P->ksplit, 0)) / sb;
    q.wide = st_wide >= st_half ? 1 : 0;
    q.stages = std::max(2, std::min(std::min(MAX_STAGES, q.wide ? st_wide : st_half), std::max(2, q.num_k_blocks)));
    q.epi_off = q.stages * sb;
    g.smem = pair_smem_bytes(P->ksplit, bn, q.stages, q.wide);
    g.cluster = 2;
    g.grid = 2 * (int)std::min<long long>((long long)q.num_m_pairs * q.num_n_tiles, P->num_sms / 2);
    for (int sp = 0; sp < P->ksplit; ++sp)
      err |= make_map(P, &g.tmO[sp], d_base(d_idx, sp), out_cols, B, 2ll * P->ldACT, 32, q.wide ? BK : 32);
    if (P->ksplit == 1) g.tmO[1] = g.tmO[0];
  }

  // dW_g = dY_g^T X_g, db_g = colsum(dY_g) for the `groups` nets of one layer
  void wgrad(char* const y_sp[2], long long y_cols, long long ld_y, int dy_col0, int dy_group_stride, int rows_p,
             int rows_valid, char* const x_sp[2], long long x_cols, long long ld_x, int x_col0, int x_group_stride,
             int kin_p, int seg0_valid, int seg1_start, int seg1_valid, int ld_gw, int groups, const int* lin) {
    Op& o = add_op(OP_WGRAD);
    prog->wgrads.emplace_back();
    o.wgrad = (int)prog->wgrads.size() - 1;
    WgradOp& wgo = prog->wgrads.back();
    std::memset(&wgo, 0, sizeof(WgradOp));
    WgradParams& q = wgo.p;
    q.B = B;
    q.num_kb = (B + 63) / 64;
    q.G = groups;
    q.m_tiles = (rows_p + BM - 1) / BM;
    int bn = kin_p >= 128 ? 128 : 64;
    if (P->ksplit == 2 && kin_p >= 128) bn = 128;
    const int alias = 0;      // 128 x 256 tiles aliased onto the pipeline stages: parity green, no gain (r01 v11) -- not built
    // large batches: CTA pairs on 256 x 128 tiles (2.7 x fewer operand bytes per FLOP, wgrad_kernel.cuh) when the
    // launch fills at least one wave of pairs
    const bool pair = P->pair_wgrad && P->trace_buf == nullptr && q.num_kb >= 16 && kin_p % 128 == 0 && rows_p % 256 == 0 &&
                      (long long)groups * (rows_p / 256) * (kin_p / 128) * 2 >= P->num_sms;
    if (pair) {
      // 256 x 256 when the width allows: 32 KB per CTA and k-block for 8 MMAs balances one SM's ingest (~120 GB/s) against
      // the tensor pipe; 256 x 128 (24 KB for 4 MMAs) is still ingest-bound (measured: 48 -> 43 us per hidden x hidden launch)
      bn = (kin_p % 256 == 0 && P->pair_wgrad_bn >= 256) ? 256 : 128;
      q.m_tiles = rows_p / 256;                  // 256-row tiles: one per CTA pair
    }
    const int epi_bn = pair ? 128 : bn;          // staging tile width (256-wide pair tiles drain in two passes)
    wgo.pair = pair ? 1 : 0;
    q.BN = bn;
    q.epi_alias = alias;
    q.n_tiles = (kin_p + bn - 1) / bn;
    const int tiles = groups * q.m_tiles * q.n_tiles;
    int ks = 1;
    if (!pair && tiles * 2 <= P->num_sms && q.num_kb >= 8) {
      ks = std::min(P->num_sms / tiles, q.num_kb / 4);
      ks = std::max(ks, 1);
      const int per = (q.num_kb + ks - 1) / ks;
      ks = (q.num_kb + per - 1) / per;                  // no empty slice
    }
    q.kslices = ks;
    const int avail = kMaxDynSmem - 1024 - SMEM_BARRIER_BYTES - WG_ONES_BYTES - (alias ? 0 : ((wgrad_epi_bytes(epi_bn) + 15) & ~15));
    // a pair CTA stages half of the X^T columns: its stage is the one of a tile half as wide
    const int stage_b = wgrad_stage_bytes(P->ksplit, pair ? bn / 2 : bn);
    q.stages = std::max(2, std::min(MAX_STAGES, avail / stage_b));
    if (!alias) q.stages = std::min(q.stages, std::max(2, (q.num_kb + ks - 1) / ks));
    while (alias && q.stages * stage_b < wgrad_epi_bytes(bn)) q.stages++;   // room for the staging tile
    const int acc_cols = bn + 16;
    q.nacc = (2 * acc_cols <= 512) ? 2 : 1;
    int cols = 32;
    while (cols < q.nacc * acc_cols) cols <<= 1;
    q.tmem_cols = cols;
    q.dy_col0 = dy_col0;
    q.dy_group_stride = dy_group_stride;
    q.x_col0 = x_col0;
    q.x_group_stride = x_group_stride;
    q.rows_valid = rows_valid;
    q.kin_p = kin_p;
    q.seg0_valid = seg0_valid;
    q.seg1_start = seg1_start;
    q.seg1_valid = seg1_valid;
    q.ld_gw = ld_gw;
    for (int g = 0; g < groups; ++g) wgo.lin[g] = lin[g];
    for (int sp = 0; sp < P->ksplit; ++sp) {
      err |= make_map(P, &wgo.tmY[sp], y_sp[sp], y_cols, B, 2ll * ld_y, 64);
      err |= make_map(P, &wgo.tmX[sp], x_sp[sp], x_cols, B, 2ll * ld_x, 64);
    }
    if (P->ksplit == 1) {
      wgo.tmY[1] = wgo.tmY[0];
      wgo.tmX[1] = wgo.tmX[0];
    }
    wgo.smem = pair ? 1024 + q.stages * stage_b + WG_ONES_BYTES + ((wgrad_epi_bytes(epi_bn) + 15) & ~15) + SMEM_BARRIER_BYTES
                    : wgrad_smem_bytes(P->ksplit, bn, q.stages, alias);
    wgo.grid = pair ? 2 * std::min(tiles, P->num_sms / 2) : std::min(tiles * ks, P->num_sms);
    if (ks > 1) {
      // split-batch partial sums are added with atomics: clear the destinations first
      Op wg_op = prog->ops.back();
      prog->ops.pop_back();
      for (int g = 0; g < groups; ++g) {
        Op& z = add_op(OP_ZERO_GRAD);
        z.lin = lin[g];
        z.lin_rows = rows_valid;
        z.lin_cols = ld_gw;
      }
      prog->ops.push_back(wg_op);
    }
  }

  // backward of the hidden layers of one chain; `d_idx` holds dpre of the last hidden layer on entry
  void chain_bwd(const Block& blk, int k, int ci, int ui, const int* nets_i, const int* nets_net, int state_buf,
                 int du_state_off, int d_idx) {
    const Chain& c = blk.chains[ci];
    const int G = c.G;
    const long long rows = (long long)G * P->Hp;
    const int depth = P->d.depth;
    int lin[WG_MAX_GROUPS];
    for (int l = depth; l >= 0; --l) {
      char* y_sp[2] = {d_base(d_idx, 0), d_base(d_idx, 1)};
      for (int g = 0; g < G; ++g) lin[g] = lin_index(k, nets_i[g], nets_net[g], l);
      if (l > 0) {
        char* x_sp[2] = {act_base(k, ci, l - 1, 0), act_base(k, ci, l - 1, 1)};
        wgrad(y_sp, rows, P->ldACT, 0, P->Hp, P->Hp, P->d.H, x_sp, rows, rows, 0, P->Hp, P->Hp, P->d.H, 1 << 30, 0,
              P->d.H, G, lin);
        GemmOp& g = dgrad(EPI_DGRAD, y_sp, rows, P->ldACT, P->Hp, c.L[l], 0, P->Hp, G, P->Hp, P->Hp);
        set_dgrad_out(g, d_idx ^ 1, 0, rows, act_base(k, ci, l - 1, 0), rows);
        d_idx ^= 1;
      } else {
        char* x_sp[2] = {u_base(k, ui, 0), u_base(k, ui, 1)};
        const ipoke_inn_desc& d = P->d;
        int seg0 = 0, seg1s = 1 << 30, seg1v = 0, fan_in = 0;
        if (blk.mode == MODE_COND) { seg0 = d.E; fan_in = d.E; }
        else if (blk.mode == MODE_NORMAL) { seg0 = P->C2; seg1s = P->C2p; seg1v = d.E; fan_in = P->C2 + d.E; }
        else { seg0 = P->C2; fan_in = P->C2; }
        wgrad(y_sp, rows, P->ldACT, 0, P->Hp, P->Hp, d.H, x_sp, P->ldU, P->ldU, c.a_col_off, 0, c.Kin_p, seg0, seg1s,
              seg1v, fan_in, G, lin);
        // dU = sum over the G nets: one GEMM with contraction length G*Hp
        GemmOp& g = dgrad(EPI_DU, y_sp, rows, P->ldACT, 0, c.L[0], 0, 0, 1, (int)rows, c.Kin_p);
        g.p.du_state = w<float>(wl.S[state_buf]);
        g.p.du_ld_state = d.C;
        g.p.du_state_off = du_state_off;
        g.p.du_x_cols = blk.mode == MODE_COND ? 0 : P->C2;
        g.p.du_cond = nullptr;                        // patched per call (caller's dcond or the scratch accumulator)
        g.p.du_ld_cond = d.E;
        g.p.du_cond_col0 = blk.mode == MODE_COND ? 0 : P->C2p;
        g.p.du_E = d.E;
        prog->ops.back().du_to_user_dcond = d.E > 0;
      }
    }
  }

  // gradient of the last Linear of s_i / t_i: wgrad + dgrad into D[d_idx] at the chain's group columns
  void tail_grads(const Block& blk, int k, int i, int ci, int G, int dt_idx, int d_idx) {
    const Tail& t = blk.tails[i];
    const int depth = P->d.depth;
    char* y_sp[2] = {dt_base(dt_idx, 0), dt_base(dt_idx, 1)};
    char* x_sp[2] = {act_base(k, ci, depth, 0), act_base(k, ci, depth, 1)};
    const long long rows = (long long)G * P->Hp;
    const int lin[2] = {lin_index(k, i, 0, depth + 1), lin_index(k, i, 1, depth + 1)};
    wgrad(y_sp, 2ll * P->Npk, 2ll * P->Npk, 0, P->Npk, P->Np, P->C2, x_sp, rows, rows, t.g_s * P->Hp, P->Hp, P->Hp,
          P->d.H, 1 << 30, 0, P->d.H, 2, lin);
    GemmOp& g = dgrad(EPI_DGRAD, y_sp, 2ll * P->Npk, 2ll * P->Npk, P->Npk, t.w, 0, P->Np, 2, P->Npk, P->Hp);
    set_dgrad_out(g, d_idx, t.g_s * P->Hp, rows, act_base(k, ci, depth, 0), rows);
  }

  void coupling_bwd(const Block& blk, int k, int state_buf) {
    static const int kI_cond[4] = {0, 0, 1, 1}, kNet[4] = {0, 1, 0, 1};
    if (blk.mode == MODE_COND) {
      tail_bwd(k, 1, state_buf, 1);
      tail_bwd(k, 0, state_buf, 0);
      tail_grads(blk, k, 1, 0, 4, 1, 0);
      tail_grads(blk, k, 0, 0, 4, 0, 0);
      chain_bwd(blk, k, 0, 0, kI_cond, kNet, state_buf, 0, 0);
    } else {
      for (int i = 1; i >= 0; --i) {
        const int ii[2] = {i, i};
        tail_bwd(k, i, state_buf, 0);
        tail_grads(blk, k, i, i, 2, 0, 0);
        chain_bwd(blk, k, i, i, ii, kNet, state_buf, (i == 0) ? 0 : P->C2, 0);
      }
    }
  }

  void build_backward(int kb, int ke) {
    const ipoke_inn_desc& d = P->d;
    {
      Op& o = add_op(OP_BWD_INIT);
      o.zero_dcond = (d.E > 0 && ke == d.n_flows) ? 1 : 0;
      o.b_grid = ew_grid(std::max<long long>((long long)B * std::max(1, (int)d.E), 2ll * d.C * d.n_flows));
    }
    // un-shuffle the incoming gradient into the physical state layout of block ke-1
    {
      Op& o = add_boundary();
      o.user_in = 1;                                   // patched with dz
      o.b.src = pk<int>(P->blocks[ke - 1].map_fwd_inv);
      o.b.out = w<float>(wl.S[0]);
    }
    int cur = 0;
    for (int k = ke - 1; k >= kb; --k) {
      const Block& blk = P->blocks[k];
      coupling_bwd(blk, k, cur);
      Op& o = add_op(OP_BOUNDARY_BWD);
      BoundaryBwdParams& q = o.bb;
      q.g = w<float>(wl.S[cur]);
      q.ld_g = d.C;
      q.ld_h0 = d.C;
      q.loc = pk<float>(blk.loc);
      q.scale = pk<float>(blk.scale);
      q.leaky = d.conditional;
      q.alpha = d.alpha;
      q.ld_dst = d.C;
      q.acc = w<float>(wl.an_acc) + 2ll * d.C * k;
      q.counter = w<unsigned int>(wl.an_cnt) + k;
      q.dld_sum = w<float>(wl.dld_sum);
      q.B = B;
      q.C = d.C;
      o.h0_block = k;
      o.block_index = k;
      if (k > kb) {
        q.dst = w<float>(wl.S[cur ^ 1]);
        q.map = pk<int>(P->blocks[k - 1].map_fwd);
        cur ^= 1;
      } else {
        o.to_user_dx = 1;
      }
    }
  }
};

Program* get_program(ipoke_inn_plan* P, const ProgramKey& key, int* err) {
  for (auto& pr : P->programs)
    if (pr->key == key) return pr.get();
  std::unique_ptr<Program> prog(new Program());
  prog->key = key;
  BwdBuilder bld;
  bld.P = P;
  bld.prog = prog.get();
  bld.packed = static_cast<const char*>(key.packed);
  bld.ws = static_cast<char*>(key.ws);
  bld.wl = ws_layout(P, key.B);
  bld.B = key.B;
  bld.m_tiles = (key.B + BM - 1) / BM;
  bld.Bc = key.Bc;
  bld.gemm_rows = key.B;
  bld.gemm_m_tiles = bld.m_tiles;
  if (key.saved) {
    bld.saved = static_cast<char*>(const_cast<void*>(key.saved));
    bld.sl = saved_layout(P, key.B);
    bld.save = true;
  }
  // reserve so references stay valid while building
  const int per_block = 8 * (P->d.depth + 4) + 8;
  prog->ops.reserve((size_t)(key.ke - key.kb) * per_block + 8);
  prog->gemms.reserve((size_t)(key.ke - key.kb) * per_block);
  prog->wgrads.reserve((size_t)(key.ke - key.kb) * per_block);
  if (key.dir == 0) bld.build_forward(key.kb, key.ke, key.want_blocks != 0);
  else if (key.dir == 1) bld.build_reverse(key.kb, key.ke);
  else bld.build_backward(key.kb, key.ke);
  if (bld.err) {
    *err = IPOKE_INN_E_INTERNAL;
    return nullptr;
  }
  if (P->programs.size() >= 16) P->programs.erase(P->programs.begin());
  P->programs.push_back(std::move(prog));
  return P->programs.back().get();
}

template <int kSplit, int kMode, bool kWarpSlab = false>
cudaError_t launch_gemm_t(const GemmOp& g, const GemmParams& gp, cudaStream_t st, int pdl) {
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
  return cudaLaunchKernelEx(&cfg, inn_gemm_kernel<kSplit, kMode, kWarpSlab>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], g.tmO[0],
                            g.tmO[1], gp);
}

template <int kSplit, bool kRev>
cudaError_t launch_tail_cluster_t(const GemmOp& g, const GemmParams& gp, cudaStream_t st, int pdl) {
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)g.grid);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = (size_t)g.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = (unsigned)g.cluster;
  attr[na].val.clusterDim.y = 1;
  attr[na].val.clusterDim.z = 1;
  ++na;
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, inn_tail_cluster_kernel<kSplit, kRev>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], gp);
}

template <int kSplit>
cudaError_t launch_pair_t(const GemmOp& g, cudaStream_t st, int pdl) {
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
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = 2;
  attr[na].val.clusterDim.y = 1;
  attr[na].val.clusterDim.z = 1;
  ++na;
  cfg.attrs = attr;
  cfg.numAttrs = na;
  if (g.pair == 2)
    return cudaLaunchKernelEx(&cfg, inn_hidden_pair_kernel<kSplit, true>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], g.tmO[0],
                              g.tmO[1], g.pp);
  return cudaLaunchKernelEx(&cfg, inn_hidden_pair_kernel<kSplit, false>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], g.tmO[0],
                            g.tmO[1], g.pp);
}

template <int kSplit, bool kRev>
cudaError_t launch_tail_pair_t(const GemmOp& g, const GemmParams& gp, cudaStream_t st, int pdl) {
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
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = 2;
  attr[na].val.clusterDim.y = 1;
  attr[na].val.clusterDim.z = 1;
  ++na;
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, inn_tail_pair_kernel<kSplit, kRev>, g.tmA[0], g.tmA[1], g.tmW[0], g.tmW[1], gp, g.tp);
}

cudaError_t launch_gemm(const ipoke_inn_plan* P, const GemmOp& g, const GemmParams& gp, cudaStream_t st) {
  const int pdl = P->use_pdl;
  if (g.tail_pair) {
    const bool rev = g.mode == EPI_TAIL_REV;
    if (P->ksplit == 2)
      return rev ? launch_tail_pair_t<2, true>(g, gp, st, pdl) : launch_tail_pair_t<2, false>(g, gp, st, pdl);
    return rev ? launch_tail_pair_t<1, true>(g, gp, st, pdl) : launch_tail_pair_t<1, false>(g, gp, st, pdl);
  }
  if (g.pair) return P->ksplit == 2 ? launch_pair_t<2>(g, st, pdl) : launch_pair_t<1>(g, st, pdl);
  if (g.tail_cluster) {
    const bool rev = g.mode == EPI_TAIL_REV;
    if (P->ksplit == 2)
      return rev ? launch_tail_cluster_t<2, true>(g, gp, st, pdl) : launch_tail_cluster_t<2, false>(g, gp, st, pdl);
    return rev ? launch_tail_cluster_t<1, true>(g, gp, st, pdl) : launch_tail_cluster_t<1, false>(g, gp, st, pdl);
  }
  if (P->ksplit == 2) {
    switch (g.mode) {
      case EPI_HIDDEN: return gp.slab_warp ? launch_gemm_t<2, EPI_HIDDEN, true>(g, gp, st, pdl) : launch_gemm_t<2, EPI_HIDDEN>(g, gp, st, pdl);
      case EPI_TAIL_FWD: return launch_gemm_t<2, EPI_TAIL_FWD>(g, gp, st, pdl);
      case EPI_TAIL_REV: return launch_gemm_t<2, EPI_TAIL_REV>(g, gp, st, pdl);
      case EPI_DGRAD: return gp.slab_warp ? launch_gemm_t<2, EPI_DGRAD, true>(g, gp, st, pdl) : launch_gemm_t<2, EPI_DGRAD>(g, gp, st, pdl);
      default: return launch_gemm_t<2, EPI_DU>(g, gp, st, pdl);
    }
  }
  switch (g.mode) {
    case EPI_HIDDEN: return gp.slab_warp ? launch_gemm_t<1, EPI_HIDDEN, true>(g, gp, st, pdl) : launch_gemm_t<1, EPI_HIDDEN>(g, gp, st, pdl);
    case EPI_TAIL_FWD: return launch_gemm_t<1, EPI_TAIL_FWD>(g, gp, st, pdl);
    case EPI_TAIL_REV: return launch_gemm_t<1, EPI_TAIL_REV>(g, gp, st, pdl);
    case EPI_DGRAD: return gp.slab_warp ? launch_gemm_t<1, EPI_DGRAD, true>(g, gp, st, pdl) : launch_gemm_t<1, EPI_DGRAD>(g, gp, st, pdl);
    default: return launch_gemm_t<1, EPI_DU>(g, gp, st, pdl);
  }
}

template <typename... Args>
cudaError_t launch_ew(void (*kern)(Args...), int grid_x, int grid_y, int pdl, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid_x, (unsigned)grid_y);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

cudaError_t launch_wgrad(const ipoke_inn_plan* P, const WgradOp& wg, const WgradParams& q, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)wg.grid);
  cfg.blockDim = dim3(WG_THREADS);
  cfg.dynamicSmemBytes = (size_t)wg.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (P->use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (wg.pair) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  if (wg.pair) {
    if (P->ksplit == 2)
      return cudaLaunchKernelEx(&cfg, inn_wgrad_kernel<2, true>, wg.tmY[0], wg.tmY[1], wg.tmX[0], wg.tmX[1], q);
    return cudaLaunchKernelEx(&cfg, inn_wgrad_kernel<1, true>, wg.tmY[0], wg.tmY[1], wg.tmX[0], wg.tmX[1], q);
  }
  if (P->ksplit == 2)
    return cudaLaunchKernelEx(&cfg, inn_wgrad_kernel<2, false>, wg.tmY[0], wg.tmY[1], wg.tmX[0], wg.tmX[1], q);
  return cudaLaunchKernelEx(&cfg, inn_wgrad_kernel<1, false>, wg.tmY[0], wg.tmY[1], wg.tmX[0], wg.tmX[1], q);
}

// caller-owned pointers of one backward call
struct BwdArgs {
  const float* x; const float* block_outs; const float* dz; const float* dlogdet;
  float* dx; float* dcond;
  float* const* grad_w; float* const* grad_b; float* const* grad_loc; float* const* grad_scale;
};

int issue_backward(ipoke_inn_plan* P, Program* prog, const BwdArgs& a, cudaStream_t st) {
  const ipoke_inn_desc& d = P->d;
  const int B = prog->key.B;
  const long long BC = (long long)B * d.C;
  char* ws = static_cast<char*>(prog->key.ws);
  const WsLayout wl = ws_layout(P, B);
  float* dcond_acc = d.E > 0 ? (a.dcond ? a.dcond : reinterpret_cast<float*>(ws + wl.dcond)) : nullptr;
  const int pdl = P->use_pdl;
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
  for (Op& o : prog->ops) {
    int kind = 0;
    if (o.kind == OP_GEMM) kind = 3;
    else if (o.kind == OP_WGRAD) kind = 4;
    if (P->profile) { P->prof_kind[op_i] = kind; cudaEventRecord(P->prof_ev[2 * op_i], st); }
    cudaError_t e = cudaSuccess;
    switch (o.kind) {
      case OP_BWD_INIT: {
        float* z0 = o.zero_dcond ? dcond_acc : nullptr;
        const long long n0 = (long long)B * d.E;
        float* z1 = reinterpret_cast<float*>(ws + wl.an_acc);
        const long long n1 = (wl.an_cnt + 4ll * d.n_flows - wl.an_acc) / 4;
        e = launch_ew(inn_bwd_init_kernel, o.b_grid, 1, pdl, st, a.dlogdet, B, reinterpret_cast<float*>(ws + wl.dld_sum),
                      z0, n0, z1, n1);
        break;
      }
      case OP_BOUNDARY: {
        BoundaryParams b = o.b;
        b.in = a.dz;
        e = launch_ew(inn_boundary_kernel, o.b_grid, 1, pdl, st, (const BoundaryParams)b);
        break;
      }
      case OP_TAIL_BWD: {
        TailBwdParams t = o.tb;
        t.dlogdet = a.dlogdet;
        e = launch_ew(inn_tail_bwd_kernel, o.b_grid, 1, pdl, st, (const TailBwdParams)t);
        break;
      }
      case OP_GEMM: {
        GemmOp g = prog->gemms[o.gemm];
        if (o.du_to_user_dcond) g.p.du_cond = dcond_acc;
        e = launch_gemm(P, g, g.p, st);
        break;
      }
      case OP_ZERO_GRAD: {
        e = cudaMemsetAsync(a.grad_w[o.lin], 0, sizeof(float) * (size_t)o.lin_rows * o.lin_cols, st);
        if (e == cudaSuccess) e = cudaMemsetAsync(a.grad_b[o.lin], 0, sizeof(float) * (size_t)o.lin_rows, st);
        break;
      }
      case OP_WGRAD: {
        const WgradOp& wg = prog->wgrads[o.wgrad];
        WgradParams q = wg.p;
        for (int g = 0; g < q.G; ++g) {
          q.gw[g] = a.grad_w[wg.lin[g]];
          q.gb[g] = a.grad_b[wg.lin[g]];
        }
        e = launch_wgrad(P, wg, q, st);
        break;
      }
      case OP_BOUNDARY_BWD: {
        BoundaryBwdParams q = o.bb;
        q.h0 = o.h0_block == 0 ? a.x : a.block_outs + (long long)(o.h0_block - 1) * BC;
        if (o.to_user_dx) { q.dst = a.dx; q.map = nullptr; }
        q.gloc = a.grad_loc[o.block_index];
        q.gscale = a.grad_scale[o.block_index];
        const int gx = (d.C + 31) / 32;
        int gy = std::max(1, std::min((B + 63) / 64, (2 * P->num_sms + gx - 1) / gx));
        e = launch_ew(inn_boundary_bwd_kernel, gx, gy, pdl, st, (const BoundaryBwdParams)q);
        break;
      }
      default: return IPOKE_INN_E_INTERNAL;
    }
    if (e != cudaSuccess) return (int)e;
    if (P->profile) cudaEventRecord(P->prof_ev[2 * op_i + 1], st);
    ++op_i;
  }
  P->last_launches = prog->launches;
  return 0;
}

// Backward as a CUDA graph when every pointer of the call repeats (a steady training loop under the caching allocator):
// at the reference's batch of 64 the step is bound by the host issuing ~150 launches, not by the kernels.
int run_backward(ipoke_inn_plan* P, Program* prog, const BwdArgs& a, cudaStream_t st) {
  if (!P->use_graph || !P->bwd_graph || P->profile || prog->graph_misses > 8 + 4 * prog->graph_hits)
    return issue_backward(P, prog, a, st);
  std::vector<const void*> key;
  key.reserve(8 + 2 * prog->ops.size());
  key.push_back(a.x); key.push_back(a.block_outs); key.push_back(a.dz); key.push_back(a.dlogdet);
  key.push_back(a.dx); key.push_back(a.dcond);
  for (const Op& o : prog->ops) {
    if (o.kind == OP_ZERO_GRAD) { key.push_back(a.grad_w[o.lin]); key.push_back(a.grad_b[o.lin]); }
    else if (o.kind == OP_WGRAD) {
      const WgradOp& wg = prog->wgrads[o.wgrad];
      for (int g = 0; g < wg.p.G; ++g) { key.push_back(a.grad_w[wg.lin[g]]); key.push_back(a.grad_b[wg.lin[g]]); }
    } else if (o.kind == OP_BOUNDARY_BWD) { key.push_back(a.grad_loc[o.block_index]); key.push_back(a.grad_scale[o.block_index]); }
  }
  ++P->tick;
  for (auto& g : prog->bwd_graphs)
    if (g.key == key) {
      g.last_use = P->tick;
      ++prog->graph_hits;
      cudaError_t e = cudaGraphLaunch(g.exec, st);
      P->last_launches = prog->launches;
      return e == cudaSuccess ? 0 : (int)e;
    }
  ++prog->graph_misses;
  if (!P->cap_stream) {
    cudaError_t e = cudaStreamCreateWithFlags(&P->cap_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e = cudaStreamBeginCapture(P->cap_stream, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) return (int)e;
  const int rc = issue_backward(P, prog, a, P->cap_stream);
  cudaGraph_t graph = nullptr;
  e = cudaStreamEndCapture(P->cap_stream, &graph);
  cudaGraphExec_t exec = nullptr;
  if (rc == 0 && e == cudaSuccess && graph != nullptr) e = cudaGraphInstantiate(&exec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (rc != 0 || e != cudaSuccess || exec == nullptr) {
    cudaGetLastError();
    P->bwd_graph = 0;                 // capture unsupported here: plain launches for good
    return issue_backward(P, prog, a, st);
  }
  if (prog->bwd_graphs.size() >= 2) {
    size_t lru = prog->bwd_graphs[0].last_use <= prog->bwd_graphs[1].last_use ? 0 : 1;
    cudaGraphExecDestroy(prog->bwd_graphs[lru].exec);
    prog->bwd_graphs.erase(prog->bwd_graphs.begin() + lru);
  }
  Program::BwdGraph bg;
  bg.key = std::move(key);
  bg.exec = exec;
  bg.last_use = P->tick;
  prog->bwd_graphs.push_back(std::move(bg));
  e = cudaGraphLaunch(exec, st);
  return e == cudaSuccess ? 0 : (int)e;
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
      if (o.user_logdet) b.logdet = logdet_out;
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
      {
        // warp-per-row path (coalesced rows through smem) for larger batches when every access can be a 16-byte one
        auto a16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
        auto a8 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 7u) == 0; };
        const bool u_ok = b.u_hi == nullptr || (b.ld_u % 4 == 0 && b.u_begin % 4 == 0 && b.u_count % 4 == 0 && a8(b.u_hi) && a8(b.u_lo));
        if (P->fast_boundary && b.B >= P->fast_boundary && b.C % 4 == 0 && b.C <= 896 && b.ld_in % 4 == 0 && b.ld_out % 4 == 0 &&
            b.ld_raw % 4 == 0 && u_ok && a16(b.in) && a16(b.out) && a16(b.raw_out) && a16(b.raw_out2)) {
          b.fast_rows = 1;
          cfg.gridDim = dim3((unsigned)std::min((b.B + 7) / 8, P->num_sms * 8));
          cfg.dynamicSmemBytes = (size_t)13 * b.C * sizeof(float);
        }
      }
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = P->use_pdl ? 1 : 0;
      cudaError_t e = cudaLaunchKernelEx(&cfg, inn_boundary_kernel, b);
      if (e != cudaSuccess) return (int)e;
    } else {
      const GemmOp& g = prog->gemms[o.gemm];
      GemmParams gp = g.p;
      if (o.user_logdet) gp.logdet = logdet_out;
      if (o.user_out_final) gp.f_out = out;
      cudaError_t e = launch_gemm(P, g, gp, st);
      if (e != cudaSuccess) return (int)e;
    }
  }
  P->last_launches = prog->launches;
  return 0;
}

// Host-buffer calls (ipoke_inn_*_host): the copies around the kernels, so that a pinned-memory call replays as ONE graph
// (H2D nodes, kernels, D2H nodes) instead of one graph launch plus 3-5 separate copy calls per request.
struct HostIO {
  const float* in_host; const float* cond_host; float* out_host; float* ld_host;   // ld_host nullable (reverse)
  size_t in_bytes, cond_bytes, out_bytes, ld_bytes;
};
int issue_with_io(ipoke_inn_plan* P, Program* prog, const float* in, const float* cond, float* out, float* logdet_out,
                  float* block_outs, float* block_lds, cudaStream_t st, const HostIO* io) {
  if (io) {
    cudaError_t e = cudaMemcpyAsync(const_cast<float*>(in), io->in_host, io->in_bytes, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    if (io->cond_host) {
      e = cudaMemcpyAsync(const_cast<float*>(cond), io->cond_host, io->cond_bytes, cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return (int)e;
    }
  }
  int rc = issue_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st);
  if (rc) return rc;
  if (io) {
    cudaError_t e = cudaMemcpyAsync(io->out_host, out, io->out_bytes, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return (int)e;
    if (io->ld_host) {
      e = cudaMemcpyAsync(io->ld_host, logdet_out, io->ld_bytes, cudaMemcpyDeviceToHost, st);
      if (e != cudaSuccess) return (int)e;
    }
  }
  return 0;
}
bool host_pinned(const void* q) {
  if (!q) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, q) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

// Replay as a CUDA graph when the caller's pointers repeat (steady-state loops); plain launches otherwise.
int run_program(ipoke_inn_plan* P, Program* prog, const float* in, const float* cond, float* out, float* logdet_out,
                float* block_outs, float* block_lds, cudaStream_t st, const HostIO* io = nullptr) {
  if (!P->use_graph || P->profile || prog->graph_misses > 8 + 4 * prog->graph_hits)
    return issue_with_io(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st, io);
  // pageable host memory cannot be captured: such calls copy outside the graph
  if (io && !(host_pinned(io->in_host) && host_pinned(io->cond_host) && host_pinned(io->out_host) && host_pinned(io->ld_host))) {
    HostIO pre = *io;
    cudaError_t e = cudaMemcpyAsync(const_cast<float*>(in), pre.in_host, pre.in_bytes, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    if (pre.cond_host) {
      e = cudaMemcpyAsync(const_cast<float*>(cond), pre.cond_host, pre.cond_bytes, cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return (int)e;
    }
    int rc = run_program(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st, nullptr);
    if (rc) return rc;
    e = cudaMemcpyAsync(pre.out_host, out, pre.out_bytes, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return (int)e;
    if (pre.ld_host) e = cudaMemcpyAsync(pre.ld_host, logdet_out, pre.ld_bytes, cudaMemcpyDeviceToHost, st);
    return e == cudaSuccess ? 0 : (int)e;
  }
  const void* ptrs[6] = {in, cond, out, logdet_out, block_outs, block_lds};
  if (io) {                            // host calls are keyed by the host pointers (the device side is workspace scratch)
    ptrs[0] = io->in_host; ptrs[1] = io->cond_host; ptrs[2] = io->out_host; ptrs[3] = io->ld_host;
    ptrs[4] = in; ptrs[5] = out;
  }
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
  int rc = issue_with_io(P, prog, in, cond, out, logdet_out, block_outs, block_lds, P->cap_stream, io);
  cudaGraph_t graph = nullptr;
  e = cudaStreamEndCapture(P->cap_stream, &graph);
  if (rc != 0 || e != cudaSuccess || graph == nullptr) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    P->use_graph = 0;          // capture unsupported here: fall back to plain launches for good
    return issue_with_io(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st, io);
  }
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess || exec == nullptr) {
    cudaGetLastError();
    P->use_graph = 0;
    return issue_with_io(P, prog, in, cond, out, logdet_out, block_outs, block_lds, st, io);
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
    case IPOKE_INN_E_UNSUPPORTED: return "ipoke_inn: unsupported shape (C must be even and >= 2; H, depth, n_flows >= 1; E == 0 iff unconditional; 0 <= mlp_slope <= 1)";
    case IPOKE_INN_E_NO_DEVICE: return "ipoke_inn: no sm_100 CUDA device / driver entry point unavailable";
    case IPOKE_INN_E_INTERNAL: return "ipoke_inn: internal error (tensor map encode failed)";
    case IPOKE_INN_E_TOO_LARGE: return "ipoke_inn: B * C exceeds IPOKE_INN_FLOW_LOSS_STATIC_MAX_ELEMS (16777216), the size the library's fixed scratch array covers; use ipoke_inn_flow_loss_ws with caller-provided scratch";
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
  if (!(d.mlp_slope >= 0.f && d.mlp_slope <= 1.f)) return IPOKE_INN_E_UNSUPPORTED;   // LeakyReLU evaluated as max(a, slope * a)
  if (d.precision != IPOKE_INN_PREC_FP32_SPLIT && d.precision != IPOKE_INN_PREC_BF16) return IPOKE_INN_E_INVALID;
  ipoke_inn_plan* P = new (std::nothrow) ipoke_inn_plan();
  if (!P) return IPOKE_INN_E_INTERNAL;
  P->d = d;
  P->device = device;
  if (const char* e = std::getenv("IPOKE_INN_GRAPH")) P->use_graph = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_PDL")) P->use_pdl = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_FOLD")) P->use_fold = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_FOLD_MAXB")) P->fold_max_rows = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_FASTB")) P->fast_boundary = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_TAIL_BN")) P->tail_bn_max = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_HIDDEN_BN")) P->hidden_bn_force = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_WIDEK_COVER")) P->widek_cover = std::max(1, std::atoi(e));
  if (const char* e = std::getenv("IPOKE_INN_WIDEK_MT")) P->widek_mtiles = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_DBG")) P->dbg = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_DIRECT")) P->tail_direct = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_TAIL_CLUSTER")) P->tail_cluster = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_WARPSLAB")) P->warp_slab = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_PAIR")) P->use_pair = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_PAIR_DGRAD")) P->pair_dgrad = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_PAIR_TAIL")) P->pair_tail = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_SERPENTINE")) P->serpentine = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_PAIR_WGRAD")) P->pair_wgrad = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_PAIR_WGRAD_BN")) P->pair_wgrad_bn = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_PAIR_TAIL_BN")) P->pair_tail_bn = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_BWD_GRAPH")) P->bwd_graph = std::atoi(e);
  if (const char* e = std::getenv("IPOKE_INN_STREAMS")) P->streams = std::max(1, std::min(std::atoi(e), 8));
  if (const char* e = std::getenv("IPOKE_INN_WPREFETCH")) P->w_prefetch = std::atoi(e) != 0;
  if (const char* e = std::getenv("IPOKE_INN_STAGECAP_KB")) P->latency_smem_cap_kb = std::atoi(e);
  P->ksplit = d.precision == IPOKE_INN_PREC_FP32_SPLIT ? 2 : 1;
  P->C2 = d.C / 2;
  P->C2p = (int)round_up(P->C2, BK);
  P->Ep = d.E > 0 ? (int)round_up(d.E, BK) : 0;
  P->Hp = (int)round_up(d.H, BK);
  P->Np = (int)round_up(P->C2, 16);
  P->Npk = (int)round_up(P->Np, BK);
  P->ldU = P->C2p + P->Ep;
  P->Gmax = d.conditional ? 4 : 2;
  P->ldACT = P->Gmax * P->Hp;
  ArenaCur cur;
  P->blocks.resize(d.n_flows);
  P->block_w_begin.assign(d.n_flows, 0);
  P->block_w_end.assign(d.n_flows, 0);
  for (int k = 0; k < d.n_flows; ++k) {
    Block& b = P->blocks[k];
    cur.w = round_up(cur.w, 1024);
    P->block_w_begin[k] = cur.w / 2;
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
    P->block_w_end[k] = cur.w / 2;
    b.loc = bump(cur.small, 4ll * d.C);
    b.scale = bump(cur.small, 4ll * d.C);
    b.an_ld = bump(cur.small, 16);
    b.map_fwd = bump(cur.small, 4ll * d.C);
    b.map_rev = bump(cur.small, 4ll * d.C);
    b.map_fwd_inv = bump(cur.small, 4ll * d.C);
    b.map_rev_inv = bump(cur.small, 4ll * d.C);
  }
  const int n_lin = d.n_flows * 4 * (d.depth + 2);
  P->jobs_linear = bump(cur.small, (long long)sizeof(PackJob) * n_lin);
  P->jobs_block = bump(cur.small, (long long)sizeof(PackBlockJob) * d.n_flows);
  // rebase: hi arena at 0, lo arena after it, small arena last
  P->w_arena_bytes = round_up(cur.w, 1024);
  P->lo_arena_off = P->ksplit == 2 ? P->w_arena_bytes : -1;
  const long long small_off = P->w_arena_bytes * (P->ksplit == 2 ? 2 : 1);
  auto rebase = [&](LayerW& L) {
    if (L.lo >= 0) L.lo += P->lo_arena_off;
    L.bias += small_off;
  };
  for (Block& b : P->blocks) {
    for (int c = 0; c < b.n_chains; ++c)
      for (LayerW& L : b.chains[c].L) rebase(L);
    for (int i = 0; i < 2; ++i) {
      rebase(b.tails[i].w);
      b.tails[i].bias_t += small_off;
    }
    b.loc += small_off; b.scale += small_off; b.an_ld += small_off;
    b.map_fwd += small_off; b.map_rev += small_off; b.map_fwd_inv += small_off; b.map_rev_inv += small_off;
  }
  P->jobs_linear += small_off;
  P->jobs_block += small_off;
  P->packed_bytes = round_up(small_off + cur.small, 1024);
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
    bj.map_fwd_inv_off = b.map_fwd_inv / 4;
    bj.map_rev_inv_off = b.map_rev_inv / 4;
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
  inn_pack_linear_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const PackJob*>(pk + P->jobs_linear), packed, 0);
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
  ProgramKey key{packed, workspace, B, block_begin, block_end, 0, (block_outs || block_logdets) ? 1 : 0, nullptr};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, x, cond, z, logdet, block_outs, block_logdets, static_cast<cudaStream_t>(stream));
}

size_t ipoke_inn_saved_bytes(const ipoke_inn_plan* plan, int B) {
  if (!plan || B <= 0) return 0;
  return (size_t)saved_layout(plan, B).total;
}

int ipoke_inn_forward_train(ipoke_inn_plan* P, const void* packed, const float* x, const float* cond, float* z,
                            float* logdet, float* block_outs, float* block_logdets, int B, void* workspace,
                            void* saved, void* stream) {
  int rc = check_common(P, packed, x, z, B, 0, P ? P->d.n_flows : 0, cond, workspace);
  if (rc) return rc;
  if (!logdet || !block_outs || !saved || x == z) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, 0, P->d.n_flows, 0, 1, saved};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, x, cond, z, logdet, block_outs, block_logdets, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_backward(ipoke_inn_plan* P, const void* packed, const float* x, const float* cond,
                       const float* block_outs, const void* saved, const float* dz, const float* dlogdet, float* dx,
                       float* dcond, float* const* grad_w, float* const* grad_b, float* const* grad_loc,
                       float* const* grad_scale, int B, int block_begin, int block_end, void* workspace, void* stream) {
  int rc = check_common(P, packed, x, dz, B, block_begin, block_end, cond, workspace);
  if (rc) return rc;
  if (!block_outs || !saved || !dlogdet || !grad_w || !grad_b || !grad_loc || !grad_scale) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, block_begin, block_end, 2, 0, saved};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  BwdArgs a{x, block_outs, dz, dlogdet, dx, dcond, grad_w, grad_b, grad_loc, grad_scale};
  return run_backward(P, prog, a, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_reverse(ipoke_inn_plan* P, const void* packed, const float* z, const float* cond, float* x, int B,
                      int block_begin, int block_end, void* workspace, void* stream) {
  int rc = check_common(P, packed, z, x, B, block_begin, block_end, cond, workspace);
  if (rc) return rc;
  if (x == z) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, block_begin, block_end, 1, 0, nullptr};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, z, cond, x, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

// Multi-draw calls: draws * B rows, draw j of conditioning row b at row j * B + b.
static int check_draws(const ipoke_inn_plan* P, int B, int n_draws) {
  if (n_draws <= 0 || (long long)B * n_draws > 0x7fffffffll) return IPOKE_INN_E_INVALID;
  if (n_draws > 1 && (!P->d.conditional || P->d.E <= 0 || B % BM != 0)) return IPOKE_INN_E_UNSUPPORTED;
  return 0;
}

int ipoke_inn_reverse_draws(ipoke_inn_plan* P, const void* packed, const float* z, const float* cond, float* x, int B,
                            int n_draws, void* workspace, void* stream) {
  int rc = check_common(P, packed, z, x, B, 0, P ? P->d.n_flows : 0, cond, workspace);
  if (rc) return rc;
  if (x == z) return IPOKE_INN_E_INVALID;
  rc = check_draws(P, B, n_draws);
  if (rc) return rc;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B * n_draws, 0, P->d.n_flows, 1, 0, nullptr, n_draws > 1 ? B : 0};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, z, cond, x, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_forward_draws(ipoke_inn_plan* P, const void* packed, const float* x, const float* cond, float* z,
                            float* logdet, int B, int n_draws, void* workspace, void* stream) {
  int rc = check_common(P, packed, x, z, B, 0, P ? P->d.n_flows : 0, cond, workspace);
  if (rc) return rc;
  if (!logdet || x == z) return IPOKE_INN_E_INVALID;
  rc = check_draws(P, B, n_draws);
  if (rc) return rc;
  std::lock_guard<std::mutex> lock(P->mu);
  rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B * n_draws, 0, P->d.n_flows, 0, 0, nullptr, n_draws > 1 ? B : 0};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  return run_program(P, prog, x, cond, z, logdet, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int ipoke_inn_forward_host(ipoke_inn_plan* P, const void* packed, const float* x_host, const float* cond_host,
                           float* z_host, float* logdet_host, int B, void* workspace, void* stream) {
  if (!P || !packed || !x_host || !z_host || !logdet_host || !workspace || B <= 0) return IPOKE_INN_E_INVALID;
  if (P->d.E > 0 && !cond_host) return IPOKE_INN_E_INVALID;
  const WsLayout wl = ws_layout(P, B);
  char* ws = static_cast<char*>(workspace);
  const size_t bc = sizeof(float) * (size_t)B * P->d.C;
  float* x = reinterpret_cast<float*>(ws + wl.xin);
  float* cond = P->d.E > 0 ? reinterpret_cast<float*>(ws + wl.condin) : nullptr;
  float* z = reinterpret_cast<float*>(ws + wl.zout);
  float* ld = reinterpret_cast<float*>(ws + wl.ldout);
  std::lock_guard<std::mutex> lock(P->mu);
  int rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, 0, P->d.n_flows, 0, 0, nullptr};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  const HostIO io{x_host, P->d.E > 0 ? cond_host : nullptr, z_host, logdet_host, bc,
                  sizeof(float) * (size_t)B * std::max(0, (int)P->d.E), bc, sizeof(float) * (size_t)B};
  return run_program(P, prog, x, cond, z, ld, nullptr, nullptr, static_cast<cudaStream_t>(stream), &io);
}

int ipoke_inn_reverse_host(ipoke_inn_plan* P, const void* packed, const float* z_host, const float* cond_host,
                           float* x_host, int B, void* workspace, void* stream) {
  if (!P || !packed || !z_host || !x_host || !workspace || B <= 0) return IPOKE_INN_E_INVALID;
  if (P->d.E > 0 && !cond_host) return IPOKE_INN_E_INVALID;
  const WsLayout wl = ws_layout(P, B);
  char* ws = static_cast<char*>(workspace);
  const size_t bc = sizeof(float) * (size_t)B * P->d.C;
  float* z = reinterpret_cast<float*>(ws + wl.xin);
  float* cond = P->d.E > 0 ? reinterpret_cast<float*>(ws + wl.condin) : nullptr;
  float* x = reinterpret_cast<float*>(ws + wl.zout);
  std::lock_guard<std::mutex> lock(P->mu);
  int rc = ensure_device(P);
  if (rc) return rc;
  ProgramKey key{packed, workspace, B, 0, P->d.n_flows, 1, 0, nullptr};
  Program* prog = get_program(P, key, &rc);
  if (!prog) return rc;
  const HostIO io{z_host, P->d.E > 0 ? cond_host : nullptr, x_host, nullptr, bc,
                  sizeof(float) * (size_t)B * std::max(0, (int)P->d.E), bc, 0};
  return run_program(P, prog, z, cond, x, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream), &io);
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
                                                                          out3, dz, dlogdet, nullptr, nullptr);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_flow_loss_ref(const float* z, const float* logdet, const float* ref_sample, int B, int C, float nll_weight,
                            float logdet_weight, float* out4, float* dz, float* dlogdet, void* stream) {
  if (!z || !logdet || !ref_sample || !out4 || B <= 0 || C <= 0) return IPOKE_INN_E_INVALID;
  inn_flow_loss_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(z, logdet, B, C, nll_weight, logdet_weight,
                                                                          out4, dz, dlogdet, ref_sample, nullptr);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_flow_loss_scratch_bytes(void) { return FLOW_LOSS_SCRATCH_BYTES; }

int ipoke_inn_flow_loss_ws(const float* z, const float* logdet, const float* ref_sample, int B, int C, float nll_weight,
                           float logdet_weight, float* out4, float* dz, float* dlogdet, void* scratch, void* stream) {
  if (!z || !logdet || !out4 || B <= 0 || C <= 0) return IPOKE_INN_E_INVALID;
  if (scratch && (reinterpret_cast<uintptr_t>(scratch) & 7u)) return IPOKE_INN_E_INVALID;
  const long long total = (long long)B * C;
  int blocks = scratch ? (int)std::min<long long>(FLOW_LOSS_MAX_BLOCKS, (total + 32767) / 32768) : 1;
  blocks = std::max(blocks, 1);
  inn_flow_loss_kernel<<<blocks, 1024, 0, static_cast<cudaStream_t>(stream)>>>(
      z, logdet, B, C, nll_weight, logdet_weight, out4, dz, dlogdet, ref_sample, static_cast<double*>(scratch));
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_adam_chunk_elems(void) { return ADAM_CHUNK; }

int ipoke_inn_adam_step(const ipoke_inn_adam_job* jobs_dev, const int* chunk_job_dev, const long long* chunk_off_dev,
                        int n_chunks, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                        int amsgrad, float grad_scale, void* stream) {
  if (n_chunks == 0) return 0;
  if (!jobs_dev || !chunk_job_dev || !chunk_off_dev || n_chunks < 0 || step < 1) return IPOKE_INN_E_INVALID;
  AdamParams a;
  a.jobs = jobs_dev;
  a.chunk_job = chunk_job_dev;
  a.chunk_off = chunk_off_dev;
  a.n_chunks = n_chunks;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.weight_decay = weight_decay;
  // bias corrections in double like Python floats in torch/optim/adam.py
  const double bc1 = 1.0 - std::pow((double)beta1, (double)step);
  const double bc2 = 1.0 - std::pow((double)beta2, (double)step);
  a.step_size = (float)((double)lr / bc1);
  a.inv_sqrt_bc2 = (float)(1.0 / std::sqrt(bc2));
  a.grad_scale = grad_scale;
  a.amsgrad = amsgrad;
  inn_adam_kernel<<<n_chunks, ADAM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(a);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

// ---- packed-master training (ShardedAdam): the packed buffer as the parameter store ------------------------------
int ipoke_inn_param_layout(const ipoke_inn_plan* P, long long* out, int n) {
  if (!P || !out) return IPOKE_INN_E_INVALID;
  const ipoke_inn_desc& d = P->d;
  const int nl = d.depth + 2, n_lin = d.n_flows * 4 * nl;
  if (n < 8 + n_lin + 2 * d.n_flows) return IPOKE_INN_E_INVALID;
  // direct = every nn.Linear weight is one unpadded, contiguous [out, in] span of the hi arena and the spans tile it
  bool direct = d.H == P->Hp && P->C2 == P->C2p && P->C2 == P->Np && (d.E == 0 || d.E == P->Ep);
  long long covered = 0;
  for (int k = 0; k < d.n_flows; ++k) {
    const Block& b = P->blocks[k];
    for (int i = 0; i < 2; ++i)
      for (int net = 0; net < 2; ++net)
        for (int l = 0; l < nl; ++l) {
          const int li = ((k * 2 + i) * 2 + net) * nl + l;
          long long off, numel;
          if (l <= d.depth) {
            const LayerW& L = b.chains[b.mode == MODE_COND ? 0 : i].L[l];
            const int g = b.mode == MODE_COND ? 2 * i + net : net;
            off = L.hi / 2 + (long long)g * P->Hp * L.Kp;
            numel = (long long)P->Hp * L.Kp;
          } else {
            const LayerW& L = b.tails[i].w;
            off = L.hi / 2 + (long long)net * P->Np * L.Kp;
            numel = (long long)P->Np * L.Kp;
          }
          out[8 + li] = off;
          covered += numel;
        }
    out[8 + n_lin + 2 * k] = P->block_w_begin[k];
    out[8 + n_lin + 2 * k + 1] = P->block_w_end[k];
  }
  if (covered != P->w_arena_bytes / 2) direct = false;        // alignment gaps between matrices
  out[0] = direct ? 1 : 0;
  out[1] = 0;                                                  // hi arena byte offset
  out[2] = P->w_arena_bytes / 2;                               // bf16 elements of the hi arena
  out[3] = P->lo_arena_off;                                    // byte offset of the lo arena, -1 in the bf16 mode
  out[4] = n_lin;
  out[5] = d.n_flows;
  out[6] = out[7] = 0;
  return 0;
}

int ipoke_inn_pack_small(const ipoke_inn_plan* cplan, void* packed, void* stream) {
  ipoke_inn_plan* P = const_cast<ipoke_inn_plan*>(cplan);
  if (!P || !packed) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  int rc = ensure_device(P);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* pk = static_cast<char*>(packed);
  const int n_lin = P->d.n_flows * 4 * (P->d.depth + 2);
  inn_pack_linear_kernel<<<dim3(1, (unsigned)n_lin), 256, 0, st>>>(reinterpret_cast<const PackJob*>(pk + P->jobs_linear), packed, 1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  inn_pack_block_kernel<<<P->d.n_flows, 256, 0, st>>>(reinterpret_cast<const PackBlockJob*>(pk + P->jobs_block), packed, P->d.C);
  e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

static int shard_grid(long long n) {
  const long long per_cta = 2ll * 4 * ADAM_THREADS * 4;         // 4 double-float4 iterations per thread
  return (int)std::max<long long>(1, std::min<long long>((n + per_cta - 1) / per_cta, 148 * 16));
}
static bool al16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

int ipoke_inn_adam_shard_step(float* p, const float* g, const void* g_bf16, float* m, float* v, float* vmax, void* hi,
                              void* lo, long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                              int step, int amsgrad, float grad_scale, void* stream) {
  if (n == 0) return 0;
  if (!p || (!g) == (!g_bf16) || !m || !v || (!hi && lo) || (amsgrad && !vmax) || n < 0 || (n & 3) || step < 1) return IPOKE_INN_E_INVALID;
  if (!al16(p) || !al16(g) || !al16(m) || !al16(v) || !al16(vmax) || (reinterpret_cast<uintptr_t>(g_bf16) & 7u) ||
      (reinterpret_cast<uintptr_t>(hi) & 7u) || (reinterpret_cast<uintptr_t>(lo) & 7u)) return IPOKE_INN_E_INVALID;
  AdamShardParams q;
  q.p = p; q.g = g; q.g16 = static_cast<const __nv_bfloat16*>(g_bf16);
  q.m = m; q.v = v; q.vmax = vmax;
  q.hi = static_cast<__nv_bfloat16*>(hi); q.lo = static_cast<__nv_bfloat16*>(lo);
  q.n = n;
  std::memset(&q.a, 0, sizeof(q.a));
  q.a.beta1 = beta1; q.a.beta2 = beta2; q.a.eps = eps; q.a.weight_decay = weight_decay;
  const double bc1 = 1.0 - std::pow((double)beta1, (double)step);
  const double bc2 = 1.0 - std::pow((double)beta2, (double)step);
  q.a.step_size = (float)((double)lr / bc1);
  q.a.inv_sqrt_bc2 = (float)(1.0 / std::sqrt(bc2));
  q.a.grad_scale = grad_scale;
  q.a.amsgrad = amsgrad;
  inn_adam_shard_kernel<<<shard_grid(n), ADAM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(q);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_master_to_packed(const float* p, void* hi, void* lo, long long n, void* stream) {
  if (n == 0) return 0;
  if (!p || !hi || n < 0 || (n & 3) || !al16(p) || (reinterpret_cast<uintptr_t>(hi) & 7u) || (reinterpret_cast<uintptr_t>(lo) & 7u))
    return IPOKE_INN_E_INVALID;
  inn_master_split_kernel<<<shard_grid(n), ADAM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      p, static_cast<__nv_bfloat16*>(hi), static_cast<__nv_bfloat16*>(lo), n);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_grad_to_bf16(const float* g, void* out_bf16, long long n, float scale, void* stream) {
  if (n == 0) return 0;
  if (!g || !out_bf16 || n < 0 || (n & 3) || !al16(g) || (reinterpret_cast<uintptr_t>(out_bf16) & 7u)) return IPOKE_INN_E_INVALID;
  inn_grad_to_bf16_kernel<<<shard_grid(n), ADAM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      g, static_cast<__nv_bfloat16*>(out_bf16), n, scale);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

int ipoke_inn_last_launch_count(const ipoke_inn_plan* plan) { return plan ? plan->last_launches : 0; }

int ipoke_inn_set_option(ipoke_inn_plan* P, const char* name, int value) {
  if (!P || !name) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  if (std::strcmp(name, "graph") == 0) P->use_graph = value != 0;
  else if (std::strcmp(name, "bwd_graph") == 0) P->bwd_graph = value != 0;
  else if (std::strcmp(name, "pdl") == 0) { P->use_pdl = value != 0; P->programs.clear(); }
  else if (std::strcmp(name, "fold") == 0) { P->use_fold = value != 0; P->programs.clear(); }
  else if (std::strcmp(name, "streams") == 0) { P->streams = std::max(1, std::min(value, 8)); P->programs.clear(); }
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
    if (P->prof_kind[i] > 2) continue;           // backward kinds: ipoke_inn_profile_read_n
    ms3[P->prof_kind[i]] += ms;
    launches3[P->prof_kind[i]] += 1;
  }
  return 0;
}

int ipoke_inn_profile_read_n(ipoke_inn_plan* P, double* ms, int* launches, int n) {
  if (!P || !ms || !launches || n < 1) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  for (int k = 0; k < n; ++k) { ms[k] = 0.0; launches[k] = 0; }
  const size_t cnt = P->prof_kind.size();
  if (cnt == 0) return 0;
  cudaError_t e = cudaEventSynchronize(P->prof_ev[2 * cnt - 1]);
  if (e != cudaSuccess) return (int)e;
  for (size_t i = 0; i < cnt; ++i) {
    float t = 0.f;
    e = cudaEventElapsedTime(&t, P->prof_ev[2 * i], P->prof_ev[2 * i + 1]);
    if (e != cudaSuccess) return (int)e;
    const int k = P->prof_kind[i];
    if (k < n) { ms[k] += t; launches[k] += 1; }
  }
  return 0;
}

int ipoke_inn_debug_trace(ipoke_inn_plan* P, void* dev_buf, int n_slots) {
  if (!P) return IPOKE_INN_E_INVALID;
  std::lock_guard<std::mutex> lock(P->mu);
  P->trace_buf = static_cast<unsigned long long*>(dev_buf);
  P->trace_slots = dev_buf ? n_slots : 0;
  P->programs.clear();
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
