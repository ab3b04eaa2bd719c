/* ipoke_inn.h -- C ABI of the B200-native flat-INN hot path (libipoke_inn.so).
 *
 * Drop-in boundary underneath the reference's nn.Module interface
 * (reference: models/modules/INN/INN.py:19-88 SupervisedTransformer, :250-296
 * UnsupervisedTransformer3; models/modules/INN/flow_blocks.py:10-66 ConditionalFlow,
 * :432-465 UnconditionalFlow2).  The reference has no FFI of its own for this path:
 * every entry point below replaces a span of eager PyTorch ops, cited per function.
 *
 * Conventions
 *   - plain C types only; every pointer is a DEVICE pointer unless the name ends in
 *     _host or the comment says "host array";
 *   - tensors are fp32, row-major, contiguous: x/z [B,C], cond [B,E], logdet [B];
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*),
 *     allocates nothing on the device and never synchronises the host;
 *   - return value: 0 = ok, < 0 = IPOKE_INN_E_*, > 0 = a cudaError_t value;
 *     ipoke_inn_last_error_string() decodes both.  Nothing throws or aborts.
 *   - a plan's architecture and layouts are fixed at creation; its launch-program and CUDA-graph caches are filled
 *     lazily and guarded by a mutex held for the duration of each entry point, so calls on one plan are serialised on
 *     the host (concurrency comes from issuing them on different streams, each with its own workspace).
 */
#ifndef IPOKE_INN_H_
#define IPOKE_INN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IPOKE_INN_VERSION 111 /* major*100 + minor */

enum {
  IPOKE_INN_OK = 0,
  IPOKE_INN_E_INVALID = -1,     /* bad argument (null pointer, B <= 0, bad block range ...) */
  IPOKE_INN_E_UNSUPPORTED = -2, /* shape outside what the kernels cover (odd C, E > 0 unconditional ...) */
  IPOKE_INN_E_NO_DEVICE = -3,   /* no sm_100 device / driver entry point missing */
  IPOKE_INN_E_INTERNAL = -4,
  IPOKE_INN_E_TOO_LARGE = -5    /* ipoke_inn_flow_loss_static only: B * C beyond IPOKE_INN_FLOW_LOSS_STATIC_MAX_ELEMS */
};

enum { IPOKE_INN_PREC_FP32_SPLIT = 0, /* bf16 hi+lo operands, 3 MMAs per k-step: <= 1e-4 vs fp32 */
       IPOKE_INN_PREC_BF16 = 1 };     /* single bf16 pass: stated looser bound */

/* Architecture of one flow stack.  conditional=1: ConditionalFlow (flow_blocks.py:10-35: blocks
 * alternate 'cond'/'normal' couplings, InvLeakyRelu(alpha) after ActNorm); conditional=0:
 * UnconditionalFlow2 (flow_blocks.py:432-446, no activation, E must be 0).
 * Subnetworks: BasicFullyConnectedNet (modules.py:165-191) with `depth` hidden HxH layers. */
typedef struct ipoke_inn_desc {
  int32_t C;           /* flow_in_channels (even)                                  INN.py:26  */
  int32_t E;           /* flow_embedding_channels (conditioning width), 0 if none  INN.py:31  */
  int32_t H;           /* hidden width (mid_channels)                              INN.py:27  */
  int32_t depth;       /* flow_hidden_depth                                        INN.py:28  */
  int32_t n_flows;     /* number of flow blocks                                    INN.py:29  */
  int32_t conditional; /* 1 = ConditionalFlow, 0 = UnconditionalFlow2                         */
  int32_t precision;   /* IPOKE_INN_PREC_*                                                    */
  float alpha;         /* InvLeakyRelu slope (0.9)                       flow_blocks.py:373   */
  float mlp_slope;     /* nn.LeakyReLU slope inside the subnetworks (0.01)   modules.py:174   */
} ipoke_inn_desc;

typedef struct ipoke_inn_plan ipoke_inn_plan; /* opaque, host side */

int ipoke_inn_version(void);
const char* ipoke_inn_last_error_string(int code);

/* Build the launch plan (tile shapes, packed-weight layout) for one architecture on `device`. */
int ipoke_inn_plan_create(const ipoke_inn_desc* desc, int device, ipoke_inn_plan** out);
void ipoke_inn_plan_destroy(ipoke_inn_plan* plan);

/* Bytes of the packed parameter buffer / of the scratch workspace for batch B.
 * mode 0 = forward/reverse only, 1 = also keeps what ipoke_inn_backward needs. */
size_t ipoke_inn_packed_bytes(const ipoke_inn_plan* plan);
size_t ipoke_inn_workspace_bytes(const ipoke_inn_plan* plan, int B, int mode);
/* Number of nn.Linear layers the plan expects, = n_flows * 2 * 2 * (depth + 2). */
int ipoke_inn_num_linears(const ipoke_inn_plan* plan);

/* Repack module parameters into the kernel layout (bf16 hi/lo, s|t stacked, K padded, transposed copies
 * for dgrad), gather ActNorm loc/scale, pre-compute sum(log|scale|) and the int32 gather maps of every
 * Shuffle.  Replaces nothing in the reference (its weights are used in place by cuBLAS addmm); call once
 * after construction / load_state_dict and after every optimizer step.
 *   linear_w / linear_b : host arrays [num_linears] of device pointers, ordered
 *                         (block, half-step i, net s|t, layer)  -- sub_layers[k].coupling.{s,t}[i].main[2*l]
 *   an_loc / an_scale   : host arrays [n_flows] of device pointers to ActNorm.loc / .scale (C floats each)
 *   perm_fwd / perm_bwd : host arrays [n_flows] of device pointers to Shuffle.forward_shuffle_idx /
 *                         backward_shuffle_idx (int64[C], flow_blocks.py:318-320)                     */
int ipoke_inn_pack(const ipoke_inn_plan* plan, const float* const* linear_w, const float* const* linear_b,
                   const float* const* an_loc, const float* const* an_scale, const int64_t* const* perm_fwd,
                   const int64_t* const* perm_bwd, void* packed, void* stream);

/* Forward (z, logdet) over blocks [block_begin, block_end).  Replaces ConditionalFlow.forward
 * (flow_blocks.py:37-57) / UnconditionalFlow2.forward (:448-459) including every block's
 * ActNorm -> InvLeakyRelu -> double coupling -> Shuffle (flow_blocks.py:255-267, 513-523).
 *   cond         : [B,E] or NULL when E == 0
 *   block_outs   : nullable [block_end-block_begin, B, C]  = flow.last_outs
 *   block_logdets: nullable [block_end-block_begin, B]     = flow.last_logdets (cumulative)
 *   x is never written; z may not alias x.                                                  */
int ipoke_inn_forward(ipoke_inn_plan* plan, const void* packed, const float* x, const float* cond, float* z,
                      float* logdet, float* block_outs, float* block_logdets, int B, int block_begin,
                      int block_end, void* workspace, void* stream);

/* Reverse (sampling) over blocks [block_begin, block_end), last block first.  Replaces
 * ConditionalFlow.forward(reverse=True) (flow_blocks.py:58-63) / UnconditionalFlow2 (:460-462). */
int ipoke_inn_reverse(ipoke_inn_plan* plan, const void* packed, const float* z, const float* cond, float* x,
                      int B, int block_begin, int block_end, void* workspace, void* stream);

/* Sampling loop: n_draws latents per conditioning row in ONE call.  Replaces the reference's per-draw loop
 * `for n in range(n_samples): flow(randn, cond, reverse=True)` (models/second_stage_video_fc.py:356-378, which also
 * re-encodes the same conditioning for every draw, :362) and the repeated-label sampling of
 * SupervisedTransformer.sample (INN.py:59-63).
 *   z, x : [n_draws * B, C], draw j of conditioning row b at row j * B + b;  cond : [B, E]
 * The subnetworks of the 'cond' couplings (even blocks, flow_blocks.py:27, 200-240) read only the conditioning
 * (:224, 237), so their hidden layers run once on the B conditioning rows and every draw's coupling tail reads
 * them; the 'normal' couplings run on all n_draws * B rows.  Results are those of ipoke_inn_reverse on the
 * row-repeated conditioning (same operands and MMAs; only the launch shapes picked for the row count differ).
 * n_draws > 1 needs a conditional plan and B % 128 == 0 (IPOKE_INN_E_UNSUPPORTED otherwise; callers pad B).
 * The workspace is one for batch n_draws * B (ipoke_inn_workspace_bytes(plan, n_draws * B, 0)).           */
int ipoke_inn_reverse_draws(ipoke_inn_plan* plan, const void* packed, const float* z, const float* cond, float* x,
                            int B, int n_draws, void* workspace, void* stream);
/* Same sharing for the forward (z, logdet) pass: n_draws inputs per conditioning row. */
int ipoke_inn_forward_draws(ipoke_inn_plan* plan, const void* packed, const float* x, const float* cond, float* z,
                            float* logdet, int B, int n_draws, void* workspace, void* stream);

/* ---- training: forward that keeps what the backward pass needs, and the fused backward -----------------
 * The reference has no backward code of its own: gradients are torch.autograd over the ops of
 * flow_blocks.py:125-143, 255-267, modules.py:190-191, 292-301.  These two calls replace
 * that autograd graph: `saved` (ipoke_inn_saved_bytes(plan, B) bytes, caller-owned) receives every hidden
 * activation (bf16, as the GEMMs wrote them), the MLP inputs, and per half-step the tanh output s and the
 * pre-update kept half; block_outs ([n_flows, B, C], = flow.last_outs) is required because block k's
 * ActNorm backward reads its input from block_outs[k-1].                                                */
size_t ipoke_inn_saved_bytes(const ipoke_inn_plan* plan, int B);
int ipoke_inn_forward_train(ipoke_inn_plan* plan, const void* packed, const float* x, const float* cond, float* z,
                            float* logdet, float* block_outs, float* block_logdets, int B, void* workspace,
                            void* saved, void* stream);

/* Gradients of  sum(dz * z) + sum(dlogdet * logdet)  over blocks [block_begin, block_end), last block first.
 *   dz          : [B,C] gradient w.r.t. the output of block block_end-1 (the flow output z when block_end == n_flows)
 *   dlogdet     : [B]
 *   dx          : nullable [B,C], gradient w.r.t. the input of block block_begin (feed it as dz of the next call
 *                 when the backward pass is issued in block chunks to overlap the gradient all-reduce)
 *   dcond       : nullable [B,E]; zeroed when block_end == n_flows, accumulated otherwise
 *   grad_w/grad_b      : host arrays [num_linears] of device pointers, same order as ipoke_inn_pack; each
 *                        tensor has the nn.Linear's own shape ([out,in] / [out]) and is OVERWRITTEN
 *   grad_loc/grad_scale: host arrays [n_flows] of device pointers to C floats each (ActNorm, incl. the
 *                        d log|scale| term of the log-det), overwritten
 * Only the entries of blocks in the range are touched.                                                   */
int ipoke_inn_backward(ipoke_inn_plan* plan, const void* packed, const float* x, const float* cond,
                       const float* block_outs, const void* saved, const float* dz, const float* dlogdet, float* dx,
                       float* dcond, float* const* grad_w, float* const* grad_b, float* const* grad_loc,
                       float* const* grad_scale, int B, int block_begin, int block_end, void* workspace,
                       void* stream);

/* Same two calls with HOST buffers (pinned or pageable): host->device copies of x/cond, the kernels, and
 * device->host copies of the results are all enqueued on `stream`; the caller synchronises the stream. */
int ipoke_inn_forward_host(ipoke_inn_plan* plan, const void* packed, const float* x_host, const float* cond_host,
                           float* z_host, float* logdet_host, int B, void* workspace, void* stream);
int ipoke_inn_reverse_host(ipoke_inn_plan* plan, const void* packed, const float* z_host, const float* cond_host,
                           float* x_host, int B, void* workspace, void* stream);

/* ActNorm data-dependent initialisation (modules.py:255-275), split so that ranks can all-reduce the
 * moments: stats accumulates per-channel sum / sum of squares of x [B,C] in fp64 into moments[2*C]
 * (caller zeroes it); finalize writes loc = -mean, scale = 1/(std_unbiased + 1e-6).            */
int ipoke_inn_actnorm_stats(const float* x, int B, int C, double* moments, void* stream);
int ipoke_inn_actnorm_finalize(const double* moments, double count, int C, float* loc, float* scale, void* stream);

/* FlowLoss (loss.py:17-42, 86-97): out[0] = nll_weight*mean(0.5*sum z^2) - logdet_weight*mean(logdet),
 * out[1] = mean nll, out[2] = -mean logdet; also the gradients the backward pass starts from:
 * dz = nll_weight * z / B, dlogdet = -logdet_weight / B (either may be NULL).                 */
int ipoke_inn_flow_loss(const float* z, const float* logdet, int B, int C, float nll_weight, float logdet_weight,
                        float* out3, float* dz, float* dlogdet, void* stream);
/* Same launch plus the reference NLL the reference logs every step: out4[3] = mean(0.5 * sum ref_sample^2) for the
 * N(0, I) sample it draws with randn_like (loss.py:27, 32; the draw stays with the caller: same RNG consumption). */
int ipoke_inn_flow_loss_ref(const float* z, const float* logdet, const float* ref_sample, int B, int C, float nll_weight,
                            float logdet_weight, float* out4, float* dz, float* dlogdet, void* stream);
/* Same again for large batches: up to 32 CTAs park fp64 partial sums in `scratch` (ipoke_inn_flow_loss_scratch_bytes()
 * bytes of device memory, 8-byte aligned, zeroed ONCE by the caller; the kernel leaves it zeroed) and the last one adds
 * them in block order (deterministic).  One call at a time per scratch buffer.  ref_sample nullable; scratch NULL = one CTA. */
int ipoke_inn_flow_loss_scratch_bytes(void);
int ipoke_inn_flow_loss_ws(const float* z, const float* logdet, const float* ref_sample, int B, int C, float nll_weight,
                           float logdet_weight, float* out4, float* dz, float* dlogdet, void* scratch, void* stream);

/* The same loss for callers that must not allocate device memory (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Arguments as ipoke_inn_flow_loss_ws minus `scratch`: the fp64 partial sums live in a file-scope __device__ array
 * inside the library (6400 bytes per GPU that loads the library, held until the process exits), so the call performs
 * no cudaMalloc / cudaFree and calls no allocating routine; the caller still passes every input and output.
 * Limit: B * C <= IPOKE_INN_FLOW_LOSS_STATIC_MAX_ELEMS (256 CTAs x 65536 elements); larger sizes return
 * IPOKE_INN_E_TOO_LARGE and launch nothing.  The array is shared by every call on that GPU and keeps the last call's
 * contents (the ticket word is cleared on `stream` before each launch): calls that overlap on one GPU -- two streams,
 * two host threads -- race on it; serialise them.  The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress, with that device current): make it before starting a graph capture.
 * Results agree with ipoke_inn_flow_loss_ws to fp64-summation-order accuracy (not bitwise: the CTA count differs). */
#define IPOKE_INN_FLOW_LOSS_STATIC_MAX_BLOCKS 256
#define IPOKE_INN_FLOW_LOSS_STATIC_MAX_ELEMS 16777216LL
int ipoke_inn_flow_loss_static(const float* z, const float* logdet, const float* ref_sample, int B, int C,
                               float nll_weight, float logdet_weight, float* out4, float* dz, float* dlogdet,
                               void* stream);

/* Fused Adam / AMSGrad step over all parameter tensors in one launch -- replaces the reference's
 * torch.optim.Adam(lr, betas, weight_decay, amsgrad=True).step() (models/second_stage_video_fc.py:685-687;
 * arithmetic of torch/optim/adam.py `_single_tensor_adam`, L2 weight decay added to the gradient).
 * jobs_dev: device array of n_jobs entries; chunk tables: chunk c updates job chunk_job[c], elements
 * [chunk_off[c], chunk_off[c] + ipoke_inn_adam_chunk_elems()) (device arrays, built once by the caller).
 * step counts from 1; grad_scale multiplies the gradients first (1/world_size after a sum all-reduce).
 * vmax pointers may be NULL when amsgrad == 0.                                                         */
typedef struct ipoke_inn_adam_job {
  float* p;
  const float* g;
  float* m;
  float* v;
  float* vmax;
  long long n;
} ipoke_inn_adam_job;
int ipoke_inn_adam_chunk_elems(void);
int ipoke_inn_adam_step(const ipoke_inn_adam_job* jobs_dev, const int* chunk_job_dev, const long long* chunk_off_dev,
                        int n_chunks, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                        int amsgrad, float grad_scale, void* stream);

/* ---- packed-master training: the sharded optimizer step of the data-parallel training step --------------------
 * The reference's training step is DDP's bucketed all-reduce of every gradient followed by a replicated
 * torch.optim.Adam (experiments/experiment.py:86, models/second_stage_video_fc.py:670-700).  These entry points let the
 * host side run it as reduce-scatter -> Adam on the rank's shard -> all-gather of the low-precision weights, with the
 * packed buffer itself as the all-gather target (no fp32 -> bf16 repack pass):
 *
 * ipoke_inn_param_layout: the packed buffer is [hi arena | lo arena | biases, ActNorm, maps].  out[0] = 1 when every
 *   nn.Linear weight is one unpadded contiguous [out,in] span of the hi arena and the spans tile it (H, E, C/2 multiples
 *   of 64; true for every reference config), out[1] = byte offset of the hi arena, out[2] = its bf16 elements, out[3] =
 *   byte offset of the lo arena (-1 in the bf16 mode; a weight's lo image has the same element offset there), out[4] =
 *   num_linears, out[5] = n_flows, out[8 + li] = element offset of Linear li's weight (order of ipoke_inn_pack),
 *   out[8 + num_linears + 2k], [.. + 2k + 1] = element range of flow block k.  n >= 8 + num_linears + 2 * n_flows.   */
int ipoke_inn_param_layout(const ipoke_inn_plan* plan, long long* out, int n);
/* Refresh only the small part of the packed buffer (fp32 biases, ActNorm loc / scale / sum log|scale|, gather maps)
 * from the tensors given to the last ipoke_inn_pack on this buffer (its job tables live in the buffer). */
int ipoke_inn_pack_small(const ipoke_inn_plan* plan, void* packed, void* stream);
/* Adam / AMSGrad (arithmetic of ipoke_inn_adam_step) on one contiguous span of n weights (n % 4 == 0, 16-byte aligned):
 * p = fp32 master (updated in place), exactly one of g (fp32) / g_bf16 (bf16 wire format) non-NULL, m / v / vmax =
 * moments of the span, hi / lo = where the span's bf16 images live in the packed buffer (lo NULL in the bf16 mode; both NULL
 * for parameters that have no packed image: biases and ActNorm parameters kept in one flat buffer). */
int ipoke_inn_adam_shard_step(float* p, const float* g, const void* g_bf16, float* m, float* v, float* vmax, void* hi,
                              void* lo, long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                              int step, int amsgrad, float grad_scale, void* stream);
/* fp32 master span -> its bf16 hi / lo images (after load_state_dict / out-of-band edits of the master). */
int ipoke_inn_master_to_packed(const float* p, void* hi, void* lo, long long n, void* stream);
/* fp32 gradient span * scale -> bf16 (wire format of the reduce-scatter when the caller asks for bf16 gradients). */
int ipoke_inn_grad_to_bf16(const float* g, void* out_bf16, long long n, float scale, void* stream);

/* Number of kernels the last forward / reverse call on this plan launched (for bench accounting). */
int ipoke_inn_last_launch_count(const ipoke_inn_plan* plan);

/* Launch options: "graph" (default 1: replay a call as one CUDA graph when the caller's pointers repeat),
 * "pdl" (default 1: programmatic dependent launch between the GEMM kernels), "streams" (default 1: how many calls
 * the caller issues concurrently on different CUDA streams with different workspaces; one-wave launches are then
 * sized for num_sms / streams SMs so the concurrent kernels are co-resident).  Environment overrides at plan
 * creation: IPOKE_INN_GRAPH=0/1, IPOKE_INN_PDL=0/1.                                                     */
int ipoke_inn_set_option(ipoke_inn_plan* plan, const char* name, int value);

/* Per-launch timing for the roofline report: while enabled, every kernel of a forward/reverse call is
 * bracketed by CUDA events on the launching stream; read() waits for the last call and returns the summed
 * milliseconds and launch counts of {boundary kernels, hidden-layer GEMMs, tail GEMMs} of that call.     */
int ipoke_inn_profile_enable(ipoke_inn_plan* plan, int on);
int ipoke_inn_profile_read(ipoke_inn_plan* plan, double* ms3, int* launches3);
/* n kinds: 0 elementwise, 1 hidden GEMM, 2 tail GEMM, 3 dgrad GEMM, 4 wgrad GEMM (3, 4: backward calls). */
int ipoke_inn_profile_read_n(ipoke_inn_plan* plan, double* ms, int* launches, int n);

/* Test/diagnostic aid: byte offsets of the scratch buffers inside the workspace for batch B and the padded
 * dimensions, out[20] = {S0, S1, U_hi, U_lo, ACT_hi0, ACT_lo0, ACT_hi1, ACT_lo1, logdet, total, ldU, ldACT,
 * Hp, C2p, Ep, Np, ksplit, Gmax, packed_bytes, num_sms}.                                              */
int ipoke_inn_debug_layout(const ipoke_inn_plan* plan, int B, long long* out, int n);

/* Debug: per-launch device timeline. `dev_buf` (device memory, n_slots * 32 * 8 bytes, zeroed by the caller) gets,
 * for GEMM launch i of every program built afterwards, up to 16 %globaltimer stamps (ns) of its first CTA at
 * [32*i .. 32*i+15] and of its last CTA at [32*i+16 ..]: kernel entry, prologue done, producer past the grid
 * dependency, first operands landed, last MMA issued, accumulator ready (epilogue), epilogue stores issued, exit; 8.. = tail-epilogue phases.
 * dev_buf = NULL switches tracing off.  Not for production use (extra global stores in every role).      */
int ipoke_inn_debug_trace(ipoke_inn_plan* plan, void* dev_buf, int n_slots);

#ifdef __cplusplus
}
#endif
#endif /* IPOKE_INN_H_ */
