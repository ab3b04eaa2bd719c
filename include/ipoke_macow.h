/* ipoke_macow.h -- C ABI of the MaCow conv flow (masked-convolution autoregressive flow), sm_100a.
 *
 * Row N3 of SURVEY.md section 8(f): everything `UnsupervisedMaCowTransformer3` / `SupervisedMacowTransformer`
 * (models/modules/INN/INN.py:365-390, 447-482) compute --
 *   * `MaCowUnit` (macow2.py:925-996) = MaskedConvFlow A, B (macow2.py:25-289) -> ActNorm2dFlow (macow2.py:476-541) ->
 *     MaskedConvFlow C, D -> ActNorm2dFlow, unconditional or conditional (h_channels > 0), forward (out, logdet) and the
 *     reference's sequential line-by-line inverse (macow2.py:222-289) inside one kernel launch, one CTA per sample;
 *   * `NICE2d` couplings with the convolutional net (macow2.py:291-437, macow_utils.py:253-338);
 *   * the ActNorm2dFlow / Shuffle glue of `MaCowStep`, `MultiScalePrior`, `MultiScaleInternal` (macow2.py:543-594,
 *     821-918, 999-1117), in place on one [B, C, H, W] buffer
 * for the affine transform (macow_utils.py:42-66), dropout 0 -- the configuration `config/second_stage.yaml:53-80` selects.
 *
 * Everything is float32 on the device (the reference computes this path in fp32).  All pointers are device
 * pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.  Error codes are those of
 * ipoke_inn.h (IPOKE_INN_OK, IPOKE_INN_E_*; positive values are cudaError_t) and
 * ipoke_inn_last_error_string() decodes them.  The calls are stateless and thread safe.
 *
 * The ctypes binding a maintainer of the reference would add is in INTEGRATION.md section 7; the Python mirrors of the
 * reference modules (same constructors, state_dict keys and RNG draw order) are
 * ipoke-main-opticalflow_b200/macow.py and macow_flow.py.
 */
#ifndef IPOKE_MACOW_H_
#define IPOKE_MACOW_H_

#include <stddef.h>
#include <stdint.h>

#include "ipoke_inn.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { IPOKE_MACOW_ACT_RELU = 0, IPOKE_MACOW_ACT_ELU = 1, IPOKE_MACOW_ACT_LEAKY_RELU = 2 }; /* macow_utils.py:417-424 */

enum { IPOKE_MACOW_STAGES_PER_UNIT = 6 }; /* conv1, conv2, actnorm1, conv3, conv4, actnorm2 (macow2.py:965-982) */

typedef struct ipoke_macow_desc {
  int32_t C;          /* in_channels                                                        macow2.py:930  */
  int32_t hidden;     /* hidden channels of the MCF blocks: 4C for C <= 96, else min(2C,512) macow2.py:36-41 */
  int32_t k0;         /* kernel_size[0]: taps along the masked axis                         macow2.py:938  */
  int32_t k1;         /* kernel_size[1]: centred taps along the line, odd (1, 3 or 5)   macow_utils.py:454 */
  int32_t activation; /* IPOKE_MACOW_ACT_*                                                                  */
  float alpha;        /* Affine alpha: scale = tanh(log_scale / 2) * alpha + 1           macow_utils.py:51 */
  int32_t h_channels; /* conditioning channels concatenated to the hidden activations of every masked flow
                         (MCFBlock, macow_utils.py:409-412, 431-432); 0 = unconditional                        */
} ipoke_macow_desc;

/* Floats of one packed unit (shift weights transposed to [C][k0][k1][hidden], weight-normed 1x1 weights
 * transposed to [hidden][C][2] = (shift, log-scale) pairs, biases likewise, ActNorm log-scales and biases). */
size_t ipoke_macow_unit_floats(const ipoke_macow_desc* desc);

/* Pack one unit.  `params` is a HOST array of 20 device pointers in the reference's state_dict order of the
 * learnable tensors (MaCowUnit.state_dict(), buffers skipped), float32 contiguous:
 *   [4f+0] conv{f+1}.net.shift_conv.weight      [hidden, C, kH, kW]  (kH, kW) = (k0, k1) for conv1/2, (k1, k0) for conv3/4
 *   [4f+1] conv{f+1}.net.conv1x1.conv.bias      [2C]
 *   [4f+2] conv{f+1}.net.conv1x1.conv.weight_g  [2C]       (nn.utils.weight_norm gain, macow_utils.py:233)
 *   [4f+3] conv{f+1}.net.conv1x1.conv.weight_v  [2C, hidden]
 *   [16] actnorm1.log_scale [C]   [17] actnorm1.bias [C]   [18] actnorm2.log_scale [C]   [19] actnorm2.bias [C]
 * `packed` receives ipoke_macow_unit_floats() floats. */
int ipoke_macow_pack_unit(const ipoke_macow_desc* desc, const void* const* params, float* packed, void* stream);

/* Forward of stages [stage_begin, stage_end) of a chain of n_units units packed back to back (stage s belongs
 * to unit s / 6): x [B, C, H, W] -> out [B, C, H, W] (may alias x), logdet [B] (nullable; accumulate != 0 adds
 * to the existing values).  The full chain is stage_begin = 0, stage_end = 6 * n_units. */
int ipoke_macow_forward(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* x, float* out,
                        float* logdet, int accumulate, int B, int H, int W, int stage_begin, int stage_end,
                        void* stream);

/* Inverse of the same stage range, applied last stage first (macow2.py:983-996): z -> x. */
int ipoke_macow_reverse(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* z, float* out,
                        int B, int H, int W, int stage_begin, int stage_end, void* stream);

/* Dynamic shared memory (bytes) and threads per CTA the two calls above would launch with (for batches below 2048;
 * the inverse uses half the threads from there on), for reporting. */
int ipoke_macow_launch_info(const ipoke_macow_desc* desc, int H, int W, int reverse, int* smem_bytes, int* threads,
                            int* tile_lines);

/* The two calls above with explicit sample strides (floats between consecutive samples of x / out, >= C * H * W):
 * the multi-scale flow (MultiScaleInternal, macow2.py:872-918) runs every level in place on the first channels of
 * one [B, C_total, H, W] buffer, so its split2d / unsplit2d (macow2.py:15-23) cost nothing. */
int ipoke_macow_units_forward(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* x,
                              long long x_stride, float* out, long long out_stride, float* logdet, int accumulate,
                              int B, int H, int W, int stage_begin, int stage_end, void* stream);
int ipoke_macow_units_reverse(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* z,
                              long long z_stride, float* out, long long out_stride, int B, int H, int W,
                              int stage_begin, int stage_end, void* stream);

/* ---- Conditional units (h_channels > 0: SupervisedMacowTransformer, INN.py:447-482, as models/second_stage_video.py:122
 * builds it).  The conditioning h [B, h_channels, H, W] is concatenated to the hidden activations of every masked flow
 * before the activation and the weight-normed 1x1 convolution (macow_utils.py:429-436), so its contribution to
 * (shift, log-scale) does not depend on the image: for a chain of n_units units it is ONE product
 *   hb [B H W, n_units * 4 * 2C] = act(h) [B H W, h_channels] x Wh [h_channels, n_units * 4 * 2C]
 * computed once per call and added by the unit kernel in both directions. */
/* Floats of Wh for a chain of n_units units: h_channels rows of round4(n_units * 8 C) columns. */
size_t ipoke_macow_cond_floats(const ipoke_macow_desc* desc, int n_units);
/* ipoke_macow_pack_unit for a conditional unit (weight_v is [2C, hidden + h_channels]); also writes the unit's columns
 * of Wh (`wh`: the chain's matrix, `unit_index` < n_units). */
int ipoke_macow_pack_unit_cond(const ipoke_macow_desc* desc, const void* const* params, float* packed, float* wh,
                               int n_units, int unit_index, void* stream);
/* hb [B * H * W][round4(n_units * 8 C)] from h (sample stride h_stride >= h_channels * H * W). */
int ipoke_macow_cond_bias(const ipoke_macow_desc* desc, const float* wh, int n_units, const float* h, long long h_stride,
                          float* hb, int B, int H, int W, void* stream);
/* ipoke_macow_units_forward / _reverse with the conditioning product `hb` of the chain. */
int ipoke_macow_units_forward_cond(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* hb,
                                   const float* x, long long x_stride, float* out, long long out_stride, float* logdet,
                                   int accumulate, int B, int H, int W, int stage_begin, int stage_end, void* stream);
int ipoke_macow_units_reverse_cond(const ipoke_macow_desc* desc, const float* packed, int n_units, const float* hb,
                                   const float* z, long long z_stride, float* out, long long out_stride, int B, int H,
                                   int W, int stage_begin, int stage_end, void* stream);

/* ---- NICE2d coupling with the convolutional net (macow2.py:291-437; NICEConvBlock macow_utils.py:253-338 without
 * normalisation layers, attention or conditioning; Affine macow_utils.py:42-66).  The net reads image channels
 * z_off + z_stride * i (i < c_in) and transforms channels p_off + p_stride * i (i < c_t) in place:
 *   split 'continuous', order 'up'   (macow2.py:366-367, 410):  z_off 0,         z_stride 1, p_off C - C / factor, p_stride 1
 *   split 'continuous', order 'down':                           z_off C / factor, z_stride 1, p_off 0,              p_stride 1
 *   split 'skip' (even C), 'up'      (macow2.py:368-373):       z_off 0,         z_stride 2, p_off 1,              p_stride 2
 *   split 'skip' (even C), 'down':                              z_off 1,         z_stride 2, p_off 0,              p_stride 2
 * with c_t = C / factor, c_in = C - c_t. */
typedef struct ipoke_macow_nice_desc {
  int32_t c_in, z_off, z_stride;
  int32_t c_t, p_off, p_stride;
  int32_t hidden;     /* hidden_channels of the coupling net (flow_mid_channels)                macow2.py:312-313 */
  int32_t activation; /* IPOKE_MACOW_ACT_*                                                                         */
  float alpha;
} ipoke_macow_nice_desc;

/* Floats of one packed coupling net (the three convolutions transposed to [K][N] with N padded to 4, the last one
 * weight-normed, its nine taps side by side with (shift, log-scale) column pairs, and its bias). */
size_t ipoke_macow_nice_floats(const ipoke_macow_nice_desc* desc);

/* `params`: HOST array of 5 device pointers, float32 contiguous, in state_dict order:
 *   net.conv1.weight [hidden, c_in, 3, 3], net.conv2.weight [hidden, hidden, 1, 1], net.conv3.conv.bias [2 c_t],
 *   net.conv3.conv.weight_g [2 c_t], net.conv3.conv.weight_v [2 c_t, hidden, 3, 3]. */
int ipoke_macow_pack_nice(const ipoke_macow_nice_desc* desc, const void* const* params, float* packed, void* stream);

/* Preferred workspace (floats) for batch B: two pixel-major hidden-activation buffers and the nine tap outputs of the
 * last convolution for as many samples as keep a pass in L2.  Any size >= H * W * (2 * hidden + 9 * round4(2 * c_t))
 * works (fewer samples per pass). */
size_t ipoke_macow_nice_workspace_floats(const ipoke_macow_nice_desc* desc, int B, int H, int W);

/* Forward (reverse == 0: transformed channels <- scale * x + shift, logdet[b] += sum log scale; logdet nullable) or
 * inverse (reverse != 0: (x - shift) / (scale + 1e-12)) of one coupling on x [B, *, H, W] with sample stride x_stride. */
int ipoke_macow_nice(const ipoke_macow_nice_desc* desc, const float* packed, float* x, long long x_stride, float* logdet,
                     int reverse, int B, int H, int W, float* ws, size_t ws_floats, void* stream);

/* Per-channel glue of MaCowStep / MultiScalePrior / MultiScaleInternal, in place on the first channels of every sample:
 * an optional ActNorm2dFlow (macow2.py:502-521; forward x * exp(log_scale) + bias with logdet[b] += H W sum(log_scale),
 * inverse (x - bias) / (exp(log_scale) + 1e-8)) on channels [norm_c0, norm_c0 + norm_n), and an optional Shuffle
 * (flow_blocks.py:322-326) out[c] = in[perm[c]] of the first n_perm channels (perm: the module's int64 buffer on the
 * device); norm_first != 0 applies the ActNorm before the permutation, else after it. */
int ipoke_macow_chanmap(float* x, long long x_stride, int B, int H, int W, const long long* perm, int n_perm,
                        const float* log_scale, const float* bias, int norm_c0, int norm_n, int norm_inverse,
                        int norm_first, float* logdet, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IPOKE_MACOW_H_ */
