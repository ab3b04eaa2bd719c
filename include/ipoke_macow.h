/* ipoke_macow.h -- C ABI of the MaCow unit (masked-convolution autoregressive flow), sm_100a.
 *
 * Row N3 of SURVEY.md section 8(f), first slice: the reference's `MaCowUnit`
 * (models/modules/INN/macow2.py:925-996) = MaskedConvFlow A, B (macow2.py:25-289) -> ActNorm2dFlow
 * (macow2.py:476-541) -> MaskedConvFlow C, D -> ActNorm2dFlow, unconditional (h_channels = 0), affine
 * transform (macow_utils.py:42-66), dropout 0 -- the configuration `config/second_stage.yaml:53-80` selects.
 * Forward returns (out, logdet); the inverse is the reference's sequential line-by-line loop
 * (macow2.py:222-289) run inside one kernel launch, one CTA per sample.
 *
 * Everything is float32 on the device (the reference computes this path in fp32).  All pointers are device
 * pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.  Error codes are those of
 * ipoke_inn.h (IPOKE_INN_OK, IPOKE_INN_E_*; positive values are cudaError_t) and
 * ipoke_inn_last_error_string() decodes them.  The calls are stateless and thread safe.
 *
 * The ctypes binding a maintainer of the reference would add is in INTEGRATION.md; the Python mirror of the
 * reference modules (same constructor, state_dict keys and RNG draw order) is
 * ipoke-main-opticalflow_b200/macow.py.
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

#ifdef __cplusplus
}
#endif
#endif /* IPOKE_MACOW_H_ */
