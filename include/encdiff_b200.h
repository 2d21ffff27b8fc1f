/*
 * encdiff_b200 — C ABI of the B200-native EncDiff denoising hot path.
 *
 * Drop-in boundary: everything the Python classes encdiff_b200.UNetModel / SpatialTransformer /
 * CrossAttention / DDIMSampler do on the GPU goes through the entry points below.  Plain pointers and sizes only;
 * every pointer is a DEVICE pointer unless it says "host".  All calls are enqueue-only on `stream` (a cudaStream_t
 * passed as void*), never synchronise, never allocate device memory and are therefore CUDA-graph capturable.
 * Return value: 0 on success, non-zero on failure (encdiff_last_error() gives the message; the Python side raises
 * RuntimeError, mirroring the reference's assert/exception convention, openaimodel_enc.py:472-488,722-724).
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   encdiff_op_sgemm            nn.Conv2d 3x3/1x1 and nn.Linear forward + input-gradient,
 *                               ldm/modules/diffusionmodules/openaimodel_enc.py:201-241, ldm/modules/attention.py:37-64,152-193
 *   encdiff_op_wgrad            autograd weight gradient of the same layers (convolution_backward / addmm backward)
 */
#ifndef ENCDIFF_B200_H_
#define ENCDIFF_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

/* Message of the last failing call on this thread ("" if none). */
const char* encdiff_last_error(void);

/* ---------------------------------------------------------------------------------------------------------------
 * Operator level: tensor-core GEMM families (tcgen05 + TMEM + TMA).
 * Activations are NHWC bf16: [B, H, W, C] == token matrix [B*H*W, C].
 * ------------------------------------------------------------------------------------------------------------- */

/* out[M = B*H*W, N] = act( sum_s conv_taps(A_s) * Wp^T + bias ) + res
 *   a_ptrs[s]     bf16 [B,H,W,a_channels[s]]; a_taps[s] = 1 (1x1 conv / linear) or 9 (3x3 conv, zero pad 1)
 *   w_packed      bf16 [w_rows, K], K = sum_s 64*ceil(a_channels[s]/64)*a_taps[s]; K order = source, tap (ky*3+kx),
 *                 channel; rows >= N padded to the N tile `bn` (0 = pick)
 *   tap_sign      +1 forward convolution, -1 mirrored taps (input gradient with transposed weights)
 *   bias          fp32 [N] or NULL; res bf16 [M, ldres] or NULL; act 0 none / 1 SiLU
 *   out0/out1     bf16 (out_fp32 = 0) or fp32 destinations; columns < split_col go to out0 (leading dim ld0), the
 *                 rest to out1 (ld1); out1 = NULL for a single destination                                      */
int encdiff_op_sgemm(int nsrc, const void* const* a_ptrs /*host array*/, const int* a_channels /*host*/,
                     const int* a_taps /*host*/, int H, int W, int B, const void* w_packed, long long w_rows, int N,
                     int tap_sign, const float* bias, const void* res, int ldres, int act, int out_fp32, void* out0,
                     int ld0, int split_col, void* out1, int ld1, int bn, void* stream);

/* out[p*stride_p + (q+q_off)*stride_q + tap*stride_tap] += sum_pixels dy[pix, p] * x(shifted by tap)[pix, q]
 *   dy bf16 [B,H,W,Cp], x bf16 [B,H,W,Cq], out fp32 (atomically accumulated), taps 1 or 9,
 *   splits = CTAs sharing the pixel reduction (>= 1), bn = Cq tile (0 = pick)                                   */
int encdiff_op_wgrad(const void* dy, int Cp, const void* x, int Cq, int H, int W, int B, int taps, float* out,
                     long long stride_p, long long stride_q, long long stride_tap, int q_off, int bn, int splits,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ENCDIFF_B200_H_ */
