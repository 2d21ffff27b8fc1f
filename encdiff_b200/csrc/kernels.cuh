// Launchers of the non-GEMM kernels (HBM-bound: norms, attention, elementwise glue).  All tensors NHWC h16
// unless stated.  Every launcher returns a cudaError_t as int and only enqueues work on `stream`.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "types.cuh"

namespace edb {


enum { RESAMPLE_NONE = 0, RESAMPLE_DOWN = 1, RESAMPLE_UP = 2 };

// GroupNorm(32) over the channel-concatenation of up to two sources, optionally followed by FiLM
// (y*(1+scale)+shift), SiLU and 2x average-pool / nearest-upsample.
// Reference: GroupNorm32 util.py:227-244; ResBlock._forward openaimodel_enc.py:255-271; Normalize attention.py:76-77.
struct GnParams {
  const h16* x0; int C0;      // [B, H*W, C0]
  const h16* x1; int C1;      // [B, H*W, C1] or null/0
  int H, W;                    // input spatial size
  const float* gamma; const float* beta;   // [C0+C1]
  const float* film; int ldfilm; int film_off;  // scale at film[b*ldfilm+film_off+c], shift at +C; null = no FiLM
  float eps; int silu; int resample;
  float* stats;                // [B, 32, 2] (mean, rstd): written by fwd, read by bwd
  // forward outputs
  h16* act;                   // [B, H'*W', C]
  h16* xs;                    // resampled raw input [B, H'*W', C] (only when resample != NONE), may be null
  // backward inputs/outputs
  const h16* dact;            // grad of act
  const h16* dxs;             // grad arriving on the (resampled) raw-input path or null; same resolution as act
  h16* dx0; int acc0;         // grad of x0 (acc: add to existing content)
  h16* dx1; int acc1;
  float* dgamma; float* dbeta; // fp32 atomics
  h16* dfilm; int lddfilm;    // grad of film (same column layout), plain store
  const float* gscale;         // device {scale, 1/scale} of the activation gradients (null = 1): fp32 outputs are unscaled
  int stash_off;               // backward, set by the launcher: float offset of the per-thread dy' stash in shared memory (0 = none)
};
int launch_gn_fwd(const GnParams& p, int B, cudaStream_t s);
int launch_gn_bwd(const GnParams& p, int B, cudaStream_t s);

// LayerNorm over the last dim (attention.py:203-205, eps 1e-5).  stats [M,2] = (mean, rstd).
int launch_ln_fwd(const h16* x, const float* gamma, const float* beta, float eps, h16* y, float* stats, int M, int C,
                  cudaStream_t s);
// dx = dres + LN-backward(dy)   (dres = gradient of the residual branch that shares x, may be null)
int launch_ln_bwd(const h16* x, const float* stats, const float* gamma, const h16* dy, const h16* dres, h16* dx,
                  float* dgamma, float* dbeta, const float* gscale, int M, int C, cudaStream_t s);

// Self-attention over N tokens per sample (CrossAttention.forward with context=None, attention.py:170-193).
// qkv [B*N, 3C] (q | k | v, head h at columns h*D), out [B*N, C], lse [B*heads*N] (log2 domain).
int launch_self_attn_fwd(const h16* qkv, h16* out, float* lse, int B, int N, int heads, int D, cudaStream_t s);
// tensor-core (mma.sync) variants, attn_mma.cu; return -1 when the shape is not handled (N % 16, N > 256, D not 8/16/32)
int launch_self_attn_fwd_mma(const h16* qkv, h16* out, float* lse, int B, int N, int heads, int D, cudaStream_t s);
int launch_self_attn_bwd_mma(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                             int heads, int D, cudaStream_t s);
int launch_self_attn_bwd(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                         int heads, int D, cudaStream_t s);

// Cross-attention against T concept tokens of width Dc (attention.py:170-193 with context; openaimodel_enc.py:729).
// q [B*N, C] h16, ctx [B, T, Dc] fp32, wk/wv [C, Dc] fp32 (reference to_k/to_v weights), out [B*N, C].
// kv (optional): K | V [B, T, 2C] fp32 already projected by launch_cross_kv_project (then ctx / wk / wv are not read)
int launch_cross_attn_fwd(const h16* q, const float* ctx, const float* wk, const float* wv, h16* out, int B, int N,
                          int heads, int D, int T, int Dc, cudaStream_t s, const float* kv = nullptr);
struct CrossKvEntry {
  const float* wk;       // [C, Dc] reference to_k / to_v weights
  const float* wv;
  float* kv;             // [B, T, 2C]
  int C;
  int pad;
};
int launch_cross_kv_project(const CrossKvEntry* tab_dev, int n_layers, const float* ctx, int B, int T, int Dc, cudaStream_t s);
int launch_cross_attn_bwd(const h16* q, const float* ctx, const float* wk, const float* wv, const h16* dout, h16* dq,
                          float* dwk, float* dwv, float* dctx, const float* gscale, float* dkv_scratch, int B, int N,
                          int heads, int D, int T, int Dc, cudaStream_t s, bool defer_proj = false,
                          const float* kv = nullptr);
// tensor-core (mma.sync) variants over pre-projected K | V, attn_mma.cu; -1 when the shape is not handled
int launch_cross_attn_fwd_mma(const h16* q, const float* kv, h16* out, int B, int N, int heads, int D, int T, cudaStream_t s);
// register-resident variants (attn_reg.cu); -1 = shape not covered
int launch_cross_attn_fwd_reg(const h16* q, const float* kv, h16* out, int B, int N, int heads, int D, int T, cudaStream_t s);
int launch_cross_attn_bwd_reg(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads,
                              int D, int T, cudaStream_t s);
int launch_cross_attn_bwd_mma(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads,
                              int D, int T, cudaStream_t s);
// deferred form of the projection half (d to_k, d to_v, d context) for all layers at once: dkv of every layer must still
// be alive; dctx is accumulated with atomics (caller zeroes it)
struct CrossProjEntry {
  const float* dkv;      // [B, T, 2C]
  const float* wk;       // [C, Dc]
  const float* wv;
  float* dwk;
  float* dwv;
  int C;
  int pad;
};
int launch_cross_attn_bwd_proj_group(const CrossProjEntry* tab_dev, int n_layers, int max_C, const float* ctx, float* dctx,
                                     const float* gscale, int B, int T, int Dc, cudaStream_t s);
// bytes of the fp32 scratch [B, T, 2*heads*D] the backward needs (per-sample dK | dV before the to_k / to_v transpose)
size_t cross_attn_bwd_scratch_bytes(int B, int heads, int D, int T);

// ---- elementwise glue -------------------------------------------------------------------------------------------
// sinusoidal timestep embedding, cos half first (util.py:179-199): t [B] fp32 -> h16 [B, dim]
int launch_temb(const float* t, h16* out, int B, int dim, cudaStream_t s);
// y = silu(x): x fp32 [n] -> h16;  dx = dy * silu'(x): dy h16, x fp32 -> dx h16
int launch_silu_fwd(const float* x, h16* y, long long n, cudaStream_t s);
int launch_silu_bwd(const float* x, const h16* dy, h16* dx, long long n, cudaStream_t s);
// GEGLU (attention.py:37-45): g [M, 2*inner] (value | gate) -> out [M, inner] = value * gelu_erf(gate)
int launch_geglu_fwd(const h16* g, h16* out, long long M, int inner, cudaStream_t s);
int launch_geglu_bwd(const h16* g, const h16* dout, h16* dg, long long M, int inner, cudaStream_t s);
// bias gradient: dst0[n] (+= and dst1[n] +=) sum_m x[m, n]; x h16 [M, ld] columns [col0, col0+N)
int launch_colsum(const h16* x, long long M, int ld, int col0, int N, float* dst0, float* dst1, const float* gscale,
                  cudaStream_t s);
// input stem: x NCHW fp32 [B,Cin,H,W] (optionally q_sample'd: sa[b]*x + sb[b]*noise) -> im2col h16 [B*H*W, 64],
// column ci*9 + (ky*3+kx), zero padded; also writes x_t (NCHW fp32) when xt != null
int launch_stem_im2col(const float* x, const float* noise, const float* sa, const float* sb, float* xt, h16* cols,
                       int B, int Cin, int H, int W, cudaStream_t s);
// transpose of the above for the input gradient: dcols h16 [B*H*W, 64] -> dx NCHW fp32 (overwrite)
int launch_stem_col2im(const h16* dcols, float* dx, const float* gscale, int B, int Cin, int H, int W, cudaStream_t s);
// head: y fp32 [B*H*W, C] (NHWC) -> out NCHW fp32
int launch_nhwc_to_nchw(const float* y, float* out, int B, int C, int H, int W, cudaStream_t s);
// dout NCHW fp32 [B,C,H,W] -> h16 [B*H*W, Cpad] (zero padded)
int launch_nchw_to_nhwc_h16(const float* x, h16* out, const float* gscale, int B, int C, int H, int W, int Cpad,
                            cudaStream_t s);
// Range scale of the backward pass: gscale[0] = 2^k with 2^k * max|x| in [2^-4, 2^-3) (1 when x == 0 or in the bf16
// build), gscale[1] = 2^-k.  The h16 activation gradients are carried multiplied by gscale[0]; every fp32 gradient
// output (weights, biases, norms, d context, dx) is multiplied by gscale[1].  One block; n = B*C*H*W.
int launch_grad_scale(const float* x, long long n, float* gscale, cudaStream_t s);
// L1 loss of p_losses (ddpm_enc.py:297-301,1198-1213): loss += mean|eps_hat-noise| ; deps = sign(eps_hat-noise)/numel
int launch_l1_loss(const float* eps_hat, const float* noise, float* loss, float* deps, long long n, cudaStream_t s);
// DDIM update (ddim.py:197-206); coef = {a_t, a_prev, sigma_t, sqrt_one_minus_at} (device fp32[4]); noise may be null
int launch_ddim_update(const float* x, const float* e_t, const float* noise, const float* coef, float temperature,
                       float* x_prev, float* pred_x0, long long n, cudaStream_t s);

// device-resident DDIM loop state (one captured step graph replayed S times): tables are in loop order (i = 0 is the
// first, noisiest step); cursor/idx are device ints
int launch_ddim_select(const float* t_table, int* cursor, float* ts, int* idx, int B, cudaStream_t s);
int launch_ddim_update_indexed(const float* x, const float* e_t, const float* noise_table, const float* coef_table,
                               const int* idx, float temperature, float* x_prev, float* pred_x0, long long n,
                               cudaStream_t s);

// ---- concept-token MLPs (Encoder4.warp, openaimodel_enc.py:1017-1041): latent_unit independent MLPs 1 -> 64 -> 128 -> Dc ----
// all fp32; parameters stacked over the units; x [B, L] scalar codes, y [B, L * Dc] tokens (unit-major, the reference's
// th.cat(out, dim=1)); backward ACCUMULATES into the d* parameter gradients (fp32 atomics) and overwrites dx (may be null)
struct ConceptParams {
  const float* x;
  const float *w1, *b1;      // [L, 64], [L, 64]          (Linear(1, 64).weight squeezed)
  const float *w2, *b2;      // [L, 128, 64], [L, 128]
  const float *w3, *b3;      // [L, Dc, 128], [L, Dc]
  float* y;
  const float* dy;
  float* dx;
  float *dw1, *db1, *dw2, *db2, *dw3, *db3;
  int B, L, Dc;
};
int launch_concept_mlp_fwd(const ConceptParams& p, cudaStream_t s);
int launch_concept_mlp_bwd(const ConceptParams& p, cudaStream_t s);

// ---- weight packing ---------------------------------------------------------------------------------------------
struct PackEntry {
  const float* src;      // fp32 parameter, reference layout [rows, cin, taps]
  const float* src2;     // second addend for kind 3
  long long dst_off;     // element offset in the destination arena (h16 for kinds 0/1, fp32 for kind 2/3)
  long long prefix;      // first global element index of this entry
  int kind;              // 0: dst[row0+r][col0 + tap*cin_pad + ci]   (forward, K = tap-major)
                         // 1: dst[row0+ci][col0 + tap*rows_pad + r]  (transposed, for the input gradient)
                         // 2: fp32 copy  dst[col0 + i] = src[i]
                         // 3: fp32 sum   dst[col0 + i] = src[i] + src2[i]
                         // 4: dst[row0+r][col0 + ci*taps + tap]      (stem: flat reference order)
                         // 5: dst[row0 + ci*taps + tap][col0 + r]    (stem transposed)
                         // 6: dst[row0 + tap*cin_pad + ci][col0 + r]  (transposed, tap-major rows: dgrad of an im2col conv)
  int rows, cin, taps;
  int cin_pad, rows_pad;
  int dst_ld, row0, col0;
  int blk0;              // first block of this entry in the pack launch (pack_entry_blocks() blocks each)
  int pad_;
};
int pack_entry_blocks(const PackEntry& e);   // taps <= 9 for kinds 0 / 1
int launch_pack(const PackEntry* table_dev, int n_entries, int n_blocks, h16* arena_h16, float* arena_f32,
                cudaStream_t s);

// fused AdamW (+ optional EMA) over a flat fp32 parameter buffer (ddpm_enc.py:1598-1639, ema.py:25-44)
int launch_adamw(float* p, const float* g, float* m, float* v, float* ema, long long n, float lr, float beta1,
                 float beta2, float eps, float wd, float bc1, float bc2, float ema_decay, cudaStream_t s);

// hyper (device) = {lr, 1-beta1^t, 1-beta2^t, ema_decay}; g is multiplied by grad_scale (1/world_size after a sum
// all-reduce) before use
int launch_adamw_dev(float* p, const float* g, float* m, float* v, float* ema, long long n, const float* hyper,
                     float beta1, float beta2, float eps, float wd, float grad_scale, cudaStream_t s);
// step[0] += 1; hyper = {lr[0], 1-beta1^step, 1-beta2^step, min(ema_decay_max, (1+step)/(10+step))}   (all device)
int launch_adamw_hyper_advance(int* step, const float* lr, float* hyper, float beta1, float beta2, float ema_decay_max,
                               cudaStream_t s);

// ---- first-stage VQ encoder (vq_kernels.cu) ----------------------------------------------------------------------
// Downsample-with-conv im2col: x NHWC [B, H, H, C] -> cols [B*(H/2)^2, 9 * ceil64(C)] tap-major (pad right / bottom)
int launch_vq_down_im2col(const h16* x, h16* cols, int B, int H, int C, cudaStream_t s);
// single-head attention of AttnBlock: qkv [B*N, 3D] -> out [B*N, D], D in {64, 128}, N <= 256 multiple of 16
int launch_vq_attn(const h16* qkv, h16* out, int B, int N, int D, cudaStream_t s);
// decoder: nearest x2 upsample (NHWC h16), and the front of VQModelInterface.decode (quantise / concat / post_quant_conv)
int launch_vq_upsample2x(const h16* x, h16* y, int B, int H, int C, cudaStream_t s);
int launch_vq_post_quant(const float* h, const float* sdis, const float* codebook, int n_embed, const float* w,
                         const float* bias, float* out, int B, int HW, int ed, int dd, int zc, cudaStream_t s);
// quant_conv (1x1, zc -> ed) + NHWC fp32 -> NCHW fp32
int launch_vq_tail(const float* h, const float* wq, const float* bq, float* z, int B, int HW, int zc, int ed, cudaStream_t s);

// ---- concept-encoder trunk (enc4_kernels.cu) ---------------------------------------------------------------------
// tap-major im2col of a k x k / stride / pad convolution over NHWC h16, its image (fp32 NCHW, flat (ci, tap) order, K <= 64)
// variant, and the col2im gather of the input gradient (optional ReLU gate on `gate` > 0 and added skip gradient)
int launch_e4_im2col(const h16* x, h16* cols, int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s);
int launch_e4_im2col_image(const float* x, h16* cols, int B, int Cin, int H, int W, int k, int stride, int pad, cudaStream_t s);
int launch_e4_col2im(const h16* dcols, h16* dx, const h16* gate, const h16* add, const float* add_from, const float* add_to,
                     int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s);
// BatchNorm2d: column sums (forward: x, x^2; backward: masked dy, dy * xhat), forward apply (+ running statistics, + ReLU),
// backward apply (+ dgamma / dbeta)
int launch_e4_bn_sums(const h16* x, const h16* dy, const h16* out, const float* stat, float* sums, long long M, int C, int bwd,
                      cudaStream_t s);
int launch_e4_bn_apply(const h16* x, const float* sums, const float* gamma, const float* beta, float* run_mean, float* run_var,
                       float* stat, h16* y, long long M, int C, float eps, float momentum, int training, int relu, cudaStream_t s);
int launch_e4_bn_bwd_apply(const h16* x, const h16* dy, const h16* out, const float* stat, const float* sums, const float* gamma,
                           float* dgamma, float* dbeta, const float* gs_in, float* gs_out, h16* dx, long long M, int C,
                           int training, cudaStream_t s);
int launch_e4_flatten(const h16* in, h16* out, int B, int HW, int C, int to_flat, cudaStream_t s);
struct E4Finalize {            // the tap-major staging matrices of one backward and the reference-layout gradients they fold into
  float* stage[8];
  float* grad[8];
  int KK[8];
  int Cout, C, n;
};
int launch_e4_wgrad_finalize(const E4Finalize& f, cudaStream_t s);

}  // namespace edb
