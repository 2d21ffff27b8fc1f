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
/* The 16-bit storage / tensor-core operand type the library was built with: 0 = fp16 (default), 1 = bf16
 * (-DEDB_ACT_BF16).  Wherever this header says "h16" it means that type. */
int encdiff_act_format(void);

/* ---------------------------------------------------------------------------------------------------------------
 * Operator level: tensor-core GEMM families (tcgen05 + TMEM + TMA).
 * Activations are NHWC bf16: [B, H, W, C] == token matrix [B*H*W, C].
 * ------------------------------------------------------------------------------------------------------------- */

/* out[M = B*H*W, N] = act( sum_s conv_taps(A_s) * Wp^T + bias ) + res
 *   a_ptrs[s]     h16 [B,H,W,a_channels[s]]; a_taps[s] = 1 (1x1 conv / linear) or 9 (3x3 conv, zero pad 1)
 *   w_packed      h16 [w_rows, K], K = sum_s 64*ceil(a_channels[s]/64)*a_taps[s]; K order = source, tap (ky*3+kx),
 *                 channel; rows >= N padded to the N tile `bn` (0 = pick)
 *   tap_sign      +1 forward convolution, -1 mirrored taps (input gradient with transposed weights)
 *   bias          fp32 [N] or NULL; res h16 [M, ldres] or NULL; act 0 none / 1 SiLU
 *   out0/out1     h16 (out_fp32 = 0) or fp32 destinations; columns < split_col go to out0 (leading dim ld0), the
 *                 rest to out1 (ld1); out1 = NULL for a single destination                                      */
int encdiff_op_sgemm(int nsrc, const void* const* a_ptrs /*host array*/, const int* a_channels /*host*/,
                     const int* a_taps /*host*/, int H, int W, int B, const void* w_packed, long long w_rows, int N,
                     int tap_sign, const float* bias, const void* res, int ldres, int act, int out_fp32, void* out0,
                     int ld0, int split_col, void* out1, int ld1, int bn, void* stream);

/* out[p*stride_p + (q+q_off)*stride_q + tap*stride_tap] += sum_pixels dy[pix, p] * x(shifted by tap)[pix, q]
 *   dy h16 [B,H,W,Cp], x h16 [B,H,W,Cq], out fp32 (atomically accumulated), taps 1 or 9,
 *   splits = CTAs sharing the pixel reduction (>= 1), bn = Cq tile (0 = pick)                                   */
int encdiff_op_wgrad(const void* dy, int Cp, const void* x, int Cq, int H, int W, int B, int taps, float* out,
                     long long stride_p, long long stride_q, long long stride_tap, int q_off, int bn, int splits,
                     void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Operator level: the HBM-bound kernels of SpatialTransformer / CrossAttention (ldm/modules/attention.py), exposed so
 * each can be checked against its reference call site on its own.  h16 tensors, fp32 parameters.
 * ------------------------------------------------------------------------------------------------------------- */

/* GroupNorm(32 groups) (+ optional SiLU): Normalize attention.py:76-77 (eps 1e-6), GroupNorm32 util.py:227-244 (1e-5).
 *   x, out h16 [B, H*W, C] (NHWC); gamma, beta fp32 [C]; stats fp32 [B, 32, 2] receives (mean, rstd).          */
int encdiff_op_group_norm(const void* x, const float* gamma, const float* beta, float eps, int silu, void* out,
                          float* stats, int B, int H, int W, int C, void* stream);
/* nn.LayerNorm over the last dim (attention.py:203-205): x, y h16 [M, C]; stats fp32 [M, 2] = (mean, rstd).    */
int encdiff_op_layer_norm(const void* x, const float* gamma, const float* beta, float eps, void* y, float* stats,
                          long long M, int C, void* stream);
/* CrossAttention.forward with context=None between the projections (attention.py:176-190):
 *   qkv h16 [B*N, 3*heads*D] (q | k | v, head h at columns h*D), out h16 [B*N, heads*D],
 *   lse fp32 [B*heads*N] (log2 domain, consumed by the backward).  Scale D^-1/2.                                */
int encdiff_op_self_attention(const void* qkv, void* out, float* lse, int B, int N, int heads, int D, void* stream);
int encdiff_op_self_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                  int B, int N, int heads, int D, void* stream);
/* CrossAttention.forward with a context of T concept tokens of width Dc, INCLUDING to_k / to_v (attention.py:174-190):
 *   q h16 [B*N, heads*D], ctx fp32 [B, T, Dc], wk / wv fp32 [heads*D, Dc] (reference weight layout), out h16.
 * Backward: dq h16, dwk / dwv / dctx fp32 (accumulated: caller zeroes), scratch fp32 of ..._scratch_bytes().     */
/* kv (optional, may be NULL): K | V [B, T, 2*heads*D] fp32 produced by encdiff_op_cross_kv_project (what the UNet does
 * once per forward for all its layers); with it the tensor-core path runs and ctx / wk / wv are not read by the forward. */
int encdiff_op_cross_kv_project(const float* ctx, const float* wk, const float* wv, float* kv, void* table_scratch /*>=64 B*/,
                                int B, int C, int T, int Dc, void* stream);
int encdiff_op_cross_attention(const void* q, const float* ctx, const float* wk, const float* wv, const float* kv, void* out,
                               int B, int N, int heads, int D, int T, int Dc, void* stream);
long long encdiff_op_cross_attention_bwd_scratch_bytes(int B, int heads, int D, int T);
int encdiff_op_cross_attention_bwd(const void* q, const float* ctx, const float* wk, const float* wv, const float* kv,
                                   const void* dout, void* dq, float* dwk, float* dwv, float* dctx, float* scratch, int B,
                                   int N, int heads, int D, int T, int Dc, void* stream);
/* Allocation-free variant of encdiff_op_cross_attention_bwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and arithmetic, but the fp32 dK | dV scratch lives in a static device array of
 * ENCDIFF_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it until exit); the call performs no
 * cudaMalloc / cudaFree / allocating library routine.  B*T*2*heads*D*4 bytes must fit, otherwise the call returns a
 * non-zero status and encdiff_last_error() names both sizes.  All calls on a device share the one array: calls that
 * overlap on the GPU (different streams) race -- serialise them, or use the caller-scratch entry point. */
#define ENCDIFF_NOALLOC_SCRATCH_BYTES (16u << 20)
int encdiff_op_cross_attention_bwd_noalloc(const void* q, const float* ctx, const float* wk, const float* wv,
                                           const float* kv, const void* dout, void* dq, float* dwk, float* dwv,
                                           float* dctx, int B, int N, int heads, int D, int T, int Dc, void* stream);
/* GEGLU (attention.py:37-45): g h16 [M, 2*inner] (value | gate) -> out h16 [M, inner] = value * gelu_erf(gate)  */
int encdiff_op_geglu(const void* g, void* out, long long M, int inner, void* stream);

/* Concept-token MLPs of the encoder: replaces the per-unit Python loop of Encoder4.forward / Encoder4.warp
 * (ldm/modules/diffusionmodules/openaimodel_enc.py:1024-1041: for i in range(latent_unit): net[i](x[:, i][:, None]);
 * th.cat(out, dim=1)) with one launch.  All fp32 device pointers; parameters stacked over the L units:
 * w1/b1 [L,64] (Linear(1,64)), w2 [L,128,64], b2 [L,128], w3 [L,Dc,128], b3 [L,Dc]; x [B,L]; y [B,L*Dc].           */
int encdiff_concept_mlp_forward(const float* x, const float* w1, const float* b1, const float* w2, const float* b2,
                                const float* w3, const float* b3, float* y, int B, int L, int Dc, void* stream);
/* Backward of the above for dy [B,L*Dc]: overwrites dx [B,L] (may be NULL) and ADDS to dw1..db3 (same shapes as the
 * parameters; the caller zeroes them); the hidden layers are recomputed from x.                                       */
int encdiff_concept_mlp_backward(const float* x, const float* w1, const float* b1, const float* w2, const float* b2,
                                 const float* w3, const float* b3, const float* dy, float* dx, float* dw1, float* db1,
                                 float* dw2, float* db2, float* dw3, float* db3, int B, int L, int Dc, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Model level: the UNet denoiser.
 *   Replaces UNetModel.__init__/forward (ldm/modules/diffusionmodules/openaimodel_enc.py:443-748) and everything
 *   it calls: ResBlock (:179-275), Upsample/Downsample (:109-160), GroupNorm32 / timestep_embedding
 *   (ldm/modules/diffusionmodules/util.py:179-244), SpatialTransformer / BasicTransformerBlock / CrossAttention /
 *   FeedForward / GEGLU (ldm/modules/attention.py:37-64,152-261); the backward replaces what torch autograd
 *   (+ CheckpointFunction, util.py:147-176) computes for them.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct encdiff_unet_config {   /* the `params` of unet_config, configs/latent-diffusion/shapes3d-vq-4-16-encdiff.yaml:35-48 */
  int image_size, in_channels, model_channels, out_channels, num_res_blocks;
  int n_attention_resolutions;
  int attention_resolutions[8];
  int n_channel_mult;
  int channel_mult[8];
  int num_heads, context_dim, latent_unit, transformer_depth;
  int use_scale_shift_norm, resblock_updown;
} encdiff_unet_config;

typedef struct encdiff_unet encdiff_unet;

/* Build the architecture (host only, no CUDA call).  Unsupported combinations fail with a message. */
int encdiff_unet_create(const encdiff_unet_config* cfg /*host*/, encdiff_unet** out /*host*/);
int encdiff_unet_destroy(encdiff_unet* h);

/* Parameter inventory in the order of the reference UNetModel.named_parameters(); names are state_dict keys. */
int encdiff_unet_num_params(const encdiff_unet* h);
int encdiff_unet_param_info(const encdiff_unet* h, int i, char* name /*host*/, int name_cap, long long* shape4 /*host*/,
                            int* ndim /*host*/);

/* Sizes of the two caller-owned device buffers (bytes): the packed h16 weight arena and the activation workspace
 * for a batch (training != 0 keeps every activation and adds gradient buffers). */
long long encdiff_unet_packed_bytes(const encdiff_unet* h);
long long encdiff_unet_workspace_bytes(encdiff_unet* h, int batch, int training);

/* Attach fp32 parameter tensors (reference layouts), optional fp32 gradient tensors (same shapes; gradients are
 * ACCUMULATED into them) and the packed arena (1 KB aligned).  params/grads are host arrays of device pointers in
 * inventory order.  Synchronises `stream` once (not a hot call). */
int encdiff_unet_bind(encdiff_unet* h, const float* const* params /*host*/, float* const* grads /*host, may be NULL*/,
                      void* packed, void* stream);
/* Re-pack fp32 parameters into h16 GEMM operands (forward and transposed layouts).  Call after every optimiser /
 * EMA update of the parameters (the reference mutates them in place, ema.py:46-76). */
int encdiff_unet_pack_weights(encdiff_unet* h, void* stream);
/* Assign the workspace (1 KB aligned, >= workspace_bytes(batch, training)); cheap when unchanged. */
int encdiff_unet_plan(encdiff_unet* h, int batch, void* workspace, int training);

/* eps = UNet(x_t, t, context).  x [B,Cin,S,S] fp32 NCHW, t [B] fp32 (timesteps, may be fractional), ctx
 * [B, latent_unit*context_dim] fp32, out [B,Cout,S,S] fp32.  When noise != NULL the q_sample of p_losses
 * (ldm/models/diffusion/ddpm_enc.py:292-295) is fused in front: x_t = sqrt_acp[b]*x + sqrt_1macp[b]*noise (per-sample
 * fp32 coefficients); x_t is also written to x_t_out when non-NULL. */
int encdiff_unet_forward(encdiff_unet* h, int batch, const float* x, const float* noise, const float* sqrt_acp,
                         const float* sqrt_1macp, const float* t, const float* ctx, float* out, float* x_t_out,
                         void* stream);
/* fp32 VALIDATION forward (the north star's "fp32 validation build", eps max relative error <= 1e-4 against the
 * reference's fp32 result): the same UNetModel.forward (openaimodel_enc.py:712-748) with fp32 tensors, fp32 FMA
 * contractions on the CUDA cores and the bound fp32 master parameters read in the reference's layouts -- no 16-bit
 * operands, no tensor cores, slow by design.  A checker that runs on the GPU at any batch (tests compare the 16-bit
 * product path with it at B = 256); needs bind() only, not plan().  workspace: 256-byte aligned,
 * >= encdiff_unet_fp32_workspace_bytes(h, batch). */
long long encdiff_unet_fp32_workspace_bytes(encdiff_unet* h, int batch);
int encdiff_unet_forward_fp32(encdiff_unet* h, int batch, const float* x, const float* t, const float* ctx, float* out,
                              void* workspace, long long workspace_bytes, void* stream);
/* Backward of the last forward on this handle (training plan).  dout [B,Cout,S,S] fp32; writes dx (may be NULL) and
 * dctx [B, latent_unit*context_dim] fp32 (overwritten); parameter gradients are accumulated into the bound tensors. */
int encdiff_unet_backward(encdiff_unet* h, int batch, const float* dout, const float* ctx, float* dx, float* dctx,
                          void* stream);
/* Data-parallel overlap (what DistributedDataParallel's bucketed all-reduce gives the reference, main_val.py:656-660).
 * The weight gradients of a step only become complete in the deferred launches at the very end of the backward; with a split
 * pointer into the caller's flat gradient bucket, encdiff_unet_backward() completes every gradient at or above `split`
 * (group A: in registration order the bucket's tail, i.e. the output blocks, whose activation gradients are produced
 * first) and leaves the rest to encdiff_unet_backward_tail(), so the caller can all-reduce [split, end) on another stream
 * while group B is computed.  split = NULL (default): encdiff_unet_backward() does everything, the tail call is a no-op.
 * Takes effect at the next backward (which rebuilds its tables: not capturable, run it eagerly once). */
int encdiff_unet_set_grad_split(encdiff_unet* h, const float* split /*device*/);
int encdiff_unet_backward_tail(encdiff_unet* h, void* stream);

/* Live per-launch timing for bench.py's roofline figure: while enabled, forward/backward record one CUDA event after
 * every kernel launch on the launching stream.  profile_read synchronises on the last event, writes one text line
 * "family,launches,total_ms,total_flops,total_algorithmic_bytes" per kernel family into buf and clears the records; returns the byte count,
 * or -(needed) when cap is too small, or -1 on error. */
int encdiff_unet_profile(encdiff_unet* h, int enable);
long long encdiff_unet_profile_read(encdiff_unet* h, char* buf /*host*/, long long cap);

/* Debug / parity-test hook: device pointer and shape of the NHWC h16 output of one block of the LAST forward.
 * section 0 = input_blocks[index], 1 = middle_block, 2 = output_blocks[index]; the tensor is [B, H, W, C].  The
 * reference exposes the same tensors through forward hooks on input_blocks / middle_block / output_blocks. */
int encdiff_unet_debug_block(encdiff_unet* h, int section, int index, void** ptr /*host*/, int* C /*host*/,
                             int* H /*host*/, int* W /*host*/);

/* ---------------------------------------------------------------------------------------------------------------
 * Diffusion glue around the UNet.
 * ------------------------------------------------------------------------------------------------------------- */
/* loss[0] = mean|eps_hat - noise| (loss_type l1, ddpm_enc.py:297-301,1198-1213); deps (may be NULL) = d loss / d eps_hat */
int encdiff_l1_loss(const float* eps_hat, const float* noise, float* loss, float* deps, long long n, void* stream);
/* One DDIM update (DDIMSampler.p_sample_ddim, ldm/models/diffusion/ddim.py:192-206), fp32, same operation order.
 * coef4 = {a_t, a_prev, sigma_t, sqrt_one_minus_at} on the device; noise may be NULL (eta = 0). */
int encdiff_ddim_update(const float* x, const float* e_t, const float* noise, const float* coef4, float temperature,
                        float* x_prev, float* pred_x0, long long n, void* stream);
/* Device-resident loop state of DDIMSampler.ddim_sampling (ddim.py:141-166) so that one captured step graph can be
 * replayed S times without host updates.  Tables are in loop order (row 0 = first, noisiest step): t_table [S] fp32,
 * coef_table [S,4] = {a_t, a_prev, sigma_t, sqrt_one_minus_at}, noise_table [S,n] or NULL.
 * ddim_select: i = cursor[0]; ts[0..batch) = t_table[i] (the `ts = torch.full((b,), step)` of ddim.py:143); idx[0] = i;
 * cursor[0] = i + 1.  ddim_update_indexed: encdiff_ddim_update with row idx[0] of the tables. */
int encdiff_ddim_select(const float* t_table, int* cursor, float* ts, int* idx, int batch, void* stream);
int encdiff_ddim_update_indexed(const float* x, const float* e_t, const float* noise_table, const float* coef_table,
                                const int* idx, float temperature, float* x_prev, float* pred_x0, long long n,
                                void* stream);
/* timestep_embedding (ldm/modules/diffusionmodules/util.py:179-199): t [B] fp32 -> h16 [B, dim], cosine half first */
int encdiff_timestep_embedding(const float* t, void* out_h16, int batch, int dim, void* stream);
/* Fused AdamW (+EMA when ema != NULL) over flat fp32 buffers: torch.optim.AdamW as configured by
 * LatentDiffusion.configure_optimizers (ddpm_enc.py:1598-1639) and LitEma.forward (ldm/modules/ema.py:25-44). */
int encdiff_adamw_step(float* p, const float* g, float* m, float* v, float* ema, long long n, float lr, float beta1,
                       float beta2, float eps, float weight_decay, int step, float ema_decay, void* stream);

/* Graph-replayable form: the step-dependent scalars come from device memory, hyper = {lr, 1-beta1^t, 1-beta2^t,
 * ema_decay}; g is multiplied by grad_scale first (1/world_size after a sum all-reduce). */
int encdiff_adamw_step_dev(float* p, const float* g, float* m, float* v, float* ema, long long n, const float* hyper,
                           float beta1, float beta2, float eps, float weight_decay, float grad_scale, void* stream);

/* Device-resident optimiser clock (enqueue-only, graph capturable): step[0] += 1, then hyper = {lr[0], 1-beta1^step,
 * 1-beta2^step, min(ema_decay_max, (1+step)/(10+step))} -- the bias corrections of torch.optim.AdamW
 * (ddpm_enc.py:1598-1604) and LitEma's decay warm-up (ldm/modules/ema.py:31-33).  step (int32), lr and hyper are device
 * pointers. */
int encdiff_adamw_hyper_advance(int* step, const float* lr, float* hyper, float beta1, float beta2, float ema_decay_max,
                                void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * First-stage VQ-f4 ENCODE path: z = quant_conv(Encoder(x)), the frozen front end of every
 * training step.  Replaces VQModelInterface.encode (ldm/models/autoencoder.py:313-316) = Encoder.forward
 * (ldm/modules/diffusionmodules/model.py:437-459: conv_in, ResnetBlock :121-143, Downsample :72-79, mid AttnBlock
 * :178-204, norm_out / swish / conv_out) followed by the 1x1 quant_conv.  Inference only (the first stage is frozen,
 * ddpm_enc.py instantiate_first_stage).  The handle is an encdiff_unet: num_params / param_info (names = state_dict keys
 * `encoder.*`, `quant_conv.*`, reference registration order) / packed_bytes / bind (grads NULL) / pack_weights / destroy
 * are the calls above.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct encdiff_vqenc_config {   /* first_stage_config.params of shapes3d-vq-4-16-encdiff.yaml:52-75 */
  int ch, n_ch_mult;
  int ch_mult[4];
  int num_res_blocks, in_channels, resolution, z_channels, embed_dim;
} encdiff_vqenc_config;
int encdiff_vqenc_create(const encdiff_vqenc_config* cfg /*host*/, encdiff_unet** out /*host*/);
long long encdiff_vqenc_workspace_bytes(encdiff_unet* h, int batch);
/* x [B, in_channels, R, R] fp32 NCHW in [-1, 1] -> z [B, embed_dim, R/2^(levels-1), R/2^(levels-1)] fp32 NCHW.
 * workspace: 1 KB aligned, >= encdiff_vqenc_workspace_bytes(h, batch).  Enqueue-only. */
int encdiff_vqenc_forward(encdiff_unet* h, int batch, const float* x, float* z, void* workspace, void* stream);

/* First-stage DECODE path: img = Decoder(post_quant_conv([quantize(h) | s])).  Replaces
 * VQModelInterface.decode (ldm/models/autoencoder.py:328-369): nearest-codebook-entry quantisation (taming's
 * VectorQuantizer2 forward value; skipped with force_not_quantize), channel concatenation of the disentangled
 * representation s [B, disentangled_dim] broadcast over the latent grid (zeros when NULL; only when the model is built with
 * disentangled_dim > 0, i.e. use_disentangled_concat), the 1x1 post_quant_conv, and Decoder.forward
 * (ldm/modules/diffusionmodules/model.py:532-568: conv_in, mid ResnetBlock / AttnBlock / ResnetBlock, per level
 * num_res_blocks + 1 ResnetBlocks and Upsample = nearest x2 + conv3x3 :41-57, norm_out / swish / conv_out).  Same handle
 * type and generic calls as the encoder; parameter names `decoder.*`, `quantize.embedding.weight` (when n_embed > 0),
 * `post_quant_conv.*` in the reference's registration order (Decoder PREPENDS its levels: up.0 .. up.L-1). */
typedef struct encdiff_vqdec_config {
  encdiff_vqenc_config dd;          /* the shared ddconfig (in_channels unused) */
  int out_ch, n_embed, disentangled_dim;
} encdiff_vqdec_config;
int encdiff_vqdec_create(const encdiff_vqdec_config* cfg /*host*/, encdiff_unet** out /*host*/);
long long encdiff_vqdec_workspace_bytes(encdiff_unet* h, int batch);
/* h [B, embed_dim, r, r] fp32 NCHW latent; s [B, disentangled_dim] fp32 or NULL; quantize != 0 needs n_embed > 0;
 * img [B, out_ch, R, R] fp32 NCHW.  Enqueue-only; workspace 1 KB aligned. */
int encdiff_vqdec_forward(encdiff_unet* h, int batch, const float* latent, const float* s, int quantize, float* img,
                          void* workspace, void* stream);

/* Concept encoder trunk (the convolutional half): codes = Encoder4.encoder(images)
 * (ldm/modules/diffusionmodules/openaimodel_enc.py:995-1014: four 4x4 stride-2 convolutions with BatchNorm2d / ReLU, two
 * EncResBlocks :969-989 with a BatchNorm between them, View, Linear(16 d -> latent_unit)), forward AND backward (the concept
 * encoder is trained with the UNet).  Same handle type and generic calls as the UNet: parameter names `encoder.N...` in the
 * reference's registration order; bind() with gradient tensors for training.  BatchNorm follows torch.nn.BatchNorm2d
 * (eps 1e-5): training != 0 normalises with batch statistics and updates the bound running statistics with `momentum`;
 * training == 0 normalises with the running statistics.  The per-unit MLPs behind the trunk are encdiff_concept_mlp_*. */
typedef struct encdiff_enc4_config {   /* cond_stage_config.params of shapes3d-vq-4-16-encdiff.yaml:77-82 (+ image size) */
  int d, latent_unit, num_channels, resolution, bn;
} encdiff_enc4_config;
int encdiff_enc4_create(const encdiff_enc4_config* cfg /*host*/, encdiff_unet** out /*host*/);
int encdiff_enc4_num_batchnorm(const encdiff_unet* h);
/* running_mean / running_var device pointers of every BatchNorm2d in module order: stats[2 i], stats[2 i + 1] */
int encdiff_enc4_bind_stats(encdiff_unet* h, float* const* stats /*host*/);
long long encdiff_enc4_workspace_bytes(encdiff_unet* h, int batch);
/* images [B, num_channels, R, R] fp32 NCHW -> codes [B, latent_unit] fp32; keeps what backward needs in the workspace */
int encdiff_enc4_forward(encdiff_unet* h, int batch, const float* images, float* codes, void* workspace, int training,
                         float momentum, void* stream);
/* backward of the last forward (same workspace, same training flag): dcodes [B, latent_unit] fp32; parameter gradients are
 * ACCUMULATED into the bound gradient tensors */
int encdiff_enc4_backward(encdiff_unet* h, int batch, const float* dcodes, void* workspace, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ENCDIFF_B200_H_ */
