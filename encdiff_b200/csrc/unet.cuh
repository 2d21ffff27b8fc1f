// Host-side description of the EncDiff UNet denoiser: architecture walk, parameter inventory, packed-weight layout,
// activation plan and the forward / backward launch sequences.  Mirrors UNetModel.__init__ / forward
// (ldm/modules/diffusionmodules/openaimodel_enc.py:443-748) for the spatial-transformer + resblock_updown branch.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/encdiff_b200.h"
#include "gemm.cuh"
#include "kernels.cuh"

namespace edb {

struct ParamInfo {
  std::string name;
  int ndim;
  long long shape[4];
  long long numel;
};

// a block-level activation (NHWC h16) with its gradient buffer
struct Tn {
  h16* p = nullptr;
  h16* g = nullptr;
  int C = 0, H = 0, W = 0;
  bool ginit = false;
};

// one packed GEMM operand: h16 [rows_pad, K] at element offset `off` of the packed h16 arena
struct Packed {
  long long off = 0;
  long long rows = 0;   // padded rows
  long long K = 0;
};

struct ResLayer {
  std::string prefix;
  int c0 = 0, c1 = 0, cout = 0;       // input channels (main, concat skip), output channels
  bool up = false, down = false, skipconv = false;
  int Hin = 0, Win = 0, Hout = 0, Wout = 0;
  int film_off = 0;
  // parameter indices
  int gn1_w, gn1_b, conv1_w, conv1_b, emb_w, emb_b, gn2_w, gn2_b, conv2_w, conv2_b, skip_w = -1, skip_b = -1;
  Packed w1f, w1t, w2f, w2t, wst;     // conv1 fwd/T, conv2(+skip) fwd, conv2 T, skip T
  long long bias2_off = -1;           // fused conv2+skip bias in the fp32 arena (skipconv only)
  // plan
  Tn *x0 = nullptr, *x1 = nullptr, *y = nullptr;
  h16 *a1 = nullptr, *xs = nullptr, *h1 = nullptr, *a2 = nullptr;
  float *st1 = nullptr, *st2 = nullptr;
};

struct StLayer {
  std::string prefix;
  int C = 0, heads = 0, D = 0, H = 0, W = 0;
  int norm_w, norm_b, pin_w, pin_b, q1_w, k1_w, v1_w, o1_w, o1_b, ff1_w, ff1_b, ff2_w, ff2_b, q2_w, k2_w, v2_w, o2_w, o2_b,
      ln1_w, ln1_b, ln2_w, ln2_b, ln3_w, ln3_b, pout_w, pout_b;
  Packed pinf, pint, qkvf, qkvt, o1f, o1t, q2f, q2t, o2f, o2t, ff1f, ff1t, ff2f, ff2t, poutf, poutt;
  Tn *x = nullptr, *y = nullptr;
  h16 *n = nullptr, *t0 = nullptr, *l1 = nullptr, *qkv = nullptr, *att = nullptr, *t1 = nullptr, *l2 = nullptr,
       *q2 = nullptr, *att2 = nullptr, *t2 = nullptr, *l3 = nullptr, *g = nullptr, *gg = nullptr, *t3 = nullptr;
  float *stn = nullptr, *ls1 = nullptr, *ls2 = nullptr, *ls3 = nullptr, *lse = nullptr;
  float* kv = nullptr;    // [B, T, 2C] to_k | to_v of the context, projected once per forward for all layers
};

// first-stage VQ encoder layers (vq_enc.inc): the same UNet object carries their parameters / packed operands
struct VqLayer {
  int kind = 0;            // 1 ResnetBlock, 2 Downsample (3x3 stride 2, pad right/bottom), 3 AttnBlock, 4 Upsample (nearest x2 + 3x3)
  int cin = 0, cout = 0, H = 0;     // H = input spatial size (square)
  int n1w = -1, n1b = -1, c1w = -1, c1b = -1, n2w = -1, n2b = -1, c2w = -1, c2b = -1, sw = -1, sb = -1;   // res / down (c1*) / attn
  int qw = -1, qb = -1, kw = -1, kb = -1, vw = -1, vb = -1, pw = -1, pb = -1;
  Packed w1, w2;           // res: conv1, conv2 (+ nin_shortcut slab); down: conv; attn: qkv, proj_out
  long long bias_off = -1; // fp32 arena: res with shortcut: conv2.bias + nin.bias; attn: q|k|v biases
};

struct Layer {
  int kind;   // 0 stem conv, 1 res, 2 st
  int idx;    // index into res / st vectors
};

// live per-launch timing (bench / tuning): one CUDA event after every launch of a step, on the launching stream;
// a launch's duration is the difference to the previous event (launches of one stream are serialised)
struct ProfRec {
  const char* name;
  double flops;
  double bytes;
  cudaEvent_t ev;
};

struct UNet {
  encdiff_unet_config cfg;
  bool prof = false;
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  size_t prof_used = 0;
  std::vector<ParamInfo> params;
  std::vector<ResLayer> res;
  std::vector<StLayer> st;
  std::vector<std::vector<Layer>> in_blocks, out_blocks;
  std::vector<Layer> mid;
  int head_ch = 0;
  int film_cols = 0;         // F = sum 2*cout
  int emb_ch = 0;            // 4 * model_channels
  // stem / head / time-embed parameters
  int stem_w, stem_b, te0_w, te0_b, te2_w, te2_b, hn_w, hn_b, hc_w, hc_b;
  Packed stemf, stemt, te0f, te0t, te2f, te2t, filmf, filmt, headf, headt;
  long long film_bias_off = 0;
  // packed arena layout
  std::vector<PackEntry> table;     // src pointers resolved at bind()
  std::vector<int> table_src, table_src2;   // parameter index of each entry
  long long pack_total = 0;         // total source elements walked by the pack kernel
  int pack_blocks = 0;              // blocks of the pack launch
  size_t table_bytes = 0, h16_bytes = 0, f32_bytes = 0;
  // binding
  std::vector<const float*> pptr;
  std::vector<float*> gptr;
  char* packed = nullptr;
  bool bound = false;
  // plan (per batch / workspace)
  int planB = -1;
  char* plan_ws = nullptr;
  bool plan_training = false;
  size_t ws_bytes_persist = 0, ws_bytes_scratch = 0;
  std::vector<Tn> tn;               // block-level tensors; [0] = stem output
  std::vector<int> hs_of_in_block;  // tn index of each input block's output
  h16 *temb = nullptr, *a1e = nullptr, *a2e = nullptr, *cols = nullptr, *aH = nullptr, *dyH = nullptr, *dfilm = nullptr;
  float *z1 = nullptr, *z2 = nullptr, *film = nullptr, *stH = nullptr, *yH = nullptr;
  float* gs = nullptr;       // {scale, 1/scale} of the activation gradients of the current backward (device)
  char* scratch = nullptr;
  size_t scratch_half = 0;   // the backward alternates between two scratch halves (layer parity)
  // side stream of the backward: weight-gradient GEMMs run concurrently with the input-gradient chain
  bool side_on = true;       // ENCDIFF_NO_SIDE_STREAM disables
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_side_done[2] = {nullptr, nullptr};
  // forward: the timestep-embedding / FiLM chain and the context K|V projection run on the side stream next to the stem and
  // the first layers; their first consumers wait for ev_side_done[0] (FiLM) / ev_side_done[1] (K|V) while these flags are set
  bool pend_film = false, pend_kv = false;
  // tensor-map caches
  std::map<std::tuple<const void*, int, int, int, int>, CUtensorMap> act_maps;
  std::map<std::tuple<const void*, long long, long long, int>, CUtensorMap> w_maps;
  std::map<std::tuple<const void*, long long, int, int>, CUtensorMap> out_maps;
  // conv3x3 weight-gradient staging (tap-major fp32, wgrad2.cu) + the finalize table that folds it into the grads
  struct ConvStage { int param; long long Cp, Cq; float* ptr; };
  std::vector<ConvStage> conv_stage;
  std::vector<WgradFinalizeEntry> fin_host;
  WgradFinalizeEntry* fin_dev = nullptr;
  long long fin_total = 0;
  bool fin_dirty = true;      // staging needs zeroing / table needs (re)building: after plan() or bind()
  std::map<std::tuple<const void*, int, int, int>, CUtensorMap> grad_maps;
  // deferred, grouped weight gradients (wgrad2_group_kernel): the backward only RECORDS its wgrad problems (their operand
  // pointers are fixed by plan()/bind(), so the table is built once, on the first backward after either) and launches them
  // all at once at the end; the per-layer activation-gradient scratch is therefore not reused inside one backward
  bool group_wgrad = true;    // ENCDIFF_NO_WGRAD_GROUP: launch per layer (side stream) instead
  std::vector<Wgrad2Params> wg_host;
  std::vector<int> wg_prefix;
  std::vector<const float*> wg_dst;   // reference-layout gradient tensor each recorded problem accumulates into
  // Gradient split for data parallelism (encdiff_unet_set_grad_split): problems / finalize entries whose gradient lies at or
  // above `grad_split` in the caller's flat bucket form group A, run by unet_backward(); the rest form group B, run by
  // unet_backward_tail(), so the caller can all-reduce the bucket's tail while group B is still being computed.
  const float* grad_split = nullptr;
  int wg_nA = 0, wg_totalA = 0, wg_totalB = 0, fin_nA = 0;
  long long fin_totalA = 0, fin_totalB = 0;
  Wgrad2Params* wg_dev = nullptr;
  int* wg_prefix_dev = nullptr;
  int wg_total = 0;
  std::vector<CrossKvEntry> ckv_host;    // K | V projection table (built on the first forward after plan() / bind())
  CrossKvEntry* ckv_dev = nullptr;
  bool ckv_dirty = true;
  std::vector<CrossProjEntry> cp_host;   // deferred cross-attention projection gradients (same life cycle as wg_host)
  CrossProjEntry* cp_dev = nullptr;
  int cp_max_C = 0;
  static constexpr int kWgCap = 640;   // problems per backward (the Shapes3D plan records 273)
  size_t scratch_total = 0;
  bool force_old_gemm = false;   // parity A/B switch: route every GEMM through the one-tile-per-CTA kernel
  // first-stage VQ encoder (model_kind 1): config, layer list, stem / head operands live in stemf / headf
  int model_kind = 0;
  encdiff_vqenc_config vqcfg;
  std::vector<VqLayer> vq;          // registration order (= parameter order)
  std::vector<int> vq_order;        // execution order (the decoder registers its levels in reverse)
  // concept-encoder trunk (model_kind 3, enc4.inc): parameter indices and operands of the fixed Encoder4.encoder sequence
  struct E4 {
    encdiff_enc4_config cfg;
    int cw[4], cb[4];            // the four 4x4 stride-2 convolutions (encoder.0 / 3 / 6 / 8)
    int bw[7], bb[7];            // BatchNorm2d: encoder.1, 4, 7, 9, 11.convs.2, 12, 14.convs.2
    int r3w[2], r3b[2], r1w[2], r1b[2];   // EncResBlock 3x3 / 1x1 convolutions (encoder.11, encoder.14)
    int lw, lb;                  // encoder.16 Linear
    Packed c0f, cf[4], ct[4], r3f[2], r3t[2], r1f[2], r1t[2], lf, lt;
    float* run[14] = {nullptr};  // running_mean / running_var of the 7 BatchNorms (bind_stats)
    int last_training = 1;
  } e4;
  int vq_n_embed = 0, vq_dd = 0, vq_emb_w = -1, vq_pq_w = -1, vq_pq_b = -1, vq_img_ch = 0;
  int vq_in_w = -1, vq_in_b = -1, vq_no_w = -1, vq_no_b = -1, vq_co_w = -1, vq_co_b = -1, vq_qc_w = -1, vq_qc_b = -1;
  int vq_out_res = 0, vq_out_ch = 0;

  h16* pbf() const { return reinterpret_cast<h16*>(packed + table_bytes); }
  float* pf32() const { return reinterpret_cast<float*>(packed + table_bytes + h16_bytes); }
};

int unet_build(UNet* u, const encdiff_unet_config* cfg);          // architecture + params + pack table
int unet_plan(UNet* u, int B, char* ws, bool training);           // assign workspace pointers
size_t unet_workspace_bytes(UNet* u, int B, bool training);
int unet_forward(UNet* u, int B, const float* x, const float* noise, const float* sa, const float* sb, const float* t,
                 const float* ctx, float* out, float* xt, cudaStream_t s);
int unet_backward(UNet* u, int B, const float* dout, const float* ctx, float* dx, float* dctx, cudaStream_t s);
// fp32 validation forward (fp32_ref.cu): fp32 tensors, CUDA-core FMA contractions, reference-layout master parameters
int unet_backward_tail(UNet* u, cudaStream_t s);
int vqenc_build(UNet* u, const encdiff_vqenc_config* cfg);
int vqdec_build(UNet* u, const encdiff_vqdec_config* cfg);
int enc4_build(UNet* u, const encdiff_enc4_config* cfg);
size_t enc4_workspace_bytes(UNet* u, int B);
int enc4_forward(UNet* u, int B, const float* img, float* codes, char* ws, int training, float momentum, cudaStream_t s);
int enc4_backward(UNet* u, int B, const float* dcodes, char* ws, cudaStream_t s);
size_t vqdec_workspace_bytes(UNet* u, int B);
int vqdec_forward(UNet* u, int B, const float* latent, const float* sdis, int quantize, float* img, char* ws, cudaStream_t s);
size_t vqenc_workspace_bytes(UNet* u, int B);
int vqenc_forward(UNet* u, int B, const float* x, float* z, char* ws, cudaStream_t s);
size_t unet_fp32_workspace_floats(UNet* u, int B);
int unet_forward_fp32(UNet* u, int B, const float* x, const float* t, const float* ctx, float* out, float* ws,
                      size_t ws_floats, cudaStream_t s);

}  // namespace edb
