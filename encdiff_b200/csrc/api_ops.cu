// C-ABI, operator level: the two tensor-core GEMM families exposed with raw pointers so that each can be checked
// against the reference nn.Conv2d / nn.Linear call sites on its own (tests/test_gemm_gpu.py).
#include "../../include/encdiff_b200.h"
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "gemm.cuh"
#include "host.cuh"
#include "kernels.cuh"

using namespace edb;

extern "C" const char* encdiff_last_error(void) { return last_error(); }

extern "C" int encdiff_op_sgemm(int nsrc, const void* const* a_ptrs, const int* a_channels, const int* a_taps, int H,
                                int W, int B, const void* w_packed, long long w_rows, int N, int tap_sign,
                                const float* bias, const void* res, int ldres, int act, int out_fp32, void* out0,
                                int ld0, int split_col, void* out1, int ld1, int bn, void* stream) {
  EDB_REQUIRE(nsrc >= 1 && nsrc <= 3, "sgemm: 1..3 A sources");
  SgemmParams p;
  memset(&p, 0, sizeof(p));
  EDB_REQUIRE(make_geom(H, W, &p.g), "sgemm: H*W must be a power of two with W <= 128");
  long long K = 0;
  for (int s = 0; s < nsrc; ++s) {
    EDB_REQUIRE(a_taps[s] == 1 || a_taps[s] == 9, "sgemm: taps must be 1 or 9");
    EDB_REQUIRE(a_channels[s] % 8 == 0, "sgemm: channels must be a multiple of 8");
    int rc = make_act_tmap(&p.amap[s], a_ptrs[s], a_channels[s], H, W, B, p.g);
    if (rc) return rc;
    p.src_chunks[s] = (a_channels[s] + 63) / 64;
    p.src_taps[s] = a_taps[s];
    K += 64LL * p.src_chunks[s] * a_taps[s];
  }
  const bool old_only = getenv("ENCDIFF_OLD_GEMM") != nullptr;
  if (bn <= 0 && !out_fp32 && !out1 && ld0 == N && !old_only) {
    const int sms = device_sm_count();
    const int m_tiles = (B * H * W + 127) / 128;
    const int bn2 = sgemm2_pick_bn(m_tiles, N, sms, static_cast<int>(K / 64));
    if (bn2) {   // persistent kernel (gemm2.cu)
      Sgemm2Params q;
      memset(&q, 0, sizeof(q));
      for (int s = 0; s < nsrc; ++s) {
        q.amap[s] = p.amap[s];
        q.src_chunks[s] = p.src_chunks[s];
        q.src_taps[s] = p.src_taps[s];
      }
      q.nsrc = nsrc;
      q.g = p.g;
      int rc2 = make_w_tmap(&q.bmap, w_packed, K, w_rows, bn2);
      if (rc2) return rc2;
      rc2 = make_out_tmap(&q.omap, out0, static_cast<long long>(B) * H * W, N, ld0);
      if (rc2) return rc2;
      q.tap_sign = tap_sign;
      q.M = B * H * W;
      q.N = N;
      q.m_tiles = m_tiles;
      q.bias = bias;
      q.res = res;
      q.ldres = ldres;
      q.act = act;
      EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_sgemm2(q, bn2, sms, static_cast<cudaStream_t>(stream))));
      return EDB_OK;
    }
  }
  if (bn <= 0) bn = sgemm_pick_bn(N);
  EDB_REQUIRE(w_rows >= ((N + bn - 1) / bn) * bn, "sgemm: packed weight rows must be padded to the N tile");
  int rc = make_w_tmap(&p.bmap, w_packed, K, w_rows, bn);
  if (rc) return rc;
  p.nsrc = nsrc;
  p.tap_sign = tap_sign;
  p.M = B * H * W;
  p.N = N;
  p.bias = bias;
  p.res = res;
  p.ldres = ldres;
  p.act = act;
  p.out_fp32 = out_fp32;
  p.out0 = out0;
  p.ld0 = ld0;
  p.split_col = out1 ? split_col : (1 << 30);
  p.out1 = out1;
  p.ld1 = ld1;
  EDB_REQUIRE(!out1 || (split_col % 32) == 0, "sgemm: split_col must be a multiple of 32");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_sgemm(p, bn, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_wgrad(const void* dy, int Cp, const void* x, int Cq, int H, int W, int B, int taps,
                                float* out, long long stride_p, long long stride_q, long long stride_tap, int q_off,
                                int bn, int splits, void* stream) {
  WgradParams p;
  memset(&p, 0, sizeof(p));
  EDB_REQUIRE(taps == 1 || taps == 9, "wgrad: taps must be 1 or 9");
  EDB_REQUIRE(make_geom(H, W, &p.g), "wgrad: H*W must be a power of two with W <= 128");
  int rc = make_act_tmap(&p.pmap, dy, Cp, H, W, B, p.g);
  if (rc) return rc;
  rc = make_act_tmap(&p.qmap, x, Cq, H, W, B, p.g);
  if (rc) return rc;
  p.taps = taps;
  p.m_tiles = (B * H * W + 127) / 128;
  p.Cp = Cp;
  p.Cq = Cq;
  p.out = out;
  p.stride_p = stride_p;
  p.stride_q = stride_q;
  p.stride_tap = stride_tap;
  p.q_off = q_off;
  if (bn <= 0) bn = Cq <= 64 ? 64 : 128;
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_wgrad(p, bn, splits, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

// 0 = fp16, 1 = bf16: the 16-bit storage / operand type this library was built with (types.cuh)
extern "C" int encdiff_act_format(void) { return EDB_FMT; }

// ---- HBM-bound operators on their own (tests/test_ops_gpu.py; encdiff_b200/attention.py) ---------------------------
extern "C" int encdiff_op_group_norm(const void* x, const float* gamma, const float* beta, float eps, int silu,
                                     void* out, float* stats, int B, int H, int W, int C, void* stream) {
  EDB_REQUIRE(C % 32 == 0 && C % 8 == 0, "group_norm: C must be a multiple of 32");
  GnParams p;
  memset(&p, 0, sizeof(p));
  p.x0 = static_cast<const h16*>(x);
  p.C0 = C;
  p.H = H;
  p.W = W;
  p.gamma = gamma;
  p.beta = beta;
  p.eps = eps;
  p.silu = silu;
  p.resample = RESAMPLE_NONE;
  p.stats = stats;
  p.act = static_cast<h16*>(out);
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_gn_fwd(p, B, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_layer_norm(const void* x, const float* gamma, const float* beta, float eps, void* y,
                                     float* stats, long long M, int C, void* stream) {
  EDB_REQUIRE(C % 8 == 0, "layer_norm: C must be a multiple of 8");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_ln_fwd(static_cast<const h16*>(x), gamma, beta, eps, static_cast<h16*>(y),
                                                        stats, static_cast<int>(M), C, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_self_attention(const void* qkv, void* out, float* lse, int B, int N, int heads, int D,
                                         void* stream) {
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_self_attn_fwd(static_cast<const h16*>(qkv), static_cast<h16*>(out), lse,
                                                               B, N, heads, D, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_self_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                                             void* dqkv, int B, int N, int heads, int D, void* stream) {
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_self_attn_bwd(static_cast<const h16*>(qkv), static_cast<const h16*>(out),
                                                               static_cast<const h16*>(dout), lse,
                                                               static_cast<h16*>(dqkv), B, N, heads, D,
                                                               static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_cross_kv_project(const float* ctx, const float* wk, const float* wv, float* kv,
                                           void* table_scratch, int B, int C, int T, int Dc, void* stream) {
  EDB_REQUIRE(ctx && wk && wv && kv && table_scratch, "cross_kv_project: null argument");
  EDB_REQUIRE(2 * C <= 1024, "cross_kv_project: heads * d_head must be <= 512");
  const CrossKvEntry en{wk, wv, kv, C, 0};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  EDB_CHECK_CUDA(cudaMemcpyAsync(table_scratch, &en, sizeof(en), cudaMemcpyHostToDevice, s));
  EDB_CHECK_CUDA(cudaStreamSynchronize(s));      // `en` lives on this stack frame
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_cross_kv_project(static_cast<const CrossKvEntry*>(table_scratch), 1, ctx, B, T,
                                                                  Dc, s)));
  return EDB_OK;
}

extern "C" int encdiff_op_cross_attention(const void* q, const float* ctx, const float* wk, const float* wv,
                                          const float* kv, void* out, int B, int N, int heads, int D, int T, int Dc,
                                          void* stream) {
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_cross_attn_fwd(static_cast<const h16*>(q), ctx, wk, wv,
                                                                static_cast<h16*>(out), B, N, heads, D, T, Dc,
                                                                static_cast<cudaStream_t>(stream), kv)));
  return EDB_OK;
}

extern "C" long long encdiff_op_cross_attention_bwd_scratch_bytes(int B, int heads, int D, int T) {
  return static_cast<long long>(cross_attn_bwd_scratch_bytes(B, heads, D, T));
}

extern "C" int encdiff_op_cross_attention_bwd(const void* q, const float* ctx, const float* wk, const float* wv,
                                              const float* kv, const void* dout, void* dq, float* dwk, float* dwv,
                                              float* dctx,
                                              float* scratch, int B, int N, int heads, int D, int T, int Dc,
                                              void* stream) {
  EDB_REQUIRE(scratch != nullptr, "cross_attention_bwd: scratch of encdiff_op_cross_attention_bwd_scratch_bytes() needed");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_cross_attn_bwd(static_cast<const h16*>(q), ctx, wk, wv,
                                                                static_cast<const h16*>(dout), static_cast<h16*>(dq), dwk,
                                                                dwv, dctx, nullptr, scratch, B, N, heads, D, T, Dc,
                                                                static_cast<cudaStream_t>(stream), false, kv)));
  return EDB_OK;
}

static ConceptParams concept_params(const float* x, const float* w1, const float* b1, const float* w2, const float* b2,
                                    const float* w3, const float* b3, int B, int L, int Dc) {
  ConceptParams p;
  memset(&p, 0, sizeof(p));
  p.x = x;
  p.w1 = w1;
  p.b1 = b1;
  p.w2 = w2;
  p.b2 = b2;
  p.w3 = w3;
  p.b3 = b3;
  p.B = B;
  p.L = L;
  p.Dc = Dc;
  return p;
}

extern "C" int encdiff_concept_mlp_forward(const float* x, const float* w1, const float* b1, const float* w2,
                                           const float* b2, const float* w3, const float* b3, float* y, int B, int L,
                                           int Dc, void* stream) {
  EDB_REQUIRE(x && w1 && b1 && w2 && b2 && w3 && b3 && y, "concept_mlp_forward: null argument");
  EDB_REQUIRE(Dc >= 1 && Dc <= 32, "concept_mlp: context_dim must be in 1..32");
  ConceptParams p = concept_params(x, w1, b1, w2, b2, w3, b3, B, L, Dc);
  p.y = y;
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_concept_mlp_fwd(p, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_concept_mlp_backward(const float* x, const float* w1, const float* b1, const float* w2,
                                            const float* b2, const float* w3, const float* b3, const float* dy,
                                            float* dx, float* dw1, float* db1, float* dw2, float* db2, float* dw3,
                                            float* db3, int B, int L, int Dc, void* stream) {
  EDB_REQUIRE(x && w1 && b1 && w2 && b2 && w3 && b3 && dy && dw1 && db1 && dw2 && db2 && dw3 && db3,
              "concept_mlp_backward: null argument");
  EDB_REQUIRE(Dc >= 1 && Dc <= 32, "concept_mlp: context_dim must be in 1..32");
  ConceptParams p = concept_params(x, w1, b1, w2, b2, w3, b3, B, L, Dc);
  p.dy = dy;
  p.dx = dx;
  p.dw1 = dw1;
  p.db1 = db1;
  p.dw2 = dw2;
  p.db2 = db2;
  p.dw3 = dw3;
  p.db3 = db3;
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_concept_mlp_bwd(p, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_op_geglu(const void* g, void* out, long long M, int inner, void* stream) {
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_geglu_fwd(static_cast<const h16*>(g), static_cast<h16*>(out), M, inner,
                                                           static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}
