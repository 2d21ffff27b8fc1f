// C-ABI, model level: the UNet denoiser handle and the diffusion glue kernels (see include/encdiff_b200.h).
#include <stdlib.h>
#include <string.h>

#include <new>

#include "host.cuh"
#include "unet.cuh"

using namespace edb;

struct encdiff_unet {
  UNet u;
};

extern "C" int encdiff_unet_create(const encdiff_unet_config* cfg, encdiff_unet** out) {
  EDB_REQUIRE(cfg && out, "unet_create: null argument");
  encdiff_unet* h = new (std::nothrow) encdiff_unet();
  EDB_REQUIRE(h, "unet_create: out of host memory");
  h->u.force_old_gemm = getenv("ENCDIFF_OLD_GEMM") != nullptr;
  h->u.side_on = getenv("ENCDIFF_NO_SIDE_STREAM") == nullptr;
  h->u.group_wgrad = getenv("ENCDIFF_NO_WGRAD_GROUP") == nullptr && !h->u.force_old_gemm;
  int rc = unet_build(&h->u, cfg);
  if (rc) {
    delete h;
    return rc;
  }
  *out = h;
  return EDB_OK;
}

extern "C" int encdiff_unet_destroy(encdiff_unet* h) {
  if (h) {
    UNet& u = h->u;
    for (cudaEvent_t ev : u.prof_pool) cudaEventDestroy(ev);
    for (cudaEvent_t ev : {u.ev_fork, u.ev_join, u.ev_side_done[0], u.ev_side_done[1]})
      if (ev) cudaEventDestroy(ev);
    if (u.side) cudaStreamDestroy(u.side);
  }
  delete h;
  return EDB_OK;
}

extern "C" int encdiff_unet_num_params(const encdiff_unet* h) { return h ? static_cast<int>(h->u.params.size()) : -1; }

extern "C" int encdiff_unet_param_info(const encdiff_unet* h, int i, char* name, int name_cap, long long* shape4,
                                       int* ndim) {
  EDB_REQUIRE(h && i >= 0 && i < static_cast<int>(h->u.params.size()), "param_info: bad index");
  const ParamInfo& p = h->u.params[i];
  if (name && name_cap > 0) {
    strncpy(name, p.name.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
  }
  if (shape4)
    for (int k = 0; k < 4; ++k) shape4[k] = p.shape[k];
  if (ndim) *ndim = p.ndim;
  return EDB_OK;
}

extern "C" long long encdiff_unet_packed_bytes(const encdiff_unet* h) {
  return h ? static_cast<long long>(h->u.table_bytes + h->u.h16_bytes + h->u.f32_bytes) : -1;
}

extern "C" long long encdiff_unet_workspace_bytes(encdiff_unet* h, int batch, int training) {
  if (!h || batch <= 0) return -1;
  return static_cast<long long>(unet_workspace_bytes(&h->u, batch, training != 0));
}

extern "C" int encdiff_unet_bind(encdiff_unet* h, const float* const* params, float* const* grads, void* packed,
                                 void* stream) {
  EDB_REQUIRE(h && params && packed, "unet_bind: null argument");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 1023) == 0, "unet_bind: packed arena must be 1 KB aligned");
  UNet& u = h->u;
  const size_t n = u.params.size();
  u.pptr.assign(params, params + n);
  if (grads) u.gptr.assign(grads, grads + n);
  else u.gptr.clear();
  for (size_t i = 0; i < n; ++i) {
    EDB_REQUIRE(u.pptr[i] != nullptr, "unet_bind: null parameter pointer");
    EDB_REQUIRE((reinterpret_cast<uintptr_t>(u.pptr[i]) & 3) == 0, "unet_bind: parameter pointer not 4-byte aligned");
  }
  for (size_t i = 0; i < u.table.size(); ++i) {
    u.table[i].src = u.pptr[u.table_src[i]];
    u.table[i].src2 = u.table_src2[i] >= 0 ? u.pptr[u.table_src2[i]] : nullptr;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  u.packed = static_cast<char*>(packed);
  EDB_CHECK_CUDA(cudaMemsetAsync(packed, 0, u.table_bytes + u.h16_bytes + u.f32_bytes, s));
  EDB_CHECK_CUDA(cudaMemcpyAsync(packed, u.table.data(), u.table.size() * sizeof(PackEntry), cudaMemcpyHostToDevice, s));
  EDB_CHECK_CUDA(cudaStreamSynchronize(s));   // the host table may be re-bound later; bind is not a hot call
  u.w_maps.clear();
  u.grad_maps.clear();
  u.fin_dirty = true;
  u.ckv_dirty = true;
  u.bound = true;
  return EDB_OK;
}

extern "C" int encdiff_unet_pack_weights(encdiff_unet* h, void* stream) {
  EDB_REQUIRE(h && h->u.bound, "pack_weights before bind");
  UNet& u = h->u;
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_pack(reinterpret_cast<const PackEntry*>(u.packed),
                                                      static_cast<int>(u.table.size()), u.pack_blocks, u.pbf(), u.pf32(),
                                                      static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_unet_plan(encdiff_unet* h, int batch, void* workspace, int training) {
  EDB_REQUIRE(h && workspace && batch > 0, "unet_plan: bad argument");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, "unet_plan: workspace must be 1 KB aligned");
  return unet_plan(&h->u, batch, static_cast<char*>(workspace), training != 0);
}

extern "C" int encdiff_vqenc_create(const encdiff_vqenc_config* cfg, encdiff_unet** out) {
  EDB_REQUIRE(cfg && out, "vqenc_create: null argument");
  encdiff_unet* h = new (std::nothrow) encdiff_unet();
  EDB_REQUIRE(h, "vqenc_create: out of host memory");
  int rc = vqenc_build(&h->u, cfg);
  if (rc) {
    delete h;
    return rc;
  }
  *out = h;
  return EDB_OK;
}

extern "C" long long encdiff_vqenc_workspace_bytes(encdiff_unet* h, int batch) {
  if (!h || batch <= 0 || h->u.model_kind != 1) return -1;
  return static_cast<long long>(vqenc_workspace_bytes(&h->u, batch));
}

extern "C" int encdiff_vqenc_forward(encdiff_unet* h, int batch, const float* x, float* z, void* workspace, void* stream) {
  EDB_REQUIRE(h && x && z && workspace && batch > 0, "vqenc_forward: bad argument");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, "vqenc_forward: workspace must be 1 KB aligned");
  return vqenc_forward(&h->u, batch, x, z, static_cast<char*>(workspace), static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_vqdec_create(const encdiff_vqdec_config* cfg, encdiff_unet** out) {
  EDB_REQUIRE(cfg && out, "vqdec_create: null argument");
  encdiff_unet* h = new (std::nothrow) encdiff_unet();
  EDB_REQUIRE(h, "vqdec_create: out of host memory");
  int rc = vqdec_build(&h->u, cfg);
  if (rc) {
    delete h;
    return rc;
  }
  *out = h;
  return EDB_OK;
}

extern "C" long long encdiff_vqdec_workspace_bytes(encdiff_unet* h, int batch) {
  if (!h || batch <= 0 || h->u.model_kind != 2) return -1;
  return static_cast<long long>(vqdec_workspace_bytes(&h->u, batch));
}

extern "C" int encdiff_vqdec_forward(encdiff_unet* h, int batch, const float* latent, const float* s, int quantize,
                                     float* img, void* workspace, void* stream) {
  EDB_REQUIRE(h && latent && img && workspace && batch > 0, "vqdec_forward: bad argument");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, "vqdec_forward: workspace must be 1 KB aligned");
  return vqdec_forward(&h->u, batch, latent, s, quantize, img, static_cast<char*>(workspace),
                       static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_enc4_create(const encdiff_enc4_config* cfg, encdiff_unet** out) {
  EDB_REQUIRE(cfg && out, "enc4_create: null argument");
  encdiff_unet* h = new (std::nothrow) encdiff_unet();
  EDB_REQUIRE(h, "enc4_create: out of host memory");
  int rc = enc4_build(&h->u, cfg);
  if (rc) {
    delete h;
    return rc;
  }
  h->u.side_on = false;        // one stream: the im2col matrices are overwritten by their gradients in place
  h->u.group_wgrad = false;
  *out = h;
  return EDB_OK;
}

extern "C" int encdiff_enc4_num_batchnorm(const encdiff_unet* h) { return h && h->u.model_kind == 3 ? 7 : -1; }

extern "C" int encdiff_enc4_bind_stats(encdiff_unet* h, float* const* stats) {
  EDB_REQUIRE(h && h->u.model_kind == 3, "enc4_bind_stats: not a concept-encoder handle");
  for (int i = 0; i < 14; ++i) h->u.e4.run[i] = stats ? stats[i] : nullptr;
  return EDB_OK;
}

extern "C" long long encdiff_enc4_workspace_bytes(encdiff_unet* h, int batch) {
  if (!h || batch <= 0 || h->u.model_kind != 3) return -1;
  return static_cast<long long>(enc4_workspace_bytes(&h->u, batch));
}

extern "C" int encdiff_enc4_forward(encdiff_unet* h, int batch, const float* images, float* codes, void* workspace,
                                    int training, float momentum, void* stream) {
  EDB_REQUIRE(h && images && codes && workspace && batch > 0, "enc4_forward: bad argument");
  EDB_REQUIRE(h->u.model_kind == 3, "enc4_forward: not a concept-encoder handle");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, "enc4_forward: workspace must be 1 KB aligned");
  return enc4_forward(&h->u, batch, images, codes, static_cast<char*>(workspace), training, momentum,
                      static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_enc4_backward(encdiff_unet* h, int batch, const float* dcodes, void* workspace, void* stream) {
  EDB_REQUIRE(h && dcodes && workspace && batch > 0, "enc4_backward: bad argument");
  EDB_REQUIRE(h->u.model_kind == 3, "enc4_backward: not a concept-encoder handle");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, "enc4_backward: workspace must be 1 KB aligned");
  return enc4_backward(&h->u, batch, dcodes, static_cast<char*>(workspace), static_cast<cudaStream_t>(stream));
}

extern "C" long long encdiff_unet_fp32_workspace_bytes(encdiff_unet* h, int batch) {
  if (!h || batch <= 0) return -1;
  return static_cast<long long>(unet_fp32_workspace_floats(&h->u, batch) * sizeof(float));
}

extern "C" int encdiff_unet_forward_fp32(encdiff_unet* h, int batch, const float* x, const float* t, const float* ctx,
                                         float* out, void* workspace, long long workspace_bytes, void* stream) {
  EDB_REQUIRE(h && x && t && ctx && out && workspace && batch > 0, "unet_forward_fp32: bad argument");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "unet_forward_fp32: workspace must be 256-byte aligned");
  return unet_forward_fp32(&h->u, batch, x, t, ctx, out, static_cast<float*>(workspace),
                           static_cast<size_t>(workspace_bytes) / sizeof(float), static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_unet_forward(encdiff_unet* h, int batch, const float* x, const float* noise, const float* sqrt_acp,
                                    const float* sqrt_1macp, const float* t, const float* ctx, float* out, float* x_t,
                                    void* stream) {
  EDB_REQUIRE(h && x && t && ctx && out, "unet_forward: null argument");
  EDB_REQUIRE(!noise || (sqrt_acp && sqrt_1macp), "unet_forward: q_sample needs both coefficient arrays");
  return unet_forward(&h->u, batch, x, noise, sqrt_acp, sqrt_1macp, t, ctx, out, x_t, static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_unet_backward(encdiff_unet* h, int batch, const float* dout, const float* ctx, float* dx,
                                     float* dctx, void* stream) {
  EDB_REQUIRE(h, "unet_backward: null handle");
  return unet_backward(&h->u, batch, dout, ctx, dx, dctx, static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_unet_set_grad_split(encdiff_unet* h, const float* split) {
  EDB_REQUIRE(h, "unet_set_grad_split: null handle");
  if (h->u.grad_split != split) {
    h->u.grad_split = split;
    h->u.fin_dirty = true;           // the next backward re-partitions its deferred-gradient tables
  }
  return EDB_OK;
}

extern "C" int encdiff_unet_backward_tail(encdiff_unet* h, void* stream) {
  EDB_REQUIRE(h, "unet_backward_tail: null handle");
  return unet_backward_tail(&h->u, static_cast<cudaStream_t>(stream));
}

extern "C" int encdiff_unet_profile(encdiff_unet* h, int enable) {
  EDB_REQUIRE(h, "unet_profile: null handle");
  h->u.prof = enable != 0;
  h->u.prof_recs.clear();
  h->u.prof_used = 0;
  return EDB_OK;
}

extern "C" long long encdiff_unet_profile_read(encdiff_unet* h, char* buf, long long cap) {
  if (!h || !buf || cap <= 0) return -1;
  UNet& u = h->u;
  if (!u.prof_recs.empty() && cudaEventSynchronize(u.prof_recs.back().ev) != cudaSuccess) return -1;
  std::map<std::string, std::tuple<long long, double, double, double>> agg;   // launches, ms, flops, bytes
  for (size_t i = 1; i < u.prof_recs.size(); ++i) {
    const ProfRec& r = u.prof_recs[i];
    if (r.name[0] == '(') continue;   // a "(begin)" marker: the gap before it is host time between calls
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, u.prof_recs[i - 1].ev, r.ev) != cudaSuccess) continue;
    auto& a = agg[r.name];
    std::get<0>(a) += 1;
    std::get<1>(a) += ms;
    std::get<2>(a) += r.flops;
    std::get<3>(a) += r.bytes;
  }
  std::string out;
  char line[256];
  for (auto& kv : agg) {
    snprintf(line, sizeof(line), "%s,%lld,%.6f,%.6e,%.6e\n", kv.first.c_str(), std::get<0>(kv.second),
             std::get<1>(kv.second), std::get<2>(kv.second), std::get<3>(kv.second));
    out += line;
  }
  u.prof_recs.clear();
  u.prof_used = 0;
  const long long n = static_cast<long long>(out.size());
  if (n + 1 > cap) return -(n + 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  return n;
}

extern "C" int encdiff_unet_debug_block(encdiff_unet* h, int section, int index, void** ptr, int* C, int* H, int* W) {
  EDB_REQUIRE(h && ptr && C && H && W, "debug_block: null argument");
  UNet& u = h->u;
  EDB_REQUIRE(u.planB > 0 && u.plan_ws, "debug_block before plan");
  const std::vector<Layer>* ls = nullptr;
  if (section == 0 && index >= 0 && index < static_cast<int>(u.in_blocks.size())) ls = &u.in_blocks[index];
  else if (section == 1) ls = &u.mid;
  else if (section == 2 && index >= 0 && index < static_cast<int>(u.out_blocks.size())) ls = &u.out_blocks[index];
  EDB_REQUIRE(ls && !ls->empty(), "debug_block: bad section / index");
  const Layer& l = ls->back();
  const Tn* t = l.kind == 0 ? &u.tn[0] : (l.kind == 1 ? u.res[l.idx].y : u.st[l.idx].y);
  *ptr = t->p;
  *C = t->C;
  *H = t->H;
  *W = t->W;
  return EDB_OK;
}

// ---- diffusion glue ---------------------------------------------------------------------------------------------
extern "C" int encdiff_l1_loss(const float* eps_hat, const float* noise, float* loss, float* deps, long long n,
                               void* stream) {
  EDB_REQUIRE(eps_hat && noise && loss && n > 0, "l1_loss: bad argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  EDB_CHECK_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), s));
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_l1_loss(eps_hat, noise, loss, deps, n, s)));
  return EDB_OK;
}

extern "C" int encdiff_ddim_update(const float* x, const float* e_t, const float* noise, const float* coef4,
                                   float temperature, float* x_prev, float* pred_x0, long long n, void* stream) {
  EDB_REQUIRE(x && e_t && coef4 && x_prev && n > 0, "ddim_update: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(
      launch_ddim_update(x, e_t, noise, coef4, temperature, x_prev, pred_x0, n, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_ddim_select(const float* t_table, int* cursor, float* ts, int* idx, int batch, void* stream) {
  EDB_REQUIRE(t_table && cursor && ts && idx && batch > 0, "ddim_select: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(
      launch_ddim_select(t_table, cursor, ts, idx, batch, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_ddim_update_indexed(const float* x, const float* e_t, const float* noise_table,
                                           const float* coef_table, const int* idx, float temperature, float* x_prev,
                                           float* pred_x0, long long n, void* stream) {
  EDB_REQUIRE(x && e_t && coef_table && idx && x_prev && n > 0, "ddim_update_indexed: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_ddim_update_indexed(
      x, e_t, noise_table, coef_table, idx, temperature, x_prev, pred_x0, n, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_timestep_embedding(const float* t, void* out_h16, int batch, int dim, void* stream) {
  EDB_REQUIRE(t && out_h16 && batch > 0 && dim > 1, "timestep_embedding: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(
      launch_temb(t, static_cast<h16*>(out_h16), batch, dim, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_adamw_step(float* p, const float* g, float* m, float* v, float* ema, long long n, float lr,
                                  float beta1, float beta2, float eps, float weight_decay, int step, float ema_decay,
                                  void* stream) {
  EDB_REQUIRE(p && g && m && v && n > 0 && step >= 1, "adamw_step: bad argument");
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2 = 1.f - powf(beta2, static_cast<float>(step));
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_adamw(p, g, m, v, ema, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                                       ema_decay, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_adamw_step_dev(float* p, const float* g, float* m, float* v, float* ema, long long n,
                                      const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                                      float grad_scale, void* stream) {
  EDB_REQUIRE(p && g && m && v && hyper && n > 0, "adamw_step_dev: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_adamw_dev(p, g, m, v, ema, n, hyper, beta1, beta2, eps, weight_decay,
                                                           grad_scale, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}

extern "C" int encdiff_adamw_hyper_advance(int* step, const float* lr, float* hyper, float beta1, float beta2,
                                           float ema_decay_max, void* stream) {
  EDB_REQUIRE(step && lr && hyper, "adamw_hyper_advance: bad argument");
  EDB_CHECK_CUDA(static_cast<cudaError_t>(
      launch_adamw_hyper_advance(step, lr, hyper, beta1, beta2, ema_decay_max, static_cast<cudaStream_t>(stream))));
  return EDB_OK;
}
