// Attention kernels.  Self-attention: flash-style, K/V of a (sample, head) pair resident in shared memory, one
// query per thread, online softmax in the exp2 domain, scores never touch HBM.  Cross-attention: the T (~20)
// concept tokens' K/V are projected inside the kernel and kept in shared memory; softmax over T in registers.
// Reference: CrossAttention.forward, ldm/modules/attention.py:170-193.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.cuh"

namespace edb {

static constexpr float kLog2e = 1.4426950408889634f;

// ---------------------------------------------------------------------------------------------------------------
// self-attention forward
// ---------------------------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256) self_attn_fwd_kernel(const h16* __restrict__ qkv, h16* __restrict__ out,
                                                            float* __restrict__ lse, int N, int heads, int TN, int G,
                                                            int total_pairs, float scale_log2) {
  pdl_enter();
  extern __shared__ float sm[];
  float* Ks = sm;
  float* Vs = sm + static_cast<size_t>(G) * N * D;
  const int C = heads * D;
  const int pair0 = blockIdx.x * G;
  constexpr int DC = D / 8;
  for (int idx = threadIdx.x; idx < G * N * DC; idx += blockDim.x) {
    const int g = idx / (N * DC);
    const int rem = idx - g * N * DC;
    const int j = rem / DC, dc = rem - j * DC;
    const int pair = pair0 + g;
    if (pair < total_pairs) {
      const int b = pair / heads, h = pair - b * heads;
      const h16* base = qkv + (static_cast<long long>(b) * N + j) * 3 * C + h * D + dc * 8;
      unpack8(*reinterpret_cast<const uint4*>(base + C), Ks + (static_cast<size_t>(g) * N + j) * D + dc * 8);
      unpack8(*reinterpret_cast<const uint4*>(base + 2 * C), Vs + (static_cast<size_t>(g) * N + j) * D + dc * 8);
    }
  }
  __syncthreads();
  const int g = threadIdx.x / TN;
  const int i = blockIdx.y * TN + threadIdx.x % TN;
  const int pair = pair0 + g;
  if (g >= G || pair >= total_pairs || i >= N) return;
  const int b = pair / heads, h = pair - b * heads;
  const long long row = static_cast<long long>(b) * N + i;
  float q[D], acc[D];
#pragma unroll
  for (int dc = 0; dc < DC; ++dc) unpack8(*reinterpret_cast<const uint4*>(qkv + row * 3 * C + h * D + dc * 8), q + dc * 8);
#pragma unroll
  for (int d = 0; d < D; ++d) {
    q[d] *= scale_log2;
    acc[d] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  const float* Kg = Ks + static_cast<size_t>(g) * N * D;
  const float* Vg = Vs + static_cast<size_t>(g) * N * D;
  for (int j0 = 0; j0 < N; j0 += 4) {
    float s[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4* kr = reinterpret_cast<const float4*>(Kg + (j0 + k) * D);
      float a = 0.f;
#pragma unroll
      for (int d4 = 0; d4 < D / 4; ++d4) {
        const float4 kv = kr[d4];
        a += q[4 * d4] * kv.x + q[4 * d4 + 1] * kv.y + q[4 * d4 + 2] * kv.z + q[4 * d4 + 3] * kv.w;
      }
      s[k] = a;
    }
    const float mn = fmaxf(m, fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])));
    const float corr = ex2_approx(m - mn);
    float psum = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      s[k] = ex2_approx(s[k] - mn);
      psum += s[k];
    }
    l = l * corr + psum;
#pragma unroll
    for (int d = 0; d < D; ++d) acc[d] *= corr;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4* vr = reinterpret_cast<const float4*>(Vg + (j0 + k) * D);
#pragma unroll
      for (int d4 = 0; d4 < D / 4; ++d4) {
        const float4 vv = vr[d4];
        acc[4 * d4] += s[k] * vv.x;
        acc[4 * d4 + 1] += s[k] * vv.y;
        acc[4 * d4 + 2] += s[k] * vv.z;
        acc[4 * d4 + 3] += s[k] * vv.w;
      }
    }
    m = mn;
  }
  const float inv = 1.f / l;
#pragma unroll
  for (int d = 0; d < D; ++d) acc[d] *= inv;
#pragma unroll
  for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(out + row * C + h * D + dc * 8) = pack8(acc + dc * 8);
  lse[static_cast<long long>(pair) * N + i] = m + log2f(l);
}

// ---------------------------------------------------------------------------------------------------------------
// self-attention backward: phase A (thread = query) -> dQ ; phase B (thread = key) -> dK, dV.  Probabilities are
// recomputed from the saved log-sum-exp.
// ---------------------------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256) self_attn_bwd_kernel(const h16* __restrict__ qkv, const h16* __restrict__ out,
                                                            const h16* __restrict__ dout, const float* __restrict__ lse,
                                                            h16* __restrict__ dqkv, int N, int heads, int TN, int G,
                                                            int total_pairs, float scale) {
  pdl_enter();
  extern __shared__ float sm[];
  const size_t sz = static_cast<size_t>(G) * N * D;
  float* Qs = sm;
  float* Ks = Qs + sz;
  float* Vs = Ks + sz;
  float* Os = Vs + sz;            // dO
  float* Ls = Os + sz;            // [G*N] lse (log2 domain)
  float* Ds = Ls + static_cast<size_t>(G) * N;  // [G*N] delta = rowsum(dO * O)
  const int C = heads * D;
  const int pair0 = blockIdx.x * G;
  constexpr int DC = D / 8;
  const float scale_log2 = scale * kLog2e;
  for (int idx = threadIdx.x; idx < G * N * DC; idx += blockDim.x) {
    const int g = idx / (N * DC);
    const int rem = idx - g * N * DC;
    const int j = rem / DC, dc = rem - j * DC;
    const int pair = pair0 + g;
    if (pair < total_pairs) {
      const int b = pair / heads, h = pair - b * heads;
      const long long row = static_cast<long long>(b) * N + j;
      const h16* base = qkv + row * 3 * C + h * D + dc * 8;
      const size_t o = (static_cast<size_t>(g) * N + j) * D + dc * 8;
      unpack8(*reinterpret_cast<const uint4*>(base), Qs + o);
      unpack8(*reinterpret_cast<const uint4*>(base + C), Ks + o);
      unpack8(*reinterpret_cast<const uint4*>(base + 2 * C), Vs + o);
      unpack8(*reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8), Os + o);
    }
  }
  for (int r = threadIdx.x; r < G * N; r += blockDim.x) {
    const int g = r / N, j = r - g * N;
    const int pair = pair0 + g;
    if (pair < total_pairs) {
      const int b = pair / heads, h = pair - b * heads;
      const long long row = static_cast<long long>(b) * N + j;
      float dl = 0.f;
#pragma unroll
      for (int dc = 0; dc < DC; ++dc) {
        float o[8], d[8];
        unpack8(*reinterpret_cast<const uint4*>(out + row * C + h * D + dc * 8), o);
        unpack8(*reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8), d);
#pragma unroll
        for (int k = 0; k < 8; ++k) dl += o[k] * d[k];
      }
      Ds[r] = dl;
      Ls[r] = lse[static_cast<long long>(pair) * N + j];
    }
  }
  __syncthreads();
  const int g = threadIdx.x / TN;
  const int i = blockIdx.y * TN + threadIdx.x % TN;
  const int pair = pair0 + g;
  if (g >= G || pair >= total_pairs || i >= N) return;
  const int b = pair / heads, h = pair - b * heads;
  const long long row = static_cast<long long>(b) * N + i;
  const float* Qg = Qs + static_cast<size_t>(g) * N * D;
  const float* Kg = Ks + static_cast<size_t>(g) * N * D;
  const float* Vg = Vs + static_cast<size_t>(g) * N * D;
  const float* Og = Os + static_cast<size_t>(g) * N * D;
  const float* Lg = Ls + static_cast<size_t>(g) * N;
  const float* Dg = Ds + static_cast<size_t>(g) * N;
  {  // phase A: this thread is query i
    float q[D], dq[D], dO[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      q[d] = Qg[i * D + d] * scale_log2;
      dO[d] = Og[i * D + d];
      dq[d] = 0.f;
    }
    const float li = Lg[i], di = Dg[i];
    for (int j = 0; j < N; ++j) {
      float s = 0.f, dp = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) {
        s += q[d] * Kg[j * D + d];
        dp += dO[d] * Vg[j * D + d];
      }
      const float ds = ex2_approx(s - li) * (dp - di) * scale;
#pragma unroll
      for (int d = 0; d < D; ++d) dq[d] += ds * Kg[j * D + d];
    }
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(dqkv + row * 3 * C + h * D + dc * 8) = pack8(dq + dc * 8);
  }
  {  // phase B: this thread is key i
    float k[D], v[D], dk[D], dv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      k[d] = Kg[i * D + d] * scale_log2;
      v[d] = Vg[i * D + d];
      dk[d] = dv[d] = 0.f;
    }
    for (int j = 0; j < N; ++j) {  // j runs over queries
      float s = 0.f, dp = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) {
        s += k[d] * Qg[j * D + d];
        dp += v[d] * Og[j * D + d];
      }
      const float p = ex2_approx(s - Lg[j]);
      const float ds = p * (dp - Dg[j]) * scale;
#pragma unroll
      for (int d = 0; d < D; ++d) {
        dv[d] += p * Og[j * D + d];
        dk[d] += ds * Qg[j * D + d];
      }
    }
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) {
      *reinterpret_cast<uint4*>(dqkv + row * 3 * C + C + h * D + dc * 8) = pack8(dk + dc * 8);
      *reinterpret_cast<uint4*>(dqkv + row * 3 * C + 2 * C + h * D + dc * 8) = pack8(dv + dc * 8);
    }
  }
}

struct SelfAttnShape {
  int TN, G, threads, qtiles;
};
static SelfAttnShape self_attn_shape(int N) {
  SelfAttnShape s;
  s.TN = N < 256 ? N : 256;
  s.threads = N < 128 ? 128 : s.TN;
  s.G = s.threads / s.TN;
  s.qtiles = (N + s.TN - 1) / s.TN;
  return s;
}

template <int D>
static int self_attn_fwd_t(const h16* qkv, h16* out, float* lse, int B, int N, int heads, cudaStream_t st) {
  const SelfAttnShape s = self_attn_shape(N);
  const int total = B * heads;
  const size_t smem = 2ull * s.G * N * D * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(self_attn_fwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  dim3 grid((total + s.G - 1) / s.G, s.qtiles);
  const float scale_log2 = rsqrtf(static_cast<float>(D)) * kLog2e;
  launch_k(self_attn_fwd_kernel<D>, grid, s.threads, smem, st, qkv, out, lse, N, heads, s.TN, s.G, total, scale_log2);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
static int self_attn_bwd_t(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                           int heads, cudaStream_t st) {
  const SelfAttnShape s = self_attn_shape(N);
  const int total = B * heads;
  const size_t smem = (4ull * s.G * N * D + 2ull * s.G * N) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(self_attn_bwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  dim3 grid((total + s.G - 1) / s.G, s.qtiles);
  launch_k(self_attn_bwd_kernel<D>, grid, s.threads, smem, st, qkv, out, dout, lse, dqkv, N, heads, s.TN, s.G, total,
                                                         rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

int launch_self_attn_fwd(const h16* qkv, h16* out, float* lse, int B, int N, int heads, int D, cudaStream_t s) {
  if (N < 4 || (N & 3)) return static_cast<int>(cudaErrorInvalidValue);
  static const bool cuda_core_only = getenv("ENCDIFF_ATTN_CUDA_CORE") != nullptr;
  if (!cuda_core_only) {
    const int rc = launch_self_attn_fwd_mma(qkv, out, lse, B, N, heads, D, s);
    if (rc != -1) return rc;
  }
  switch (D) {
    case 8: return self_attn_fwd_t<8>(qkv, out, lse, B, N, heads, s);
    case 16: return self_attn_fwd_t<16>(qkv, out, lse, B, N, heads, s);
    case 32: return self_attn_fwd_t<32>(qkv, out, lse, B, N, heads, s);
    case 64: return self_attn_fwd_t<64>(qkv, out, lse, B, N, heads, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

int launch_self_attn_bwd(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                         int heads, int D, cudaStream_t s) {
  if (N < 4 || (N & 3)) return static_cast<int>(cudaErrorInvalidValue);
  static const bool cuda_core_only = getenv("ENCDIFF_ATTN_CUDA_CORE") != nullptr;
  if (!cuda_core_only) {
    const int rc = launch_self_attn_bwd_mma(qkv, out, dout, lse, dqkv, B, N, heads, D, s);
    if (rc != -1) return rc;
  }
  switch (D) {
    case 8: return self_attn_bwd_t<8>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 16: return self_attn_bwd_t<16>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 32: return self_attn_bwd_t<32>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 64: return self_attn_bwd_t<64>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// cross-attention against T <= 32 concept tokens (attention.py:170-193 with context; openaimodel_enc.py:729).
// CTA = (sample, chunk of TQ <= 32 query tokens); item = (head, token) with the token fastest, so a warp works on one
// head and every K/V read from shared memory is a broadcast.  to_k / to_v (K = context_dim = 16) are applied inside
// the CTA: K, V [heads][T][D] fp32 stay resident in shared memory, softmax over T in registers.
// ---------------------------------------------------------------------------------------------------------------
static constexpr int kMaxT = 32;
static constexpr int kMaxDc = 16;

// Ks/Vs[(h*T + j)*D + d] = sum_e ctx_s[j][e] * w[h*D + d][e]      (to_k / to_v, no bias: attention.py:161-162)
template <int D>
__device__ __forceinline__ void project_ctx(const float* __restrict__ ctx_s, const float* __restrict__ wk,
                                            const float* __restrict__ wv, float* Ks, float* Vs, int T, int C, int Dc) {
  const int groups = blockDim.x >= C ? blockDim.x / C : 1;
  for (int cc = threadIdx.x; cc < C * groups; cc += blockDim.x) {
    const int c = cc % C, jg = cc / C;
    float wkr[kMaxDc], wvr[kMaxDc];
#pragma unroll
    for (int e = 0; e < kMaxDc; ++e) {
      wkr[e] = e < Dc ? __ldg(wk + c * Dc + e) : 0.f;
      wvr[e] = e < Dc ? __ldg(wv + c * Dc + e) : 0.f;
    }
    const int h = c / D, d = c - h * D;
    for (int j = jg; j < T; j += groups) {
      float k = 0.f, v = 0.f;
#pragma unroll
      for (int e = 0; e < kMaxDc; ++e) {
        const float x = e < Dc ? ctx_s[j * Dc + e] : 0.f;
        k += x * wkr[e];
        v += x * wvr[e];
      }
      Ks[(h * T + j) * D + d] = k;
      Vs[(h * T + j) * D + d] = v;
    }
  }
}

// Ks/Vs from the per-(layer, sample) projections computed once per forward (cross_kv_project_kernel):
// kv [T][2C] = K | V of this sample -> Ks/Vs[(h*T + j)*D + d]
template <int D>
__device__ __forceinline__ void load_kv(const float* __restrict__ kv, float* Ks, float* Vs, int T, int C) {
  const int c4n = 2 * C / 4;
  for (int i = threadIdx.x; i < T * c4n; i += blockDim.x) {
    const int j = i / c4n, c = (i - j * c4n) * 4;
    const float4 v = *reinterpret_cast<const float4*>(kv + static_cast<long long>(j) * 2 * C + c);
    float* dst = c < C ? Ks : Vs;
    const int cc = c < C ? c : c - C;
    const int h = cc / D, d = cc - h * D;
    *reinterpret_cast<float4*>(dst + (h * T + j) * D + d) = v;
  }
}

// K | V of every cross-attention layer for every sample in ONE launch (the context is the same for all 16 layers):
// CTA = (chunk of samples, layer); thread = output channel of [to_k ; to_v] with its weight row in registers.
__global__ void __launch_bounds__(256) cross_kv_project_kernel(const CrossKvEntry* __restrict__ tab,
                                                               const float* __restrict__ ctx, int B, int T, int Dc,
                                                               int chunk) {
  pdl_enter();
  __shared__ __align__(16) float ctx_s[kMaxT * kMaxDc];
  const CrossKvEntry en = tab[blockIdx.y];
  const int C = en.C;
  constexpr int kRows = 4;     // 2C <= 4 * 256
  float w[kRows][kMaxDc];
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int row = threadIdx.x + r * 256;
    const float* src = row < C ? en.wk + row * Dc : en.wv + (row - C) * Dc;
#pragma unroll
    for (int e = 0; e < kMaxDc; ++e) w[r][e] = (row < 2 * C && e < Dc) ? __ldg(src + e) : 0.f;
  }
  const int b_end = min(B, (static_cast<int>(blockIdx.x) + 1) * chunk);
  for (int b = blockIdx.x * chunk; b < b_end; ++b) {
    __syncthreads();
    for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
    __syncthreads();
    float* dst = en.kv + static_cast<long long>(b) * T * 2 * C;
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int row = threadIdx.x + r * 256;
      if (row < 2 * C) {
        if (Dc == kMaxDc) {
          for (int j = 0; j < T; ++j) {     // the context row as four broadcast 16-byte reads
            float a = 0.f;
#pragma unroll
            for (int e4 = 0; e4 < kMaxDc / 4; ++e4) {
              const float4 x = *reinterpret_cast<const float4*>(ctx_s + j * kMaxDc + 4 * e4);
              a += x.x * w[r][4 * e4] + x.y * w[r][4 * e4 + 1] + x.z * w[r][4 * e4 + 2] + x.w * w[r][4 * e4 + 3];
            }
            dst[j * 2 * C + row] = a;
          }
        } else {
          for (int j = 0; j < T; ++j) {
            float a = 0.f;
#pragma unroll
            for (int e = 0; e < kMaxDc; ++e)
              if (e < Dc) a += ctx_s[j * Dc + e] * w[r][e];
            dst[j * 2 * C + row] = a;
          }
        }
      }
    }
  }
}

int launch_cross_kv_project(const CrossKvEntry* tab_dev, int n_layers, const float* ctx, int B, int T, int Dc, cudaStream_t s) {
  if (n_layers <= 0) return 0;
  if (T > kMaxT || Dc > kMaxDc) return static_cast<int>(cudaErrorInvalidValue);
  const int chunk = 8;      // samples per CTA (4 and 2 measured: no gain)
  launch_k(cross_kv_project_kernel, dim3((B + chunk - 1) / chunk, n_layers), 256, 0, s, tab_dev, ctx, B, T, Dc, chunk);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
__global__ void __launch_bounds__(256) cross_attn_fwd_kernel(const h16* __restrict__ q, const float* __restrict__ ctx,
                                                             const float* __restrict__ wk, const float* __restrict__ wv,
                                                             h16* __restrict__ out, int N, int heads, int T, int Dc,
                                                             int TQ, float scale_log2, const float* __restrict__ kv) {
  pdl_enter();
  extern __shared__ float sm[];
  const int C = heads * D;
  float* Ks = sm;
  float* Vs = Ks + T * C;
  float* ctx_s = Vs + T * C;
  const int b = blockIdx.y, t0 = blockIdx.x * TQ;
  if (kv) {
    load_kv<D>(kv + static_cast<long long>(b) * T * 2 * C, Ks, Vs, T, C);
  } else {
    for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
    __syncthreads();
    project_ctx<D>(ctx_s, wk, wv, Ks, Vs, T, C, Dc);
  }
  __syncthreads();
  constexpr int DC = D / 8;
  for (int it = threadIdx.x; it < heads * TQ; it += blockDim.x) {
    const int h = it / TQ, tk = t0 + (it - h * TQ);
    if (tk >= N) continue;
    const long long row = static_cast<long long>(b) * N + tk;
    float qv[D];
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) unpack8(*reinterpret_cast<const uint4*>(q + row * C + h * D + dc * 8), qv + dc * 8);
    const float* Kh = Ks + h * T * D;
    const float* Vh = Vs + h * T * D;
    float s[kMaxT];
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        float a = 0.f;
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 kv = *reinterpret_cast<const float4*>(Kh + j * D + 4 * d4);
          a += qv[4 * d4] * kv.x + qv[4 * d4 + 1] * kv.y + qv[4 * d4 + 2] * kv.z + qv[4 * d4 + 3] * kv.w;
        }
        s[j] = a * scale_log2;
        m = fmaxf(m, s[j]);
      }
    }
    float l = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        s[j] = ex2_approx(s[j] - m);
        l += s[j];
      }
    }
    const float inv = 1.f / l;
    float acc[D];
#pragma unroll
    for (int d = 0; d < D; ++d) acc[d] = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        const float p = s[j] * inv;
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 vv = *reinterpret_cast<const float4*>(Vh + j * D + 4 * d4);
          acc[4 * d4] += p * vv.x;
          acc[4 * d4 + 1] += p * vv.y;
          acc[4 * d4 + 2] += p * vv.z;
          acc[4 * d4 + 3] += p * vv.w;
        }
      }
    }
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(out + row * C + h * D + dc * 8) = pack8(acc + dc * 8);
  }
}

// Backward, kernel 1 (same CTA / item mapping as the forward): per (head, token) recompute p, then
//   dS = p * (dO.V - delta) * scale ;  dq = dS K ;  and, through a shared-memory mini-GEMM over the chunk's tokens,
//   dK[j][c] += sum_t dS[t][j] q[t][c],  dV[j][c] += sum_t p[t][j] dO[t][c]   -> fp32 atomics into dkv[b][j][0..2C)
// (at most N / TQ <= 8 CTAs add into the same sample's rows).  All chunk-local arrays are token-fastest with a row
// pitch of TQP = TQ + 4 floats: stores by (head, token) items and float4 reads by (channel, key-group) threads are
// both bank-conflict free.
template <int D>
__global__ void __launch_bounds__(256) cross_attn_bwd_kernel(const h16* __restrict__ q, const float* __restrict__ ctx,
                                                             const float* __restrict__ wk, const float* __restrict__ wv,
                                                             const h16* __restrict__ dout, h16* __restrict__ dq,
                                                             float* __restrict__ dkv, int N, int heads, int T, int Dc,
                                                             int TQ, float scale, const float* __restrict__ kv) {
  pdl_enter();
  extern __shared__ float sm[];
  const int C = heads * D;
  const int TQP = TQ + 4;
  float* Ks = sm;                        // [heads][T][D]
  float* Vs = Ks + T * C;
  float* ctx_s = Vs + T * C;             // [T][Dc]
  float* Ps = ctx_s + T * Dc;            // [heads][T][TQP]
  float* dSs = Ps + heads * T * TQP;     // [heads][T][TQP]
  float* qs = dSs + heads * T * TQP;     // [C][TQP]
  float* dOs = qs + C * TQP;             // [C][TQP]
  const int b = blockIdx.y, t0 = blockIdx.x * TQ;
  const float scale_log2 = scale * kLog2e;
  if (kv) {
    load_kv<D>(kv + static_cast<long long>(b) * T * 2 * C, Ks, Vs, T, C);
  } else {
    for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
    __syncthreads();
    project_ctx<D>(ctx_s, wk, wv, Ks, Vs, T, C, Dc);
  }
  __syncthreads();
  constexpr int DC = D / 8;
  for (int it = threadIdx.x; it < heads * TQ; it += blockDim.x) {
    const int h = it / TQ, tl = it - h * TQ, tk = t0 + tl;
    float* Pr = Ps + h * T * TQP + tl;       // element j at Pr[j * TQP]
    float* dSr = dSs + h * T * TQP + tl;
    float* qr = qs + h * D * TQP + tl;       // element d at qr[d * TQP]
    float* dOr = dOs + h * D * TQP + tl;
    if (tk >= N) {   // tail chunk: contribute zeros to the mini-GEMM
      for (int j = 0; j < T; ++j) Pr[j * TQP] = dSr[j * TQP] = 0.f;
      for (int d = 0; d < D; ++d) qr[d * TQP] = dOr[d * TQP] = 0.f;
      continue;
    }
    const long long row = static_cast<long long>(b) * N + tk;
    float qv[D], dO[D];
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) {
      unpack8(*reinterpret_cast<const uint4*>(q + row * C + h * D + dc * 8), qv + dc * 8);
      unpack8(*reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8), dO + dc * 8);
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
      qr[d * TQP] = qv[d];
      dOr[d * TQP] = dO[d];
    }
    const float* Kh = Ks + h * T * D;
    const float* Vh = Vs + h * T * D;
    float s[kMaxT], dp[kMaxT];
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        float a = 0.f, c = 0.f;
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 kv = *reinterpret_cast<const float4*>(Kh + j * D + 4 * d4);
          const float4 vv = *reinterpret_cast<const float4*>(Vh + j * D + 4 * d4);
          a += qv[4 * d4] * kv.x + qv[4 * d4 + 1] * kv.y + qv[4 * d4 + 2] * kv.z + qv[4 * d4 + 3] * kv.w;
          c += dO[4 * d4] * vv.x + dO[4 * d4 + 1] * vv.y + dO[4 * d4 + 2] * vv.z + dO[4 * d4 + 3] * vv.w;
        }
        s[j] = a * scale_log2;
        dp[j] = c;
        m = fmaxf(m, s[j]);
      }
    }
    float l = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        s[j] = ex2_approx(s[j] - m);
        l += s[j];
      }
    }
    const float inv = 1.f / l;
    float delta = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        s[j] *= inv;
        delta += s[j] * dp[j];
      }
    }
    float dqv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) dqv[d] = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        const float ds = s[j] * (dp[j] - delta) * scale;
        Pr[j * TQP] = s[j];
        dSr[j * TQP] = ds;
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 kv = *reinterpret_cast<const float4*>(Kh + j * D + 4 * d4);
          dqv[4 * d4] += ds * kv.x;
          dqv[4 * d4 + 1] += ds * kv.y;
          dqv[4 * d4 + 2] += ds * kv.z;
          dqv[4 * d4 + 3] += ds * kv.w;
        }
      }
    }
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(dq + row * C + h * D + dc * 8) = pack8(dqv + dc * 8);
  }
  __syncthreads();
  // mini-GEMM: thread = (channel c, key group jg); q / dO columns are read once per 4 tokens and reused for its keys
  float* dst = dkv + static_cast<long long>(b) * T * 2 * C;
  const int groups = blockDim.x >= C ? blockDim.x / C : 1;
  constexpr int kJ = 8;                   // keys per pass held in registers
  for (int cc = threadIdx.x; cc < C * groups; cc += blockDim.x) {
    const int c = cc % C, jg = cc / C, h = c / D;
    const float* qc = qs + c * TQP;
    const float* oc = dOs + c * TQP;
    const float* Ph = Ps + h * T * TQP;
    const float* dSh = dSs + h * T * TQP;
    const int jper = (T + groups - 1) / groups;
    const int jbeg = jg * jper, jend = min(T, jbeg + jper);
    for (int j0 = jbeg; j0 < jend; j0 += kJ) {
      float ak[kJ], av[kJ];
#pragma unroll
      for (int k = 0; k < kJ; ++k) ak[k] = av[k] = 0.f;
      for (int t = 0; t < TQ; t += 4) {
        const float4 q4 = *reinterpret_cast<const float4*>(qc + t);
        const float4 o4 = *reinterpret_cast<const float4*>(oc + t);
#pragma unroll
        for (int k = 0; k < kJ; ++k) {
          if (j0 + k < jend) {
            const float4 d4 = *reinterpret_cast<const float4*>(dSh + (j0 + k) * TQP + t);
            const float4 p4 = *reinterpret_cast<const float4*>(Ph + (j0 + k) * TQP + t);
            ak[k] += d4.x * q4.x + d4.y * q4.y + d4.z * q4.z + d4.w * q4.w;
            av[k] += p4.x * o4.x + p4.y * o4.y + p4.z * o4.z + p4.w * o4.w;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kJ; ++k) {
        if (j0 + k < jend) {
          atomicAdd(dst + (j0 + k) * 2 * C + c, ak[k]);
          atomicAdd(dst + (j0 + k) * 2 * C + C + c, av[k]);
        }
      }
    }
  }
}

// Backward, kernel 2: CTA handles samples blockIdx.x, blockIdx.x + gridDim.x, ...
//   dctx[b][j][e] += ginv * sum_c (dK[b][j][c] wk[c][e] + dV[b][j][c] wv[c][e])           (single writer per sample)
//   dwk[c][e]     += ginv * sum_b sum_j dK[b][j][c] ctx[b][j][e]   (registers over the CTA's samples, then atomics)
__global__ void __launch_bounds__(256) cross_attn_bwd_proj_kernel(const float* __restrict__ dkv,
                                                                  const float* __restrict__ ctx,
                                                                  const float* __restrict__ wk,
                                                                  const float* __restrict__ wv, float* __restrict__ dwk,
                                                                  float* __restrict__ dwv, float* __restrict__ dctx,
                                                                  const float* __restrict__ gscale, int B, int C, int T,
                                                                  int Dc) {
  pdl_enter();
  extern __shared__ float sm[];
  float* w_s = sm;                    // [2C][Dc]: wk rows then wv rows
  float* g_s = w_s + 2 * C * Dc;      // [T][2C] dK | dV of the current sample
  float* ctx_s = g_s + T * 2 * C;     // [T][Dc]
  const float ginv = gscale ? gscale[1] : 1.f;
  for (int i = threadIdx.x; i < C * Dc; i += blockDim.x) {
    w_s[i] = wk[i];
    w_s[C * Dc + i] = wv[i];
  }
  constexpr int kRows = 4;            // rows (of 2C) owned per thread for the weight gradient: 2C <= 4 * 256
  float acc[kRows][kMaxDc];
#pragma unroll
  for (int r = 0; r < kRows; ++r)
#pragma unroll
    for (int e = 0; e < kMaxDc; ++e) acc[r][e] = 0.f;
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < T * 2 * C; i += blockDim.x) g_s[i] = dkv[static_cast<long long>(b) * T * 2 * C + i];
    for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
    __syncthreads();
    for (int o = threadIdx.x; o < T * Dc; o += blockDim.x) {
      const int j = o / Dc, e = o - j * Dc;
      float a = 0.f;
      for (int c = 0; c < 2 * C; ++c) a += g_s[j * 2 * C + c] * w_s[c * Dc + e];
      dctx[static_cast<long long>(b) * T * Dc + o] += a * ginv;
    }
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int row = threadIdx.x + r * 256;
      if (row < 2 * C) {
        for (int j = 0; j < T; ++j) {
          const float g = g_s[j * 2 * C + row];
#pragma unroll
          for (int e = 0; e < kMaxDc; ++e)
            if (e < Dc) acc[r][e] += g * ctx_s[j * Dc + e];
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int row = threadIdx.x + r * 256;
    if (row < 2 * C) {
      float* dst = row < C ? dwk + row * Dc : dwv + (row - C) * Dc;
#pragma unroll
      for (int e = 0; e < kMaxDc; ++e)
        if (e < Dc) atomicAdd(dst + e, acc[r][e] * ginv);
    }
  }
}

// All layers' projection gradients in ONE launch (the backward keeps every layer's dK | dV).  With R = B*T rows the
// buffers are plain matrices: dkv [R, 2C], ctx [R, Dc], dctx [R, Dc], and the two products are small GEMMs
//   dctx[R, Dc]  += dkv[R, 2C] . [wk ; wv][2C, Dc]          dW[2C, Dc] += dkv^T[2C, R] . ctx[R, Dc]
// run on the TF32 warp MMA (m16n8k8, fp32 accumulate; fp32 operands rounded to tf32).  CTA = (160-row chunk, layer);
// A fragments come straight from global memory (every element is used by exactly two MMAs), the small B operands
// ([wk ; wv] and the chunk's context rows) sit in shared memory with a 24-float pitch (conflict-free fragment reads).
// dctx and dW are accumulated with fp32 atomics (16 layers / 32 chunks add into the same addresses).
static constexpr int kPRows = 160;      // rows of dkv per CTA (10 m-tiles)
static constexpr int kPPitch = 24;

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(256) cross_attn_bwd_proj_group_kernel(const CrossProjEntry* __restrict__ tab,
                                                                        const float* __restrict__ ctx,
                                                                        float* __restrict__ dctx,
                                                                        const float* __restrict__ gscale, int R) {
  pdl_enter();
  extern __shared__ float sm[];
  const CrossProjEntry en = tab[blockIdx.y];
  const int C2 = 2 * en.C;
  float* w_s = sm;                          // [2C][kPPitch]: wk rows then wv rows, 16 context dims each
  float* c_s = w_s + C2 * kPPitch;          // [kPRows][kPPitch] context rows of this chunk
  const float ginv = gscale ? gscale[1] : 1.f;
  const int r0 = blockIdx.x * kPRows;
  for (int i = threadIdx.x; i < C2 * 4; i += blockDim.x) {
    const int row = i >> 2, q4 = (i & 3) * 4;
    const float* src = row < en.C ? en.wk + row * 16 : en.wv + (row - en.C) * 16;
    *reinterpret_cast<float4*>(w_s + row * kPPitch + q4) = *reinterpret_cast<const float4*>(src + q4);
  }
  for (int i = threadIdx.x; i < kPRows * 4; i += blockDim.x) {
    const int row = i >> 2, q4 = (i & 3) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r0 + row < R) v = *reinterpret_cast<const float4*>(ctx + static_cast<long long>(r0 + row) * 16 + q4);
    *reinterpret_cast<float4*>(c_s + row * kPPitch + q4) = v;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const float* dk = en.dkv;
  // ---- dctx[r0 + m][e] += sum_c dkv[r0 + m][c] * W[c][e] ----
  for (int mt = warp; mt < kPRows / 16; mt += 8) {
    const int ra = r0 + mt * 16 + g, rb = ra + 8;
    const float* pa = dk + static_cast<long long>(ra) * C2;
    const float* pb = dk + static_cast<long long>(rb) * C2;
    const bool oka = ra < R, okb = rb < R;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 4
    for (int k0 = 0; k0 < C2; k0 += 8) {
      uint32_t a[4];
      a[0] = to_tf32(oka ? pa[k0 + t] : 0.f);
      a[1] = to_tf32(okb ? pb[k0 + t] : 0.f);
      a[2] = to_tf32(oka ? pa[k0 + t + 4] : 0.f);
      a[3] = to_tf32(okb ? pb[k0 + t + 4] : 0.f);
      const float* wb = w_s + (k0 + t) * kPPitch + g;
      mma_tf32(acc[0], a, to_tf32(wb[0]), to_tf32(wb[4 * kPPitch]));
      mma_tf32(acc[1], a, to_tf32(wb[8]), to_tf32(wb[4 * kPPitch + 8]));
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      if (oka) {
        atomicAdd(dctx + static_cast<long long>(ra) * 16 + nt * 8 + 2 * t, acc[nt][0] * ginv);
        atomicAdd(dctx + static_cast<long long>(ra) * 16 + nt * 8 + 2 * t + 1, acc[nt][1] * ginv);
      }
      if (okb) {
        atomicAdd(dctx + static_cast<long long>(rb) * 16 + nt * 8 + 2 * t, acc[nt][2] * ginv);
        atomicAdd(dctx + static_cast<long long>(rb) * 16 + nt * 8 + 2 * t + 1, acc[nt][3] * ginv);
      }
    }
  }
  // ---- dW[c][e] += sum_m dkv[r0 + m][c] * ctx[r0 + m][e] ----
  for (int mt = warp; mt < C2 / 16; mt += 8) {
    const int ca = mt * 16 + g, cb = ca + 8;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 4
    for (int k0 = 0; k0 < kPRows; k0 += 8) {
      const int ka = r0 + k0 + t, kb = ka + 4;
      const float* p0 = dk + static_cast<long long>(ka) * C2;
      const float* p1 = dk + static_cast<long long>(kb) * C2;
      uint32_t a[4];
      a[0] = to_tf32(ka < R ? p0[ca] : 0.f);
      a[1] = to_tf32(ka < R ? p0[cb] : 0.f);
      a[2] = to_tf32(kb < R ? p1[ca] : 0.f);
      a[3] = to_tf32(kb < R ? p1[cb] : 0.f);
      const float* cb_s = c_s + (k0 + t) * kPPitch + g;
      mma_tf32(acc[0], a, to_tf32(cb_s[0]), to_tf32(cb_s[4 * kPPitch]));
      mma_tf32(acc[1], a, to_tf32(cb_s[8]), to_tf32(cb_s[4 * kPPitch + 8]));
    }
    float* da = ca < en.C ? en.dwk + ca * 16 : en.dwv + (ca - en.C) * 16;
    float* db = cb < en.C ? en.dwk + cb * 16 : en.dwv + (cb - en.C) * 16;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      atomicAdd(da + nt * 8 + 2 * t, acc[nt][0] * ginv);
      atomicAdd(da + nt * 8 + 2 * t + 1, acc[nt][1] * ginv);
      atomicAdd(db + nt * 8 + 2 * t, acc[nt][2] * ginv);
      atomicAdd(db + nt * 8 + 2 * t + 1, acc[nt][3] * ginv);
    }
  }
}

int launch_cross_attn_bwd_proj_group(const CrossProjEntry* tab_dev, int n_layers, int max_C, const float* ctx, float* dctx,
                                     const float* gscale, int B, int T, int Dc, cudaStream_t s) {
  if (n_layers <= 0) return 0;
  if (Dc != 16 || (max_C % 8) != 0) return static_cast<int>(cudaErrorInvalidValue);   // TF32 tiles: 16 context dims
  const size_t smem = (2ull * max_C + kPRows) * kPPitch * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_bwd_proj_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int R = B * T;
  launch_k(cross_attn_bwd_proj_group_kernel, dim3((R + kPRows - 1) / kPRows, n_layers), 256, smem, s, tab_dev, ctx, dctx, gscale,
           R);
  return static_cast<int>(cudaGetLastError());
}

static int cross_tq(int N) { return N < 32 ? N : 32; }

template <int D>
static int cross_fwd_t(const h16* q, const float* ctx, const float* wk, const float* wv, h16* out, int B, int N,
                       int heads, int T, int Dc, cudaStream_t st, const float* kv) {
  const int C = heads * D, TQ = cross_tq(N);
  const size_t smem = (2ull * T * C + T * Dc) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_fwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int threads = heads * TQ >= 256 ? 256 : ((heads * TQ + 31) / 32) * 32;
  dim3 grid((N + TQ - 1) / TQ, B);
  launch_k(cross_attn_fwd_kernel<D>, grid, threads < 64 ? 64 : threads, smem, st, q, ctx, wk, wv, out, N, heads, T, Dc, TQ,
           rsqrtf(static_cast<float>(D)) * kLog2e, kv);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
static int cross_bwd_t(const h16* q, const float* ctx, const float* wk, const float* wv, const h16* dout, h16* dq,
                       float* dkv, int B, int N, int heads, int T, int Dc, cudaStream_t st, const float* kv) {
  const int C = heads * D, TQ = cross_tq(N);
  const size_t smem = (2ull * T * C + T * Dc + 2ull * heads * T * (TQ + 4) + 2ull * C * (TQ + 4)) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_bwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int threads = heads * TQ >= 256 ? 256 : ((heads * TQ + 31) / 32) * 32;
  dim3 grid((N + TQ - 1) / TQ, B);
  launch_k(cross_attn_bwd_kernel<D>, grid, threads < 64 ? 64 : threads, smem, st, q, ctx, wk, wv, dout, dq, dkv, N, heads, T, Dc,
           TQ, rsqrtf(static_cast<float>(D)), kv);
  return static_cast<int>(cudaGetLastError());
}

int launch_cross_attn_fwd(const h16* q, const float* ctx, const float* wk, const float* wv, h16* out, int B, int N,
                          int heads, int D, int T, int Dc, cudaStream_t s, const float* kv) {
  if (T < 1 || T > kMaxT || Dc < 1 || Dc > kMaxDc) return static_cast<int>(cudaErrorInvalidValue);
  static const bool cuda_core_only = getenv("ENCDIFF_ATTN_CUDA_CORE") != nullptr;
  static const bool staged_only = getenv("ENCDIFF_ATTN_STAGED") != nullptr;   // tests: keep the attn_mma.cu kernels covered
  if (kv && !cuda_core_only) {
    int rc = staged_only ? -1 : launch_cross_attn_fwd_reg(q, kv, out, B, N, heads, D, T, s);
    if (rc == -1) rc = launch_cross_attn_fwd_mma(q, kv, out, B, N, heads, D, T, s);
    if (rc != -1) return rc;
  }
  switch (D) {
    case 8: return cross_fwd_t<8>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s, kv);
    case 16: return cross_fwd_t<16>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s, kv);
    case 32: return cross_fwd_t<32>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s, kv);
    case 64: return cross_fwd_t<64>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s, kv);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

size_t cross_attn_bwd_scratch_bytes(int B, int heads, int D, int T) {
  return static_cast<size_t>(B) * T * 2 * heads * D * sizeof(float);
}

int launch_cross_attn_bwd(const h16* q, const float* ctx, const float* wk, const float* wv, const h16* dout, h16* dq,
                          float* dwk, float* dwv, float* dctx, const float* gscale, float* dkv_scratch, int B, int N,
                          int heads, int D, int T, int Dc, cudaStream_t s, bool defer_proj, const float* kv) {
  if (T < 1 || T > kMaxT || Dc < 1 || Dc > kMaxDc || 2 * heads * D > 4 * 256) return static_cast<int>(cudaErrorInvalidValue);
  const int C = heads * D;
  static const bool cuda_core_only = getenv("ENCDIFF_ATTN_CUDA_CORE") != nullptr;
  int rc = -1;
  static const bool staged_only = getenv("ENCDIFF_ATTN_STAGED") != nullptr;
  if (kv && !cuda_core_only && !staged_only) rc = launch_cross_attn_bwd_reg(q, kv, dout, dq, dkv_scratch, B, N, heads, D, T, s);
  if (kv && !cuda_core_only && rc == -1) rc = launch_cross_attn_bwd_mma(q, kv, dout, dq, dkv_scratch, B, N, heads, D, T, s);
  cudaError_t e = cudaSuccess;
  if (rc == -1) {
  e = cudaMemsetAsync(dkv_scratch, 0, cross_attn_bwd_scratch_bytes(B, heads, D, T), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  switch (D) {
    case 8: rc = cross_bwd_t<8>(q, ctx, wk, wv, dout, dq, dkv_scratch, B, N, heads, T, Dc, s, kv); break;
    case 16: rc = cross_bwd_t<16>(q, ctx, wk, wv, dout, dq, dkv_scratch, B, N, heads, T, Dc, s, kv); break;
    case 32: rc = cross_bwd_t<32>(q, ctx, wk, wv, dout, dq, dkv_scratch, B, N, heads, T, Dc, s, kv); break;
    case 64: rc = cross_bwd_t<64>(q, ctx, wk, wv, dout, dq, dkv_scratch, B, N, heads, T, Dc, s, kv); break;
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
  }
  if (rc || defer_proj) return rc;
  const size_t smem = (2ull * C * Dc + 2ull * T * C + T * Dc) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    e = cudaFuncSetAttribute(cross_attn_bwd_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int grid = B < 148 ? B : 148;
  launch_k(cross_attn_bwd_proj_kernel, grid, 256, smem, s, dkv_scratch, ctx, wk, wv, dwk, dwv, dctx, gscale, B, C, T, Dc);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
