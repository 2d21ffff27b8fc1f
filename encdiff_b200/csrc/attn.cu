// Attention kernels.  Self-attention: flash-style, K/V of a (sample, head) pair resident in shared memory, one
// query per thread, online softmax in the exp2 domain, scores never touch HBM.  Cross-attention: the T (~20)
// concept tokens' K/V are projected inside the kernel and kept in shared memory; softmax over T in registers.
// Reference: CrossAttention.forward, ldm/modules/attention.py:170-193.
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
    const float corr = exp2f(m - mn);
    float psum = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      s[k] = exp2f(s[k] - mn);
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
      const float ds = exp2f(s - li) * (dp - di) * scale;
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
      const float p = exp2f(s - Lg[j]);
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
  self_attn_fwd_kernel<D><<<grid, s.threads, smem, st>>>(qkv, out, lse, N, heads, s.TN, s.G, total, scale_log2);
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
  self_attn_bwd_kernel<D><<<grid, s.threads, smem, st>>>(qkv, out, dout, lse, dqkv, N, heads, s.TN, s.G, total,
                                                         rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

int launch_self_attn_fwd(const h16* qkv, h16* out, float* lse, int B, int N, int heads, int D, cudaStream_t s) {
  if (N < 4 || (N & 3)) return static_cast<int>(cudaErrorInvalidValue);
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
  switch (D) {
    case 8: return self_attn_bwd_t<8>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 16: return self_attn_bwd_t<16>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 32: return self_attn_bwd_t<32>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 64: return self_attn_bwd_t<64>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// cross-attention against T <= 32 context tokens: one CTA per sample
// ---------------------------------------------------------------------------------------------------------------
static constexpr int kMaxT = 32;

// K[j][c] = sum_e ctx[b][j][e] * wk[c][e]   (to_k / to_v without bias, attention.py:161-162)
__device__ __forceinline__ void project_ctx(const float* __restrict__ ctx_s, const float* __restrict__ wk,
                                            const float* __restrict__ wv, float* Ks, float* Vs, int T, int C, int Dc) {
  for (int idx = threadIdx.x; idx < T * C; idx += blockDim.x) {
    const int j = idx / C, c = idx - j * C;
    float k = 0.f, v = 0.f;
    for (int e = 0; e < Dc; ++e) {
      const float x = ctx_s[j * Dc + e];
      k += x * __ldg(wk + c * Dc + e);
      v += x * __ldg(wv + c * Dc + e);
    }
    Ks[idx] = k;
    Vs[idx] = v;
  }
}

template <int D>
__global__ void __launch_bounds__(256) cross_attn_fwd_kernel(const h16* __restrict__ q, const float* __restrict__ ctx,
                                                             const float* __restrict__ wk, const float* __restrict__ wv,
                                                             h16* __restrict__ out, int N, int heads, int T, int Dc,
                                                             float scale_log2) {
  extern __shared__ float sm[];
  const int C = heads * D;
  float* Ks = sm;
  float* Vs = Ks + T * C;
  float* ctx_s = Vs + T * C;
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
  __syncthreads();
  project_ctx(ctx_s, wk, wv, Ks, Vs, T, C, Dc);
  __syncthreads();
  constexpr int DC = D / 8;
  for (int e = threadIdx.x; e < N * heads; e += blockDim.x) {
    const int i = e / heads, h = e - i * heads;
    const long long row = static_cast<long long>(b) * N + i;
    float qv[D];
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) unpack8(*reinterpret_cast<const uint4*>(q + row * C + h * D + dc * 8), qv + dc * 8);
    float s[kMaxT];
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) a += qv[d] * Ks[j * C + h * D + d];
        s[j] = a * scale_log2;
        m = fmaxf(m, s[j]);
      }
    }
    float l = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxT; ++j) {
      if (j < T) {
        s[j] = exp2f(s[j] - m);
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
        for (int d = 0; d < D; ++d) acc[d] += p * Vs[j * C + h * D + d];
      }
    }
#pragma unroll
    for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(out + row * C + h * D + dc * 8) = pack8(acc + dc * 8);
  }
}

template <int D>
__global__ void __launch_bounds__(256) cross_attn_bwd_kernel(const h16* __restrict__ q, const float* __restrict__ ctx,
                                                             const float* __restrict__ wk, const float* __restrict__ wv,
                                                             const h16* __restrict__ dout, h16* __restrict__ dq,
                                                             float* __restrict__ dwk, float* __restrict__ dwv,
                                                             float* __restrict__ dctx,
                                                             const float* __restrict__ gscale, int N, int heads, int T,
                                                             int Dc, float scale) {
  const float ginv = gscale ? gscale[1] : 1.f;
  extern __shared__ float sm[];
  const int C = heads * D;
  float* Ks = sm;                   // [T][C]
  float* Vs = Ks + T * C;           // [T][C]
  float* ctx_s = Vs + T * C;        // [T][Dc]
  float* dctx_s = ctx_s + T * Dc;   // [T][Dc]
  float* Ps = dctx_s + T * Dc;      // [N][T]
  float* dSs = Ps + N * T;          // [N][T]
  float* dKh = dSs + N * T;         // [T][D]
  float* dVh = dKh + T * D;         // [T][D]
  const int b = blockIdx.x;
  const float scale_log2 = scale * kLog2e;
  for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) {
    ctx_s[i] = ctx[static_cast<long long>(b) * T * Dc + i];
    dctx_s[i] = 0.f;
  }
  __syncthreads();
  project_ctx(ctx_s, wk, wv, Ks, Vs, T, C, Dc);
  __syncthreads();
  constexpr int DC = D / 8;
  for (int h = 0; h < heads; ++h) {
    // phase 1: per query row -> p, dS, dq
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
      const long long row = static_cast<long long>(b) * N + i;
      float qv[D], dO[D];
#pragma unroll
      for (int dc = 0; dc < DC; ++dc) {
        unpack8(*reinterpret_cast<const uint4*>(q + row * C + h * D + dc * 8), qv + dc * 8);
        unpack8(*reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8), dO + dc * 8);
      }
      float s[kMaxT], dp[kMaxT];
      float m = -INFINITY;
#pragma unroll
      for (int j = 0; j < kMaxT; ++j) {
        if (j < T) {
          float a = 0.f, c = 0.f;
#pragma unroll
          for (int d = 0; d < D; ++d) {
            a += qv[d] * Ks[j * C + h * D + d];
            c += dO[d] * Vs[j * C + h * D + d];
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
          s[j] = exp2f(s[j] - m);
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
          Ps[i * T + j] = s[j];
          dSs[i * T + j] = ds;
#pragma unroll
          for (int d = 0; d < D; ++d) dqv[d] += ds * Ks[j * C + h * D + d];
        }
      }
#pragma unroll
      for (int dc = 0; dc < DC; ++dc) *reinterpret_cast<uint4*>(dq + row * C + h * D + dc * 8) = pack8(dqv + dc * 8);
    }
    __syncthreads();
    // phase 2: dK_h[j][d] = sum_i dS[i][j] q[i][d],  dV_h[j][d] = sum_i P[i][j] dO[i][d]
    for (int e = threadIdx.x; e < T * D; e += blockDim.x) {
      const int j = e / D, d = e - j * D;
      float ak = 0.f, av = 0.f;
      const h16* qc = q + static_cast<long long>(b) * N * C + h * D + d;
      const h16* oc = dout + static_cast<long long>(b) * N * C + h * D + d;
      for (int i = 0; i < N; ++i) {
        ak += dSs[i * T + j] * h2f(qc[static_cast<long long>(i) * C]);
        av += Ps[i * T + j] * h2f(oc[static_cast<long long>(i) * C]);
      }
      dKh[e] = ak;
      dVh[e] = av;
    }
    __syncthreads();
    // phase 3: weight and context gradients of this head
    for (int e = threadIdx.x; e < D * Dc; e += blockDim.x) {
      const int d = e / Dc, x = e - d * Dc;
      float ak = 0.f, av = 0.f;
      for (int j = 0; j < T; ++j) {
        ak += dKh[j * D + d] * ctx_s[j * Dc + x];
        av += dVh[j * D + d] * ctx_s[j * Dc + x];
      }
      atomicAdd(dwk + (h * D + d) * Dc + x, ak * ginv);
      atomicAdd(dwv + (h * D + d) * Dc + x, av * ginv);
    }
    for (int e = threadIdx.x; e < T * Dc; e += blockDim.x) {
      const int j = e / Dc, x = e - j * Dc;
      float a = 0.f;
      for (int d = 0; d < D; ++d)
        a += dKh[j * D + d] * __ldg(wk + (h * D + d) * Dc + x) + dVh[j * D + d] * __ldg(wv + (h * D + d) * Dc + x);
      dctx_s[e] += a;  // element e is owned by the same thread for every head
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < T * Dc; i += blockDim.x) dctx[static_cast<long long>(b) * T * Dc + i] += dctx_s[i] * ginv;
}

template <int D>
static int cross_fwd_t(const h16* q, const float* ctx, const float* wk, const float* wv, h16* out, int B, int N,
                       int heads, int T, int Dc, cudaStream_t st) {
  const int C = heads * D;
  const size_t smem = (2ull * T * C + T * Dc) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_fwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  cross_attn_fwd_kernel<D><<<B, 256, smem, st>>>(q, ctx, wk, wv, out, N, heads, T, Dc,
                                                 rsqrtf(static_cast<float>(D)) * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
static int cross_bwd_t(const h16* q, const float* ctx, const float* wk, const float* wv, const h16* dout, h16* dq,
                       float* dwk, float* dwv, float* dctx, const float* gscale, int B, int N, int heads, int T, int Dc,
                       cudaStream_t st) {
  const int C = heads * D;
  const size_t smem = (2ull * T * C + 2ull * T * Dc + 2ull * N * T + 2ull * T * D) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_bwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  cross_attn_bwd_kernel<D><<<B, 256, smem, st>>>(q, ctx, wk, wv, dout, dq, dwk, dwv, dctx, gscale, N, heads, T, Dc,
                                                 rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

int launch_cross_attn_fwd(const h16* q, const float* ctx, const float* wk, const float* wv, h16* out, int B, int N,
                          int heads, int D, int T, int Dc, cudaStream_t s) {
  if (T < 1 || T > kMaxT) return static_cast<int>(cudaErrorInvalidValue);
  switch (D) {
    case 8: return cross_fwd_t<8>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s);
    case 16: return cross_fwd_t<16>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s);
    case 32: return cross_fwd_t<32>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s);
    case 64: return cross_fwd_t<64>(q, ctx, wk, wv, out, B, N, heads, T, Dc, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

int launch_cross_attn_bwd(const h16* q, const float* ctx, const float* wk, const float* wv, const h16* dout, h16* dq,
                          float* dwk, float* dwv, float* dctx, const float* gscale, int B, int N, int heads, int D, int T,
                          int Dc, cudaStream_t s) {
  if (T < 1 || T > kMaxT) return static_cast<int>(cudaErrorInvalidValue);
  switch (D) {
    case 8: return cross_bwd_t<8>(q, ctx, wk, wv, dout, dq, dwk, dwv, dctx, gscale, B, N, heads, T, Dc, s);
    case 16: return cross_bwd_t<16>(q, ctx, wk, wv, dout, dq, dwk, dwv, dctx, gscale, B, N, heads, T, Dc, s);
    case 32: return cross_bwd_t<32>(q, ctx, wk, wv, dout, dq, dwk, dwv, dctx, gscale, B, N, heads, T, Dc, s);
    case 64: return cross_bwd_t<64>(q, ctx, wk, wv, dout, dq, dwk, dwv, dctx, gscale, B, N, heads, T, Dc, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

}  // namespace edb
