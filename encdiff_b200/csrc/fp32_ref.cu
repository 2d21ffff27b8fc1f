// fp32 validation forward of the UNet (north star: "eps-prediction max relative error <= 1e-4 in an fp32 validation
// build").  Same architecture walk as unet_forward (unet_run.cu; UNetModel.forward, openaimodel_enc.py:712-748), but
// every tensor is fp32, every contraction is a plain fp32 FMA loop on the CUDA cores and the kernels read the bound fp32
// master parameters in the reference's own layouts -- no 16-bit operand arena, no tensor cores (TF32 carries the same
// 10-bit significand as fp16 and cannot meet 1e-4; the reference's own GPU precision, TF32 via main_val.py:524-525, is
// ~1e-3 from its fp32 CPU result).  This is a checker that runs ON the GPU through the C ABI: slow by design
// (~1 TFLOP/s), it validates the 16-bit product path at shapes the CPU oracle cannot reach in seconds (B = 256) and pins
// the library's layer walk / parameter wiring to the oracle at 1e-4.  Forward only; nothing in the product path calls it.
#include <math.h>

#include <vector>

#include "common.cuh"
#include "host.cuh"
#include "unet.cuh"

namespace edb {
namespace {

// ---- out[M, N] = A * Wt^T (+ bias) (+ res);  A plain [M, lda] or the implicit im2col of a 3x3 / pad 1 convolution over
// NHWC x (k = ci * 9 + ky * 3 + kx, which is the reference's [Cout, Cin, 3, 3] weight flattened row-major) -------------
struct F32Gemm {
  const float* A;
  int lda;
  int H, W, Cin;
  const float* Wt;
  const float* bias;
  const float* res;
  int ldres;
  float* out;
  int ldo;
  int M, N, K;
};

template <bool CONV>
__global__ void __launch_bounds__(256) f32_gemm_kernel(const F32Gemm p) {
  __shared__ float As[16][65], Bs[16][65];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int HW = p.H * p.W;
  for (int k0 = 0; k0 < p.K; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256;
      const int mm = idx >> 4, kk = idx & 15;
      const int m = m0 + mm, k = k0 + kk;
      float v = 0.f;
      if (m < p.M && k < p.K) {
        if (CONV) {
          const int ci = k / 9, tap = k - ci * 9, ky = tap / 3, kx = tap - ky * 3;
          const int b = m / HW, pix = m - b * HW;
          const int y = pix / p.W + ky - 1, x = pix % p.W + kx - 1;
          if (y >= 0 && y < p.H && x >= 0 && x < p.W) v = p.A[(static_cast<long long>(b) * HW + y * p.W + x) * p.Cin + ci];
        } else {
          v = p.A[static_cast<long long>(m) * p.lda + k];
        }
      }
      As[kk][mm] = v;
      const int n = n0 + mm;
      Bs[kk][mm] = (n < p.N && k < p.K) ? p.Wt[static_cast<long long>(n) * p.K + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i] = As[kk][ty * 4 + i];
        b[i] = Bs[kk][tx * 4 + i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (p.bias) v += p.bias[n];
      if (p.res) v += p.res[static_cast<long long>(m) * p.ldres + n];
      p.out[static_cast<long long>(m) * p.ldo + n] = v;
    }
  }
}

__device__ __forceinline__ float block_sum(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  for (int w = 0; w < (blockDim.x >> 5); ++w) r += sh[w];
  return r;
}

// GroupNorm(32) over the channel concatenation of x0 | x1 (NHWC), optional FiLM (1 + scale) * y + shift, optional SiLU
// (util.py:227-244, openaimodel_enc.py:264-271).  grid (32, B)
__global__ void __launch_bounds__(256) f32_gn_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1,
                                                     int HW, const float* __restrict__ gamma, const float* __restrict__ beta,
                                                     float eps, const float* __restrict__ film, int ldfilm, int silu,
                                                     float* __restrict__ out) {
  __shared__ float sh[8];
  const int C = c0 + c1, cpg = C / 32, g = blockIdx.x, b = blockIdx.y;
  const int n = HW * cpg;
  auto at = [&](int idx) {
    const int pix = idx / cpg, c = g * cpg + (idx - pix * cpg);
    return c < c0 ? x0[(static_cast<long long>(b) * HW + pix) * c0 + c] : x1[(static_cast<long long>(b) * HW + pix) * c1 + (c - c0)];
  };
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += at(i);
  const float mean = block_sum(s, sh) / n;
  float q = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float d = at(i) - mean;
    q += d * d;
  }
  const float rstd = 1.0f / sqrtf(block_sum(q, sh) / n + eps);
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int pix = i / cpg, c = g * cpg + (i - pix * cpg);
    float y = (at(i) - mean) * rstd * gamma[c] + beta[c];
    if (film) y = y * (1.f + film[static_cast<long long>(b) * ldfilm + c]) + film[static_cast<long long>(b) * ldfilm + C + c];
    if (silu) y = y / (1.f + expf(-y));
    out[(static_cast<long long>(b) * HW + pix) * C + c] = y;
  }
}

// channel concatenation + optional 2x nearest upsample / 2x2 average pool (openaimodel_enc.py:109-119,158-160), NHWC.
// H, W are the INPUT size.  mode 0 none, 1 up, 2 down
__global__ void f32_resample_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1, int B, int H,
                                    int W, int mode, float* __restrict__ out) {
  const int C = c0 + c1;
  const int Ho = mode == 1 ? 2 * H : (mode == 2 ? H / 2 : H), Wo = mode == 1 ? 2 * W : (mode == 2 ? W / 2 : W);
  const long long total = static_cast<long long>(B) * Ho * Wo * C;
  auto at = [&](int b, int y, int x, int c) {
    return c < c0 ? x0[((static_cast<long long>(b) * H + y) * W + x) * c0 + c] : x1[((static_cast<long long>(b) * H + y) * W + x) * c1 + (c - c0)];
  };
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    long long r = i / C;
    const int xo = static_cast<int>(r % Wo);
    r /= Wo;
    const int yo = static_cast<int>(r % Ho), b = static_cast<int>(r / Ho);
    float v;
    if (mode == 1) v = at(b, yo >> 1, xo >> 1, c);
    else if (mode == 2) v = 0.25f * (at(b, 2 * yo, 2 * xo, c) + at(b, 2 * yo, 2 * xo + 1, c) + at(b, 2 * yo + 1, 2 * xo, c) + at(b, 2 * yo + 1, 2 * xo + 1, c));
    else v = at(b, yo, xo, c);
    out[i] = v;
  }
}

// LayerNorm over the last axis, one warp per row (attention.py:203-205)
__global__ void f32_ln_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                              float eps, float* __restrict__ out, long long M, int C) {
  const long long row = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * C;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += xr[c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / C;
  float q = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = xr[c] - mean;
    q += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = 1.0f / sqrtf(q / C + eps);
  for (int c = lane; c < C; c += 32) out[row * C + c] = (xr[c] - mean) * rstd * gamma[c] + beta[c];
}

// softmax(q k^T * D^-1/2) v per (sample, head, query): one thread per query, two passes over the keys
// (attention.py:170-193).  D <= 64
__global__ void f32_attn_kernel(const float* __restrict__ Q, int ldq, const float* __restrict__ K, int ldk,
                                const float* __restrict__ V, int ldv, float* __restrict__ O, int ldo, int N, int Nk, int D,
                                float scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, h = blockIdx.y, b = blockIdx.z;
  if (i >= N) return;
  float q[64], o[64];
  const float* qp = Q + (static_cast<long long>(b) * N + i) * ldq + h * D;
  for (int d = 0; d < D; ++d) q[d] = qp[d] * scale;
  const float* kb = K + static_cast<long long>(b) * Nk * ldk + h * D;
  const float* vb = V + static_cast<long long>(b) * Nk * ldv + h * D;
  float m = -INFINITY;
  for (int j = 0; j < Nk; ++j) {
    float s = 0.f;
    for (int d = 0; d < D; ++d) s = fmaf(q[d], kb[static_cast<long long>(j) * ldk + d], s);
    m = fmaxf(m, s);
  }
  float l = 0.f;
  for (int d = 0; d < D; ++d) o[d] = 0.f;
  for (int j = 0; j < Nk; ++j) {
    float s = 0.f;
    for (int d = 0; d < D; ++d) s = fmaf(q[d], kb[static_cast<long long>(j) * ldk + d], s);
    const float pj = expf(s - m);
    l += pj;
    for (int d = 0; d < D; ++d) o[d] = fmaf(pj, vb[static_cast<long long>(j) * ldv + d], o[d]);
  }
  float* op = O + (static_cast<long long>(b) * N + i) * ldo + h * D;
  for (int d = 0; d < D; ++d) op[d] = o[d] / l;
}

// GEGLU: out[m, c] = g[m, c] * gelu_erf(g[m, F + c]) (attention.py:37-45)
__global__ void f32_geglu_kernel(const float* __restrict__ g, float* __restrict__ out, long long M, int F) {
  const long long total = M * F;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long m = i / F;
    const int c = static_cast<int>(i - m * F);
    const float a = g[m * 2 * F + c], x = g[m * 2 * F + F + c];
    out[i] = a * (0.5f * x * (1.f + erff(x * 0.70710678118654752440f)));
  }
}

__global__ void f32_silu_kernel(const float* __restrict__ x, float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = x[i] / (1.f + expf(-x[i]));
}

// timestep_embedding (util.py:179-199): [cos | sin], max_period 1e4
__global__ void f32_temb_kernel(const float* __restrict__ t, float* __restrict__ out, int B, int dim) {
  const int half = dim / 2;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < B * half; idx += gridDim.x * blockDim.x) {
    const int b = idx / half, i = idx - b * half;
    const float freq = expf(-9.210340371976184f * static_cast<float>(i) / static_cast<float>(half));
    const float a = t[b] * freq;
    out[b * dim + i] = cosf(a);
    out[b * dim + half + i] = sinf(a);
    if ((dim & 1) && i == 0) out[b * dim + dim - 1] = 0.f;
  }
}

// NCHW <-> NHWC
__global__ void f32_layout_kernel(const float* __restrict__ in, float* __restrict__ out, int B, int C, int HW, int to_nhwc) {
  const long long total = static_cast<long long>(B) * C * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    if (to_nhwc) {      // i indexes out [B, HW, C]
      const int c = static_cast<int>(i % C);
      const long long r = i / C;
      const int pix = static_cast<int>(r % HW), b = static_cast<int>(r / HW);
      out[i] = in[(static_cast<long long>(b) * C + c) * HW + pix];
    } else {            // i indexes out [B, C, HW]
      const int pix = static_cast<int>(i % HW);
      const long long r = i / HW;
      const int c = static_cast<int>(r % C), b = static_cast<int>(r / C);
      out[i] = in[(static_cast<long long>(b) * HW + pix) * C + c];
    }
  }
}

inline int blocks_for(long long n) {
  const long long b = (n + 255) / 256;
  return static_cast<int>(b < 1 ? 1 : (b > 148 * 32 ? 148 * 32 : b));
}

// The layer walk.  dry = true only sizes the workspace (floats).
struct Fp32Exec {
  UNet* u;
  int B;
  cudaStream_t s;
  float* ws;
  size_t cap;
  bool dry;
  size_t off = 0, peak = 0;
  int err = 0;

  float* take(size_t n) {
    off = (off + 63) & ~size_t(63);
    float* p = (dry ? reinterpret_cast<float*>(uintptr_t(1) << 30) : ws) + off;   // dry: never dereferenced
    off += n;
    peak = off > peak ? off : peak;
    if (!dry && off > cap && !err) {
      err = EDB_ERR_INVALID;
      set_last_error("unet_forward_fp32: workspace too small");
    }
    return p;
  }
  const float* P(int idx) const { return u->pptr[idx]; }
  bool skip() const { return dry || err; }
  void check() {
    if (!err && cudaGetLastError() != cudaSuccess) {
      err = EDB_ERR_CUDA;
      set_last_error("unet_forward_fp32: launch failed");
    }
  }

  void linear(const float* A, int lda, long long M, int K, const float* Wt, int N, const float* bias, const float* res,
              int ldres, float* out, int ldo) {
    if (skip()) return;
    F32Gemm p{A, lda, 1, 1, K, Wt, bias, res, ldres, out, ldo, static_cast<int>(M), N, K};
    f32_gemm_kernel<false><<<dim3((N + 63) / 64, static_cast<unsigned>((M + 63) / 64)), 256, 0, s>>>(p);
    check();
  }
  void conv3(const float* x, int H, int W, int Cin, const float* Wt, int N, const float* bias, const float* res, float* out) {
    if (skip()) return;
    const long long M = static_cast<long long>(B) * H * W;
    F32Gemm p{x, 0, H, W, Cin, Wt, bias, res, N, out, N, static_cast<int>(M), N, Cin * 9};
    f32_gemm_kernel<true><<<dim3((N + 63) / 64, static_cast<unsigned>((M + 63) / 64)), 256, 0, s>>>(p);
    check();
  }
  void gn(const float* x0, int c0, const float* x1, int c1, int HW, int gw, int gb, float eps, const float* film, int ldfilm,
          int silu, float* out) {
    if (skip()) return;
    f32_gn_kernel<<<dim3(32, B), 256, 0, s>>>(x0, c0, x1, c1, HW, P(gw), P(gb), eps, film, ldfilm, silu, out);
    check();
  }
  void resample(const float* x0, int c0, const float* x1, int c1, int H, int W, int mode, float* out) {
    if (skip()) return;
    const long long n = static_cast<long long>(B) * H * W * (c0 + c1) * (mode == 1 ? 4 : 1) / (mode == 2 ? 4 : 1);
    f32_resample_kernel<<<blocks_for(n), 256, 0, s>>>(x0, c0, x1, c1, B, H, W, mode, out);
    check();
  }
  void ln(const float* x, int gw, int gb, float* out, long long M, int C) {
    if (skip()) return;
    f32_ln_kernel<<<static_cast<unsigned>((M + 7) / 8), 256, 0, s>>>(x, P(gw), P(gb), 1e-5f, out, M, C);
    check();
  }
  void attn(const float* Q, int ldq, const float* K, int ldk, const float* V, int ldv, float* O, int ldo, int N, int Nk,
            int heads, int D) {
    if (skip()) return;
    f32_attn_kernel<<<dim3((N + 63) / 64, heads, B), 64, 0, s>>>(Q, ldq, K, ldk, V, ldv, O, ldo, N, Nk, D,
                                                                  1.0f / sqrtf(static_cast<float>(D)));
    check();
  }

  struct T {
    float* p;
    int C, H;
  };

  // ResBlock._forward (openaimodel_enc.py:255-275)
  T res_block(const ResLayer& l, T x0, const T* x1, const float* semb) {
    const int cin = l.c0 + l.c1, E = u->emb_ch;
    const long long Mi = static_cast<long long>(B) * l.Hin * l.Win, Mo = static_cast<long long>(B) * l.Hout * l.Wout;
    const float* p1 = x1 ? x1->p : nullptr;
    const int c1 = x1 ? x1->C : 0;
    float* a = take(Mi * cin);
    gn(x0.p, l.c0, p1, c1, l.Hin * l.Win, l.gn1_w, l.gn1_b, 1e-5f, nullptr, 0, 1, a);
    const int mode = l.up ? 1 : (l.down ? 2 : 0);
    float* xs = nullptr;      // the (resampled, concatenated) block input for the skip branch
    if (mode || x1) {
      xs = take(Mo * cin);
      resample(x0.p, l.c0, p1, c1, l.Hin, l.Win, mode, xs);
    }
    if (mode) {
      float* ar = take(Mo * cin);
      resample(a, cin, nullptr, 0, l.Hin, l.Win, mode, ar);
      a = ar;
    }
    const float* xin = xs ? xs : x0.p;
    float* h1 = take(Mo * l.cout);
    conv3(a, l.Hout, l.Wout, cin, P(l.conv1_w), l.cout, P(l.conv1_b), nullptr, h1);
    float* film = take(static_cast<size_t>(B) * 2 * l.cout);
    linear(semb, E, B, E, P(l.emb_w), 2 * l.cout, P(l.emb_b), nullptr, 0, film, 2 * l.cout);
    float* a2 = take(Mo * l.cout);
    gn(h1, l.cout, nullptr, 0, l.Hout * l.Wout, l.gn2_w, l.gn2_b, 1e-5f, film, 2 * l.cout, 1, a2);
    const float* skip = xin;
    if (l.skipconv) {
      float* sk = take(Mo * l.cout);
      linear(xin, cin, Mo, cin, P(l.skip_w), l.cout, P(l.skip_b), nullptr, 0, sk, l.cout);
      skip = sk;
    }
    // the block output outlives the block's temporaries: the caller reserved out_slot BEFORE them
    T y{out_slot, l.cout, l.Hout};
    conv3(a2, l.Hout, l.Wout, l.cout, P(l.conv2_w), l.cout, P(l.conv2_b), skip, y.p);
    return y;
  }

  // SpatialTransformer.forward + BasicTransformerBlock._forward (attention.py:196-215,250-261)
  T st_block(const StLayer& l, T x, const float* ctx) {
    const int C = l.C, N = l.H * l.W, Tk = u->cfg.latent_unit, Dc = u->cfg.context_dim;
    const long long M = static_cast<long long>(B) * N;
    float* n = take(M * C);
    gn(x.p, C, nullptr, 0, N, l.norm_w, l.norm_b, 1e-6f, nullptr, 0, 0, n);
    float* t0 = take(M * C);
    linear(n, C, M, C, P(l.pin_w), C, P(l.pin_b), nullptr, 0, t0, C);
    float* l1 = take(M * C);
    ln(t0, l.ln1_w, l.ln1_b, l1, M, C);
    float* qkv = take(M * 3 * C);
    linear(l1, C, M, C, P(l.q1_w), C, nullptr, nullptr, 0, qkv, 3 * C);
    linear(l1, C, M, C, P(l.k1_w), C, nullptr, nullptr, 0, qkv + C, 3 * C);
    linear(l1, C, M, C, P(l.v1_w), C, nullptr, nullptr, 0, qkv + 2 * C, 3 * C);
    float* att = take(M * C);
    attn(qkv, 3 * C, qkv + C, 3 * C, qkv + 2 * C, 3 * C, att, C, N, N, l.heads, l.D);
    float* t1 = take(M * C);
    linear(att, C, M, C, P(l.o1_w), C, P(l.o1_b), t0, C, t1, C);
    float* l2 = take(M * C);
    ln(t1, l.ln2_w, l.ln2_b, l2, M, C);
    float* q2 = take(M * C);
    linear(l2, C, M, C, P(l.q2_w), C, nullptr, nullptr, 0, q2, C);
    float* kv = take(static_cast<size_t>(B) * Tk * 2 * C);
    linear(ctx, Dc, static_cast<long long>(B) * Tk, Dc, P(l.k2_w), C, nullptr, nullptr, 0, kv, 2 * C);
    linear(ctx, Dc, static_cast<long long>(B) * Tk, Dc, P(l.v2_w), C, nullptr, nullptr, 0, kv + C, 2 * C);
    float* att2 = take(M * C);
    attn(q2, C, kv, 2 * C, kv + C, 2 * C, att2, C, N, Tk, l.heads, l.D);
    float* t2 = take(M * C);
    linear(att2, C, M, C, P(l.o2_w), C, P(l.o2_b), t1, C, t2, C);
    float* l3 = take(M * C);
    ln(t2, l.ln3_w, l.ln3_b, l3, M, C);
    float* g = take(M * 8 * C);
    linear(l3, C, M, C, P(l.ff1_w), 8 * C, P(l.ff1_b), nullptr, 0, g, 8 * C);
    float* gg = take(M * 4 * C);
    if (!skip()) {
      f32_geglu_kernel<<<blocks_for(M * 4 * C), 256, 0, s>>>(g, gg, M, 4 * C);
      check();
    }
    float* t3 = take(M * C);
    linear(gg, 4 * C, M, 4 * C, P(l.ff2_w), C, P(l.ff2_b), t2, C, t3, C);
    T y{out_slot, C, l.H};
    linear(t3, C, M, C, P(l.pout_w), C, P(l.pout_b), x.p, C, y.p, C);
    return y;
  }

  float* out_slot = nullptr;

  int run(const float* x, const float* t, const float* ctx, float* out) {
    const encdiff_unet_config& c = u->cfg;
    const int mc = c.model_channels, S = c.image_size, E = u->emb_ch;
    const long long M0 = static_cast<long long>(B) * S * S;
    // time embedding MLP: Linear -> SiLU -> Linear; every ResBlock applies SiLU -> Linear to it (openaimodel_enc.py:218-224)
    float* temb = take(static_cast<size_t>(B) * mc);
    float* z1 = take(static_cast<size_t>(B) * E);
    float* a1 = take(static_cast<size_t>(B) * E);
    float* emb = take(static_cast<size_t>(B) * E);
    float* semb = take(static_cast<size_t>(B) * E);
    if (!skip()) {
      f32_temb_kernel<<<blocks_for(static_cast<long long>(B) * mc), 256, 0, s>>>(t, temb, B, mc);
      check();
    }
    linear(temb, mc, B, mc, P(u->te0_w), E, P(u->te0_b), nullptr, 0, z1, E);
    if (!skip()) {
      f32_silu_kernel<<<blocks_for(static_cast<long long>(B) * E), 256, 0, s>>>(z1, a1, static_cast<long long>(B) * E);
      check();
    }
    linear(a1, E, B, E, P(u->te2_w), E, P(u->te2_b), nullptr, 0, emb, E);
    if (!skip()) {
      f32_silu_kernel<<<blocks_for(static_cast<long long>(B) * E), 256, 0, s>>>(emb, semb, static_cast<long long>(B) * E);
      check();
    }
    // stem
    float* xn = take(M0 * c.in_channels);
    if (!skip()) {
      f32_layout_kernel<<<blocks_for(M0 * c.in_channels), 256, 0, s>>>(x, xn, B, c.in_channels, S * S, 1);
      check();
    }
    T h{take(M0 * mc), mc, S};
    conv3(xn, S, S, c.in_channels, P(u->stem_w), mc, P(u->stem_b), nullptr, h.p);
    std::vector<T> hs;
    hs.push_back(h);
    auto run_layers = [&](const std::vector<Layer>& ls, const T* skip_t) {
      bool first = true;
      for (const Layer& ly : ls) {
        if (ly.kind == 1) {
          const ResLayer& l = u->res[ly.idx];
          out_slot = take(static_cast<size_t>(B) * l.Hout * l.Wout * l.cout);
          const size_t mark = off;
          h = res_block(l, h, (first && skip_t) ? skip_t : nullptr, semb);
          off = mark;          // the block's temporaries are dead (same stream: the next block's kernels run after)
        } else if (ly.kind == 2) {
          const StLayer& l = u->st[ly.idx];
          out_slot = take(static_cast<size_t>(B) * l.H * l.W * l.C);
          const size_t mark = off;
          h = st_block(l, h, ctx);
          off = mark;
        }
        first = false;
      }
    };
    for (size_t i = 1; i < u->in_blocks.size(); ++i) {
      run_layers(u->in_blocks[i], nullptr);
      hs.push_back(h);
    }
    run_layers(u->mid, nullptr);
    for (auto& blk : u->out_blocks) {
      const T sk = hs.back();
      hs.pop_back();
      run_layers(blk, &sk);
    }
    // head: GN -> SiLU -> conv3x3 (openaimodel_enc.py:684-688,748)
    float* aH = take(M0 * mc);
    gn(h.p, mc, nullptr, 0, S * S, u->hn_w, u->hn_b, 1e-5f, nullptr, 0, 1, aH);
    float* yH = take(M0 * c.out_channels);
    conv3(aH, S, S, mc, P(u->hc_w), c.out_channels, P(u->hc_b), nullptr, yH);
    if (!skip()) {
      f32_layout_kernel<<<blocks_for(M0 * c.out_channels), 256, 0, s>>>(yH, out, B, c.out_channels, S * S, 0);
      check();
    }
    return err;
  }
};

}  // namespace

// floats of workspace the fp32 forward needs at batch B (temporaries of a block are recycled, block outputs are kept)
size_t unet_fp32_workspace_floats(UNet* u, int B) {
  Fp32Exec e{u, B, nullptr, nullptr, 0, true};
  e.run(nullptr, nullptr, nullptr, nullptr);
  return e.peak + 64;      // the peak, not the final offset: temporaries are released after every block
}

int unet_forward_fp32(UNet* u, int B, const float* x, const float* t, const float* ctx, float* out, float* ws,
                      size_t ws_floats, cudaStream_t s) {
  EDB_REQUIRE(u->bound, "forward_fp32 before bind");
  Fp32Exec e{u, B, s, ws, ws_floats, false};
  return e.run(x, t, ctx, out);
}

}  // namespace edb
