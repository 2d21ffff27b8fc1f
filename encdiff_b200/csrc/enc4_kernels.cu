// Kernels of the concept encoder's convolutional trunk (Encoder4.encoder, ldm/modules/diffusionmodules/
// openaimodel_enc.py:995-1014): 4x4 stride-2 and 3x3 convolutions as explicit tap-major im2col + tcgen05 GEMM (the im2col
// matrix is also the weight-gradient operand), their input gradients as GEMM + col2im gather, BatchNorm2d with batch
// statistics (training) or running statistics (eval), forward and backward, and the NCHW flatten in front of the Linear.
// All activations NHWC h16 [B*H*W, C]; statistics and parameter gradients fp32.
#include "common.cuh"
#include "kernels.cuh"

namespace edb {
namespace {

inline int e4_grid(long long n, int threads, int cap = 148 * 16) {
  const long long b = (n + threads - 1) / threads;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

// cols[(b, oy, ox)][tap * C + c] = x[b][oy*stride - pad + ky][ox*stride - pad + kx][c]   (tap = ky * k + kx; 0 outside)
// thread = (output pixel, kernel row ky, 8-channel chunk): the K taps of the row are K independent 16-byte loads in flight
// (one load per thread left the kernel at a third of the HBM bandwidth: too few bytes in flight per SM)
template <int K>
__global__ void __launch_bounds__(256) e4_im2col_kernel(const h16* __restrict__ x, h16* __restrict__ cols, int B, int H, int W,
                                                        int C, int stride, int pad, int Ho, int Wo) {
  pdl_enter();
  const int cv = C / 8;
  const long long total = static_cast<long long>(B) * Ho * Wo * K * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int ky = static_cast<int>(r % K);
    r /= K;
    const int ox = static_cast<int>(r % Wo);
    const long long bo = r / Wo;               // b * Ho + oy
    const int oy = static_cast<int>(bo % Ho);
    const long long b = bo / Ho;
    const int y = oy * stride - pad + ky, x0 = ox * stride - pad;
    const bool yin = y >= 0 && y < H;
    const h16* src = x + ((b * H + y) * W + x0) * C + c8 * 8;
    uint4 v[K];
#pragma unroll
    for (int kx = 0; kx < K; ++kx) {
      v[kx] = make_uint4(0u, 0u, 0u, 0u);
      if (yin && x0 + kx >= 0 && x0 + kx < W) v[kx] = *reinterpret_cast<const uint4*>(src + static_cast<long long>(kx) * C);
    }
    h16* dst = cols + (r * (K * K) + ky * K) * C + c8 * 8;
#pragma unroll
    for (int kx = 0; kx < K; ++kx) *reinterpret_cast<uint4*>(dst + static_cast<long long>(kx) * C) = v[kx];
  }
}

// first layer: fp32 NCHW image -> cols [M, 64] in the reference's flat weight order (ci * k*k + ky * k + kx), zero padded.
// thread = (output pixel, 8-column chunk): one 16-byte store; a chunk is 8 consecutive (ci, tap) columns
__global__ void __launch_bounds__(256) e4_im2col_image_kernel(const float* __restrict__ x, h16* __restrict__ cols, int B, int Cin,
                                                              int H, int W, int k, int stride, int pad, int Ho, int Wo) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * Ho * Wo * 8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int chunk = static_cast<int>(i & 7);
    const long long m = i >> 3;
    const int ox = static_cast<int>(m % Wo);
    const int oy = static_cast<int>((m / Wo) % Ho);
    const long long b = m / (static_cast<long long>(Wo) * Ho);
    // k = 4: a chunk is two kernel rows (ky = 2 * (chunk & 1) + {0, 1}) of input channel ci = chunk >> 1
    const int ci = chunk >> 1, x0 = ox * stride - pad;
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int y = oy * stride - pad + 2 * (chunk & 1) + (j >> 2), xx = x0 + (j & 3);
      f[j] = 0.f;
      if (ci < Cin && y >= 0 && y < H && xx >= 0 && xx < W) f[j] = x[((b * Cin + ci) * H + y) * W + xx];
    }
    *reinterpret_cast<uint4*>(cols + i * 8) = pack8(f);
  }
}

// dx[b][y][x][c] = (sum over the (oy, ox, tap) that read input pixel (y, x)) dcols[(b, oy, ox)][tap * C + c]
//                  [* (gate[..] > 0)]  [+ add[..]]
// gate: the block input of an EncResBlock (its leading ReLU, openaimodel_enc.py:976); add: the gradient of the skip path
// (the skip gradient carries the range scale add_from, the result add_to: both powers of two, the factor is exact).
// Per dimension a pixel is read by T = ceil(K / S) taps: ky = ((y + pad) % S) + S * a.  All T*T loads are issued before
// the sum.
template <int K, int S>
__global__ void __launch_bounds__(256) e4_col2im_kernel(const h16* __restrict__ dcols, h16* __restrict__ dx,
                                                        const h16* __restrict__ gate, const h16* __restrict__ add,
                                                        const float* __restrict__ add_from, const float* __restrict__ add_to,
                                                        int B, int H, int W, int C, int pad, int Ho, int Wo) {
  pdl_enter();
  constexpr int T = (K + S - 1) / S;
  const int cv = C / 8;
  const float add_f = (add && add_from && add_to) ? add_to[0] * add_from[1] : 1.f;
  const long long total = static_cast<long long>(B) * H * W * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int xx = static_cast<int>(r % W);
    r /= W;
    const int y = static_cast<int>(r % H);
    const long long b = r / H;
    uint4 v[T * T];
    uint4 gv = make_uint4(0u, 0u, 0u, 0u), av = gv;
    if (gate) gv = *reinterpret_cast<const uint4*>(gate + i * 8);
    if (add) av = *reinterpret_cast<const uint4*>(add + i * 8);
#pragma unroll
    for (int a = 0; a < T; ++a) {
      const int ky = (y + pad) % S + S * a, oy = (y + pad - ky) / S;      // y + pad - ky is a non-negative multiple of S or < 0
      const bool yok = ky < K && y + pad - ky >= 0 && oy < Ho;
#pragma unroll
      for (int c = 0; c < T; ++c) {
        const int kx = (xx + pad) % S + S * c, ox = (xx + pad - kx) / S;
        const bool ok = yok && kx < K && xx + pad - kx >= 0 && ox < Wo;
        v[a * T + c] = make_uint4(0u, 0u, 0u, 0u);
        if (ok)
          v[a * T + c] = *reinterpret_cast<const uint4*>(dcols + ((b * Ho + oy) * Wo + ox) * (static_cast<long long>(K * K) * C) +
                                                         (ky * K + kx) * C + c8 * 8);
      }
    }
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int t = 0; t < T * T; ++t) {
      float f[8];
      unpack8(v[t], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
    if (gate) {
      float g[8];
      unpack8(gv, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = g[j] > 0.f ? acc[j] : 0.f;
    }
    if (add) {
      float g[8];
      unpack8(av, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += g[j] * add_f;
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8(acc);
  }
}

// ---- BatchNorm2d (torch.nn.BatchNorm2d, eps 1e-5, momentum 0.1) ---------------------------------------------------------
// per-channel sums of v and v*w over the M rows: sums[c] += sum v, sums[C + c] += sum v*w, where
//   forward statistics:  v = x,            w = x
//   backward:            v = dy (masked),  w = xhat = (x - mean) * rstd      (mask: post-activation output > 0)
//                        and sums[2C] = max |v| (float bits, atomicMax) for the range re-scaling of the layer's dx
// thread = (row lane, 8-channel chunk); partial sums through shared memory, one atomic per channel and block.
template <bool BWD>
__global__ void __launch_bounds__(256) e4_bn_sums_kernel(const h16* __restrict__ x, const h16* __restrict__ dy,
                                                         const h16* __restrict__ out, const float* __restrict__ stat,
                                                         float* __restrict__ sums, long long M, int C) {
  pdl_enter();
  extern __shared__ float sh[];                 // [rows][2][C]
  const int cv = C / 8;
  const int chunk = threadIdx.x % cv, rl = threadIdx.x / cv, rows = blockDim.x / cv;
  float s[8], q[8], mu[8], rs[8];
  float mx = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s[j] = q[j] = 0.f;
    mu[j] = BWD ? stat[chunk * 8 + j] : 0.f;
    rs[j] = BWD ? stat[C + chunk * 8 + j] : 0.f;
  }
  if (rl < rows) {
    // four rows per iteration: 4 (forward) or 12 (backward) independent 16-byte loads in flight per thread
    const long long stride = static_cast<long long>(gridDim.x) * rows;
    for (long long m0 = blockIdx.x * static_cast<long long>(rows) + rl; m0 < M; m0 += 4 * stride) {
      uint4 xv[4], dv[4], ov[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long m = m0 + u * stride;
        const bool ok = m < M;
        xv[u] = dv[u] = ov[u] = make_uint4(0u, 0u, 0u, 0u);
        if (ok) xv[u] = *reinterpret_cast<const uint4*>(x + m * C + chunk * 8);
        if (BWD && ok) dv[u] = *reinterpret_cast<const uint4*>(dy + m * C + chunk * 8);
        if (BWD && ok && out) ov[u] = *reinterpret_cast<const uint4*>(out + m * C + chunk * 8);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8(xv[u], f);
        if (BWD) {
          float d[8];
          unpack8(dv[u], d);
          if (out) {
            float o[8];
            unpack8(ov[u], o);
#pragma unroll
            for (int j = 0; j < 8; ++j) d[j] = o[j] > 0.f ? d[j] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s[j] += d[j];
            q[j] += d[j] * (f[j] - mu[j]) * rs[j];
            mx = fmaxf(mx, fabsf(d[j]));
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s[j] += f[j];
            q[j] += f[j] * f[j];
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sh[(rl * 2) * C + chunk * 8 + j] = s[j];
      sh[(rl * 2 + 1) * C + chunk * 8 + j] = q[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    const int which = c / C, ch = c - which * C;
    float a = 0.f;
    for (int r = 0; r < rows; ++r) a += sh[(r * 2 + which) * C + ch];
    atomicAdd(&sums[c], a);
  }
  if (BWD) {
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(reinterpret_cast<unsigned*>(sums + 2 * C), __float_as_uint(mx));
  }
}

// forward apply: stat[c] = mean, stat[C + c] = rstd (batch statistics from `sums`, or the running statistics in eval mode);
// y = (x - mean) * rstd * gamma + beta, optional ReLU.  Block 0 also updates the running statistics (training).
__global__ void __launch_bounds__(256) e4_bn_apply_kernel(const h16* __restrict__ x, const float* __restrict__ sums,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          float* __restrict__ run_mean, float* __restrict__ run_var,
                                                          float* __restrict__ stat, h16* __restrict__ y, long long M, int C,
                                                          float eps, float momentum, int training, int relu) {
  pdl_enter();
  extern __shared__ float sh[];                 // [2][C]: scale a, shift b
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float mean, var;
    if (training) {
      mean = sums[c] / static_cast<float>(M);
      var = fmaxf(sums[C + c] / static_cast<float>(M) - mean * mean, 0.f);
    } else {
      mean = run_mean[c];
      var = run_var[c];
    }
    const float rstd = 1.0f / sqrtf(var + eps);
    if (blockIdx.x == 0) {
      stat[c] = mean;
      stat[C + c] = rstd;
      if (training && run_mean) {          // torch: running = (1 - momentum) * running + momentum * batch (unbiased variance)
        const float unb = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unb;
      }
    }
    const float a = gamma[c] * rstd;
    sh[c] = a;
    sh[C + c] = beta[c] - mean * a;
  }
  __syncthreads();
  // blockDim is a multiple of C / 8, so a thread stays on one 8-channel chunk: coefficients in registers, four vectors
  // in flight per iteration
  const int cv = C / 8;
  const long long total = M * cv, T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % cv) * 8;
  float ca[8], cb[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    ca[j] = sh[c0 + j];
    cb[j] = sh[C + c0 + j];
  }
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < total; i0 += 4 * T) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i0 + u * T < total) v[u] = *reinterpret_cast<const uint4*>(x + (i0 + u * T) * 8);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (i0 + u * T >= total) break;
      float f[8];
      unpack8(v[u], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float w = f[j] * ca[j] + cb[j];
        f[j] = relu ? fmaxf(w, 0.f) : w;
      }
      *reinterpret_cast<uint4*>(y + (i0 + u * T) * 8) = pack8(f);
    }
  }
}

// backward apply (training statistics): dx = gamma * rstd * (dy - mean(dy) - xhat * mean(dy * xhat)); dgamma += sum dy*xhat,
// dbeta += sum dy (block 0; multiplied by 1/scale of the incoming gradient's range scale gs_in).  In eval mode the
// statistics are constants: dx = gamma * rstd * dy.
// Range re-scaling: activation gradients are 16-bit, and every BatchNorm multiplies them by gamma * rstd, which over the
// trunk's seven normalisations can move them by orders of magnitude either way.  Each backward BatchNorm therefore picks a
// power of two r that puts max|gamma rstd| * max|dy| at 2^-4..2^-3, multiplies dx by it and publishes the cumulative
// scale gs_out = gs_in * r for the consumers of dx (weight-gradient GEMMs divide by it, like the UNet's single scale).
__global__ void __launch_bounds__(256) e4_bn_bwd_apply_kernel(const h16* __restrict__ x, const h16* __restrict__ dy,
                                                              const h16* __restrict__ out, const float* __restrict__ stat,
                                                              const float* __restrict__ sums, const float* __restrict__ gamma,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                              const float* __restrict__ gs_in, float* __restrict__ gs_out,
                                                              h16* __restrict__ dx, long long M, int C, int training) {
  pdl_enter();
  extern __shared__ float sh[];                 // [4][C]: mean, rstd, k1 = mean(dy), k2 = mean(dy * xhat)
  __shared__ unsigned s_amax;
  if (threadIdx.x == 0) s_amax = 0u;
  __syncthreads();
  const float ginv = gs_in ? gs_in[1] : 1.f;
  float amax = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) amax = fmaxf(amax, fabsf(gamma[c]) * stat[C + c]);
  amax = warp_max(amax);
  if ((threadIdx.x & 31) == 0) atomicMax(&s_amax, __float_as_uint(amax));
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    sh[c] = stat[c];
    sh[C + c] = stat[C + c];
    sh[2 * C + c] = training ? sums[c] / static_cast<float>(M) : 0.f;
    sh[3 * C + c] = training ? sums[C + c] / static_cast<float>(M) : 0.f;
    if (blockIdx.x == 0) {
      if (dbeta) atomicAdd(&dbeta[c], sums[c] * ginv);
      if (dgamma) atomicAdd(&dgamma[c], sums[C + c] * ginv);
    }
  }
  __syncthreads();
  float r = 1.f;
#ifndef EDB_ACT_BF16
  if (gs_out) {
    const float est = __uint_as_float(s_amax) * sums[2 * C];      // max|gamma rstd| * max|dy| (float bits via atomicMax)
    if (est > 0.f && isfinite(est)) {
      int e;
      frexpf(est, &e);
      r = ldexpf(1.f, -3 - e);
    }
  }
#endif
  if (gs_out && blockIdx.x == 0 && threadIdx.x == 0) {
    gs_out[0] = gs_in[0] * r;
    gs_out[1] = gs_in[1] / r;
  }
  const int cv = C / 8;
  const long long total = M * cv, T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % cv) * 8;      // blockDim is a multiple of C / 8: one chunk per thread
  float mu[8], ga[8], k1[8], k2[8], rs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mu[j] = sh[c0 + j];
    rs[j] = sh[C + c0 + j];
    ga[j] = r * gamma[c0 + j] * rs[j];
    k1[j] = sh[2 * C + c0 + j];
    k2[j] = sh[3 * C + c0 + j];
  }
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < total; i0 += 2 * T) {
    uint4 xv[2], dv[2], ov[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * T;
      if (i < total) {
        xv[u] = *reinterpret_cast<const uint4*>(x + i * 8);
        dv[u] = *reinterpret_cast<const uint4*>(dy + i * 8);
        if (out) ov[u] = *reinterpret_cast<const uint4*>(out + i * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * T;
      if (i >= total) break;
      float f[8], d[8];
      unpack8(xv[u], f);
      unpack8(dv[u], d);
      if (out) {
        float o[8];
        unpack8(ov[u], o);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = o[j] > 0.f ? d[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (f[j] - mu[j]) * rs[j];
        f[j] = ga[j] * (d[j] - k1[j] - xh * k2[j]);
      }
      *reinterpret_cast<uint4*>(dx + i * 8) = pack8(f);
    }
  }
}

// NHWC [B, HW, C] <-> the reference's NCHW flatten [B, C * HW] (View((-1, 128 * 4 * 4)), openaimodel_enc.py:1012)
__global__ void e4_flatten_kernel(const h16* __restrict__ in, h16* __restrict__ out, int B, int HW, int C, int to_flat) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * HW * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const int p = static_cast<int>((i / C) % HW);
    const long long b = i / (static_cast<long long>(C) * HW);
    const long long flat = (b * C + c) * HW + p;
    if (to_flat) out[flat] = in[i];
    else out[i] = in[flat];
  }
}

// grad_ref[co][ci][tap] += stage[co][tap * C + ci]     (tap-major weight-gradient staging -> reference layout); one launch
// for all im2col convolutions of the backward (blockIdx.y = entry)
__global__ void __launch_bounds__(256) e4_wgrad_finalize_kernel(E4Finalize f) {
  pdl_enter();
  const int e = blockIdx.y;
  const float* __restrict__ stage = f.stage[e];
  float* __restrict__ grad = f.grad[e];
  const int C = f.C, KK = f.KK[e];
  const long long total = static_cast<long long>(f.Cout) * C * KK;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(i % C);
    const int tap = static_cast<int>((i / C) % KK);
    const long long co = i / (static_cast<long long>(C) * KK);
    grad[(co * C + ci) * KK + tap] += stage[i];
  }
}

}  // namespace

int launch_e4_im2col(const h16* x, h16* cols, int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s) {
  if ((C & 7) || (k != 3 && k != 4)) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int grid = e4_grid(static_cast<long long>(B) * Ho * Wo * k * (C / 8), 256);
  if (k == 4) launch_k(e4_im2col_kernel<4>, grid, 256, 0, s, x, cols, B, H, W, C, stride, pad, Ho, Wo);
  else launch_k(e4_im2col_kernel<3>, grid, 256, 0, s, x, cols, B, H, W, C, stride, pad, Ho, Wo);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_im2col_image(const float* x, h16* cols, int B, int Cin, int H, int W, int k, int stride, int pad, cudaStream_t s) {
  if (k != 4 || Cin > 4) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  launch_k(e4_im2col_image_kernel, e4_grid(static_cast<long long>(B) * Ho * Wo * 8, 256), 256, 0, s, x, cols, B, Cin, H, W, k, stride,
           pad, Ho, Wo);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_col2im(const h16* dcols, h16* dx, const h16* gate, const h16* add, const float* add_from, const float* add_to,
                     int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s) {
  if (C & 7) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int grid = e4_grid(static_cast<long long>(B) * H * W * (C / 8), 256);
  if (k == 4 && stride == 2)
    launch_k(e4_col2im_kernel<4, 2>, grid, 256, 0, s, dcols, dx, gate, add, add_from, add_to, B, H, W, C, pad, Ho, Wo);
  else if (k == 3 && stride == 1)
    launch_k(e4_col2im_kernel<3, 1>, grid, 256, 0, s, dcols, dx, gate, add, add_from, add_to, B, H, W, C, pad, Ho, Wo);
  else
    return static_cast<int>(cudaErrorInvalidValue);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_bn_sums(const h16* x, const h16* dy, const h16* out, const float* stat, float* sums, long long M, int C, int bwd,
                      cudaStream_t s) {
  if ((C & 7) || C > 2048 || 256 % (C / 8) != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int rows = 256 / (C / 8);
  const size_t smem = static_cast<size_t>(rows) * 2 * C * sizeof(float);
  const int grid = e4_grid(M, rows * 4, 148 * 6);
  if (bwd) launch_k(e4_bn_sums_kernel<true>, grid, 256, smem, s, x, dy, out, stat, sums, M, C);
  else launch_k(e4_bn_sums_kernel<false>, grid, 256, smem, s, x, dy, out, stat, sums, M, C);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_bn_apply(const h16* x, const float* sums, const float* gamma, const float* beta, float* run_mean, float* run_var,
                       float* stat, h16* y, long long M, int C, float eps, float momentum, int training, int relu, cudaStream_t s) {
  launch_k(e4_bn_apply_kernel, e4_grid(M * (C / 8), 256), 256, 2 * C * sizeof(float), s, x, sums, gamma, beta, run_mean, run_var, stat,
           y, M, C, eps, momentum, training, relu);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_bn_bwd_apply(const h16* x, const h16* dy, const h16* out, const float* stat, const float* sums, const float* gamma,
                           float* dgamma, float* dbeta, const float* gs_in, float* gs_out, h16* dx, long long M, int C,
                           int training, cudaStream_t s) {
  launch_k(e4_bn_bwd_apply_kernel, e4_grid(M * (C / 8), 256), 256, 4 * C * sizeof(float), s, x, dy, out, stat, sums, gamma, dgamma,
           dbeta, gs_in, gs_out, dx, M, C, training);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_flatten(const h16* in, h16* out, int B, int HW, int C, int to_flat, cudaStream_t s) {
  launch_k(e4_flatten_kernel, e4_grid(static_cast<long long>(B) * HW * C, 256), 256, 0, s, in, out, B, HW, C, to_flat);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_wgrad_finalize(const E4Finalize& f, cudaStream_t s) {
  if (f.n <= 0) return 0;
  int kk = 0;
  for (int e = 0; e < f.n; ++e) kk = f.KK[e] > kk ? f.KK[e] : kk;
  dim3 grid(e4_grid(static_cast<long long>(f.Cout) * f.C * kk, 256, 148 * 4), f.n);
  launch_k(e4_wgrad_finalize_kernel, grid, 256, 0, s, f);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
