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
__global__ void e4_im2col_kernel(const h16* __restrict__ x, h16* __restrict__ cols, int B, int H, int W, int C, int k,
                                 int stride, int pad, int Ho, int Wo) {
  pdl_enter();
  const int cv = C / 8, KK = k * k;
  const long long total = static_cast<long long>(B) * Ho * Wo * KK * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int tap = static_cast<int>(r % KK);
    r /= KK;
    const int ox = static_cast<int>(r % Wo);
    r /= Wo;
    const int oy = static_cast<int>(r % Ho), b = static_cast<int>(r / Ho);
    const int y = oy * stride - pad + tap / k, xx = ox * stride - pad + tap % k;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (y >= 0 && y < H && xx >= 0 && xx < W)
      v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(b) * H + y) * W + xx) * C + c8 * 8);
    *reinterpret_cast<uint4*>(cols + i * 8) = v;
  }
}

// first layer: fp32 NCHW image -> cols [M, 64] in the reference's flat weight order (ci * k*k + ky * k + kx), zero padded
__global__ void e4_im2col_image_kernel(const float* __restrict__ x, h16* __restrict__ cols, int B, int Cin, int H, int W, int k,
                                       int stride, int pad, int Ho, int Wo) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * Ho * Wo;
  const int KK = k * k;
  for (long long m = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; m < total;
       m += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ox = static_cast<int>(m % Wo);
    const int oy = static_cast<int>((m / Wo) % Ho), b = static_cast<int>(m / (static_cast<long long>(Wo) * Ho));
    h16 row[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) row[i] = f2h(0.f);
    for (int ci = 0; ci < Cin; ++ci) {
      const float* plane = x + (static_cast<long long>(b) * Cin + ci) * H * W;
      for (int t = 0; t < KK; ++t) {
        const int y = oy * stride - pad + t / k, xx = ox * stride - pad + t % k;
        if (y >= 0 && y < H && xx >= 0 && xx < W && ci * KK + t < 64) row[ci * KK + t] = f2h(plane[y * W + xx]);
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(cols + m * 64);
    const uint4* src = reinterpret_cast<const uint4*>(row);
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = src[i];
  }
}

// dx[b][y][x][c] = (sum over the (oy, ox, tap) that read input pixel (y, x)) dcols[(b, oy, ox)][tap * C + c]
//                  [* (gate[..] > 0)]  [+ add[..]]
// gate: the block input of an EncResBlock (its leading ReLU, openaimodel_enc.py:976); add: the gradient of the skip path
// (the skip gradient carries the range scale add_from, the result add_to: both powers of two, the factor is exact)
__global__ void e4_col2im_kernel(const h16* __restrict__ dcols, h16* __restrict__ dx, const h16* __restrict__ gate,
                                 const h16* __restrict__ add, const float* __restrict__ add_from,
                                 const float* __restrict__ add_to, int B, int H, int W, int C, int k, int stride, int pad,
                                 int Ho, int Wo) {
  pdl_enter();
  const int cv = C / 8, KK = k * k;
  const float add_f = (add && add_from && add_to) ? add_to[0] * add_from[1] : 1.f;
  const long long total = static_cast<long long>(B) * H * W * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int xx = static_cast<int>(r % W);
    r /= W;
    const int y = static_cast<int>(r % H), b = static_cast<int>(r / H);
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 0.f;
    for (int ky = 0; ky < k; ++ky) {
      const int ty = y + pad - ky;
      if (ty < 0 || ty % stride) continue;
      const int oy = ty / stride;
      if (oy >= Ho) continue;
      for (int kx = 0; kx < k; ++kx) {
        const int tx = xx + pad - kx;
        if (tx < 0 || tx % stride) continue;
        const int ox = tx / stride;
        if (ox >= Wo) continue;
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(dcols + ((static_cast<long long>(b) * Ho + oy) * Wo + ox) * (static_cast<long long>(KK) * C) +
                                                (ky * k + kx) * C + c8 * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += f[j];
      }
    }
    if (gate) {
      float g[8];
      unpack8(*reinterpret_cast<const uint4*>(gate + i * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = g[j] > 0.f ? a[j] : 0.f;
    }
    if (add) {
      float g[8];
      unpack8(*reinterpret_cast<const uint4*>(add + i * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += g[j] * add_f;
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8(a);
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
    for (long long m = blockIdx.x * static_cast<long long>(rows) + rl; m < M; m += static_cast<long long>(gridDim.x) * rows) {
      float f[8];
      unpack8(*reinterpret_cast<const uint4*>(x + m * C + chunk * 8), f);
      if (BWD) {
        float d[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + m * C + chunk * 8), d);
        if (out) {
          float o[8];
          unpack8(*reinterpret_cast<const uint4*>(out + m * C + chunk * 8), o);
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
  const int cv = C / 8;
  const long long total = M * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    float f[8];
    unpack8(*reinterpret_cast<const uint4*>(x + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float v = f[j] * sh[c0 + j] + sh[C + c0 + j];
      f[j] = relu ? fmaxf(v, 0.f) : v;
    }
    *reinterpret_cast<uint4*>(y + i * 8) = pack8(f);
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
  const long long total = M * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    float f[8], d[8];
    unpack8(*reinterpret_cast<const uint4*>(x + i * 8), f);
    unpack8(*reinterpret_cast<const uint4*>(dy + i * 8), d);
    if (out) {
      float o[8];
      unpack8(*reinterpret_cast<const uint4*>(out + i * 8), o);
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = o[j] > 0.f ? d[j] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = c0 + j;
      const float xh = (f[j] - sh[c]) * sh[C + c];
      f[j] = r * gamma[c] * sh[C + c] * (d[j] - sh[2 * C + c] - xh * sh[3 * C + c]);
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8(f);
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

// grad_ref[co][ci][tap] += stage[co][tap * C + ci]; stage = 0     (tap-major weight-gradient staging -> reference layout)
__global__ void e4_wgrad_finalize_kernel(float* __restrict__ stage, float* __restrict__ grad, int Cout, int C, int KK) {
  pdl_enter();
  const long long total = static_cast<long long>(Cout) * C * KK;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(i % C);
    const int tap = static_cast<int>((i / C) % KK);
    const long long co = i / (static_cast<long long>(C) * KK);
    grad[(co * C + ci) * KK + tap] += stage[i];
    stage[i] = 0.f;
  }
}

}  // namespace

int launch_e4_im2col(const h16* x, h16* cols, int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s) {
  if (C & 7) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  launch_k(e4_im2col_kernel, e4_grid(static_cast<long long>(B) * Ho * Wo * k * k * (C / 8), 256), 256, 0, s, x, cols, B, H, W, C, k,
           stride, pad, Ho, Wo);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_im2col_image(const float* x, h16* cols, int B, int Cin, int H, int W, int k, int stride, int pad, cudaStream_t s) {
  if (Cin * k * k > 64) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  launch_k(e4_im2col_image_kernel, e4_grid(static_cast<long long>(B) * Ho * Wo, 128), 128, 0, s, x, cols, B, Cin, H, W, k, stride, pad,
           Ho, Wo);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_col2im(const h16* dcols, h16* dx, const h16* gate, const h16* add, const float* add_from, const float* add_to,
                     int B, int H, int W, int C, int k, int stride, int pad, cudaStream_t s) {
  if (C & 7) return static_cast<int>(cudaErrorInvalidValue);
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  launch_k(e4_col2im_kernel, e4_grid(static_cast<long long>(B) * H * W * (C / 8), 256), 256, 0, s, dcols, dx, gate, add, add_from, add_to, B,
           H, W, C, k, stride, pad, Ho, Wo);
  return static_cast<int>(cudaGetLastError());
}
int launch_e4_bn_sums(const h16* x, const h16* dy, const h16* out, const float* stat, float* sums, long long M, int C, int bwd,
                      cudaStream_t s) {
  if ((C & 7) || C > 2048 || 256 % (C / 8) != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int rows = 256 / (C / 8);
  const size_t smem = static_cast<size_t>(rows) * 2 * C * sizeof(float);
  const int grid = e4_grid(M, rows * 4, 148 * 4);
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
int launch_e4_wgrad_finalize(float* stage, float* grad, int Cout, int C, int KK, cudaStream_t s) {
  launch_k(e4_wgrad_finalize_kernel, e4_grid(static_cast<long long>(Cout) * C * KK, 256), 256, 0, s, stage, grad, Cout, C, KK);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
