// Elementwise / layout glue kernels (all HBM-bound, vectorised where the layout allows).
#include "common.cuh"
#include "kernels.cuh"

namespace edb {

static inline int grid_for(long long n, int block, int cap = 148 * 16) {
  long long g = (n + block - 1) / block;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

// ---- timestep embedding (util.py:179-199): [cos(t f_i) | sin(t f_i)], f_i = exp(-ln(1e4) i / half) ---------------
__global__ void temb_kernel(const float* __restrict__ t, h16* __restrict__ out, int B, int dim) {
  pdl_enter();
  const int half = dim / 2;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < B * half; idx += gridDim.x * blockDim.x) {
    const int b = idx / half, i = idx - b * half;
    const float freq = expf(-9.210340371976184f * static_cast<float>(i) / static_cast<float>(half));
    const float a = t[b] * freq;
    out[b * dim + i] = f2h(cosf(a));
    out[b * dim + half + i] = f2h(sinf(a));
    if ((dim & 1) && i == 0) out[b * dim + dim - 1] = f2h(0.f);
  }
}
int launch_temb(const float* t, h16* out, int B, int dim, cudaStream_t s) {
  launch_k(temb_kernel, grid_for(static_cast<long long>(B) * (dim / 2), 256), 256, 0, s, t, out, B, dim);
  return static_cast<int>(cudaGetLastError());
}

// ---- SiLU on the (tiny) embedding MLP ------------------------------------------------------------------------------
__global__ void silu_fwd_kernel(const float* __restrict__ x, h16* __restrict__ y, long long n) {
  pdl_enter();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    y[i] = f2h(silu_f(x[i]));
}
__global__ void silu_bwd_kernel(const float* __restrict__ x, const h16* __restrict__ dy, h16* __restrict__ dx,
                                long long n) {
  pdl_enter();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    dx[i] = f2h(h2f(dy[i]) * silu_grad_f(x[i]));
}
int launch_silu_fwd(const float* x, h16* y, long long n, cudaStream_t s) {
  launch_k(silu_fwd_kernel, grid_for(n, 256), 256, 0, s, x, y, n);
  return static_cast<int>(cudaGetLastError());
}
int launch_silu_bwd(const float* x, const h16* dy, h16* dx, long long n, cudaStream_t s) {
  launch_k(silu_bwd_kernel, grid_for(n, 256), 256, 0, s, x, dy, dx, n);
  return static_cast<int>(cudaGetLastError());
}

// ---- GEGLU (attention.py:37-45) -------------------------------------------------------------------------------------
// Indexing is 32-bit with a shift when the row width is a power of two (always, for the shipped widths): a 64-bit
// division per 8 elements cost more instructions than the GELU itself.
struct RowSplit {
  unsigned nv;      // 8-element chunks per row
  int shift;        // log2(nv) or -1
  __device__ __forceinline__ void split(unsigned idx, unsigned* row, unsigned* chunk) const {
    if (shift >= 0) {
      *row = idx >> shift;
      *chunk = idx & (nv - 1);
    } else {
      *row = idx / nv;
      *chunk = idx - *row * nv;
    }
  }
};
static RowSplit make_row_split(int nv) {
  RowSplit r;
  r.nv = static_cast<unsigned>(nv);
  r.shift = -1;
  for (int s = 0; s < 31; ++s)
    if ((1 << s) == nv) r.shift = s;
  return r;
}

__global__ void __launch_bounds__(256) geglu_fwd_kernel(const h16* __restrict__ g, h16* __restrict__ out, unsigned total,
                                                        int inner, RowSplit rs) {
  pdl_enter();
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    unsigned row, ch;
    rs.split(idx, &row, &ch);
    const int c = static_cast<int>(ch) * 8;
    const h16* gr = g + static_cast<long long>(row) * 2 * inner + c;
    const uint4 rv = *reinterpret_cast<const uint4*>(gr);
    const uint4 rg = *reinterpret_cast<const uint4*>(gr + inner);
    float v[8], gt[8];
    unpack8(rv, v);
    unpack8(rg, gt);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] *= gelu_erf_f(gt[k]);
    *reinterpret_cast<uint4*>(out + static_cast<long long>(row) * inner + c) = pack8(v);
  }
}
__global__ void __launch_bounds__(256) geglu_bwd_kernel(const h16* __restrict__ g, const h16* __restrict__ dout,
                                                        h16* __restrict__ dg, unsigned total, int inner, RowSplit rs) {
  pdl_launch_dependents();
  {   // the FF1 pre-activations were saved by the forward: first iteration's lines go to L2 before the wait
    if ((threadIdx.x & 7) == 0) {
      unsigned idx0 = blockIdx.x * blockDim.x + threadIdx.x;
      for (int it = 0; it < 4 && idx0 < total; ++it, idx0 += gridDim.x * blockDim.x) {
        unsigned row, ch;
        rs.split(idx0, &row, &ch);
        const long long o2 = static_cast<long long>(row) * 2 * inner + static_cast<int>(ch) * 8;
        l2_prefetch_line(g + o2);
        l2_prefetch_line(g + o2 + inner);
      }
    }
  }
  pdl_wait();
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    unsigned row, ch;
    rs.split(idx, &row, &ch);
    const int c = static_cast<int>(ch) * 8;
    const long long o2 = static_cast<long long>(row) * 2 * inner + c;
    const uint4 rv = *reinterpret_cast<const uint4*>(g + o2);
    const uint4 rg = *reinterpret_cast<const uint4*>(g + o2 + inner);
    const uint4 rd = *reinterpret_cast<const uint4*>(dout + static_cast<long long>(row) * inner + c);
    float v[8], gt[8], d[8], dv[8], dgt[8];
    unpack8(rv, v);
    unpack8(rg, gt);
    unpack8(rd, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float y, dy;
      gelu_erf_both_f(gt[k], &y, &dy);
      dv[k] = d[k] * y;
      dgt[k] = d[k] * v[k] * dy;
    }
    *reinterpret_cast<uint4*>(dg + o2) = pack8(dv);
    *reinterpret_cast<uint4*>(dg + o2 + inner) = pack8(dgt);
  }
}
int launch_geglu_fwd(const h16* g, h16* out, long long M, int inner, cudaStream_t s) {
  if (inner % 8 || M * (inner / 8) >= (1LL << 31)) return static_cast<int>(cudaErrorInvalidValue);
  const long long total = M * (inner / 8);
  launch_k(geglu_fwd_kernel, grid_for(total, 256), 256, 0, s, g, out, static_cast<unsigned>(total), inner,
           make_row_split(inner / 8));
  return static_cast<int>(cudaGetLastError());
}
int launch_geglu_bwd(const h16* g, const h16* dout, h16* dg, long long M, int inner, cudaStream_t s) {
  if (inner % 8 || M * (inner / 8) >= (1LL << 31)) return static_cast<int>(cudaErrorInvalidValue);
  const long long total = M * (inner / 8);
  launch_k(geglu_bwd_kernel, grid_for(total, 256), 256, 0, s, g, dout, dg, static_cast<unsigned>(total), inner,
           make_row_split(inner / 8));
  return static_cast<int>(cudaGetLastError());
}

// ---- bias gradient: column sums -------------------------------------------------------------------------------------
// block = (32 column-lanes) x (8 row-lanes); grid.x tiles columns by 32, grid.y splits rows
__global__ void __launch_bounds__(256) colsum_kernel(const h16* __restrict__ x, long long M, int ld, int col0, int N,
                                                     float* __restrict__ dst0, float* __restrict__ dst1,
                                                     const float* __restrict__ gscale) {
  pdl_enter();
  __shared__ float red[8][33];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cl;
  float a = 0.f;
  if (col < N) {
    for (long long r = blockIdx.y * 8 + rl; r < M; r += static_cast<long long>(gridDim.y) * 8)
      a += h2f(x[r * ld + col0 + col]);
  }
  red[rl][cl] = a;
  __syncthreads();
  if (rl == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][cl];
    if (gscale) t *= gscale[1];
    atomicAdd(dst0 + col, t);
    if (dst1) atomicAdd(dst1 + col, t);
  }
}
int launch_colsum(const h16* x, long long M, int ld, int col0, int N, float* dst0, float* dst1, const float* gscale,
                  cudaStream_t s) {
  const int gx = (N + 31) / 32;
  long long gy = (M + 8 * 32 - 1) / (8 * 32);   // >= 32 rows per thread
  const long long cap = (148 * 8 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  launch_k(colsum_kernel, dim3(gx, static_cast<unsigned>(gy)), 256, 0, s, x, M, ld, col0, N, dst0, dst1, gscale);
  return static_cast<int>(cudaGetLastError());
}

// ---- input stem: (optional q_sample) + NCHW fp32 -> im2col h16 [B*H*W, 64], column ci*9 + tap ----------------------
__global__ void stem_im2col_kernel(const float* __restrict__ x, const float* __restrict__ noise,
                                   const float* __restrict__ sa, const float* __restrict__ sb, float* __restrict__ xt,
                                   h16* __restrict__ cols, int B, int Cin, int H, int W) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * H * W;
  for (long long pix = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; pix < total;
       pix += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(pix / (H * W));
    const int rem = static_cast<int>(pix - static_cast<long long>(b) * H * W);
    const int y = rem / W, xx = rem - y * W;
    h16 row[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) row[i] = f2h(0.f);
    for (int ci = 0; ci < Cin; ++ci) {
      const long long plane = (static_cast<long long>(b) * Cin + ci) * H * W;
      for (int t = 0; t < 9; ++t) {
        const int yy = y + t / 3 - 1, xs = xx + t % 3 - 1;
        float v = 0.f;
        if (yy >= 0 && yy < H && xs >= 0 && xs < W) {
          v = x[plane + yy * W + xs];
          if (noise) v = sa[b] * v + sb[b] * noise[plane + yy * W + xs];
        }
        if (ci * 9 + t < 64) row[ci * 9 + t] = f2h(v);
        if (t == 4 && xt) xt[plane + y * W + xx] = v;
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(cols + pix * 64);
    const uint4* src = reinterpret_cast<const uint4*>(row);
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = src[i];
  }
}
int launch_stem_im2col(const float* x, const float* noise, const float* sa, const float* sb, float* xt, h16* cols,
                       int B, int Cin, int H, int W, cudaStream_t s) {
  if (Cin * 9 > 64) return static_cast<int>(cudaErrorInvalidValue);
  launch_k(stem_im2col_kernel, grid_for(static_cast<long long>(B) * H * W, 128), 128, 0, s, x, noise, sa, sb, xt, cols, B,
                                                                                      Cin, H, W);
  return static_cast<int>(cudaGetLastError());
}

__global__ void stem_col2im_kernel(const h16* __restrict__ dcols, float* __restrict__ dx,
                                   const float* __restrict__ gscale, int B, int Cin, int H, int W) {
  pdl_enter();
  const float ginv = gscale ? gscale[1] : 1.f;
  const long long total = static_cast<long long>(B) * Cin * H * W;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int xx = static_cast<int>(idx % W);
    const int y = static_cast<int>((idx / W) % H);
    const int ci = static_cast<int>((idx / (static_cast<long long>(W) * H)) % Cin);
    const int b = static_cast<int>(idx / (static_cast<long long>(W) * H * Cin));
    float a = 0.f;
    // input pixel (y, xx) was read by output pixel (y - dy, xx - dx) through tap (dy, dx)
    for (int t = 0; t < 9; ++t) {
      const int oy = y - (t / 3 - 1), ox = xx - (t % 3 - 1);
      if (oy >= 0 && oy < H && ox >= 0 && ox < W)
        a += h2f(dcols[((static_cast<long long>(b) * H + oy) * W + ox) * 64 + ci * 9 + t]);
    }
    dx[idx] = a * ginv;
  }
}
int launch_stem_col2im(const h16* dcols, float* dx, const float* gscale, int B, int Cin, int H, int W, cudaStream_t s) {
  launch_k(stem_col2im_kernel, grid_for(static_cast<long long>(B) * Cin * H * W, 256), 256, 0, s, dcols, dx, gscale, B, Cin, H, W);
  return static_cast<int>(cudaGetLastError());
}

// ---- layout converters for the 3-channel head ------------------------------------------------------------------------
__global__ void nhwc_to_nchw_kernel(const float* __restrict__ y, float* __restrict__ out, int B, int C, int HW) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * C * HW;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int p = static_cast<int>(idx % HW);
    const int c = static_cast<int>((idx / HW) % C);
    const long long b = idx / (static_cast<long long>(HW) * C);
    out[idx] = y[(b * HW + p) * C + c];
  }
}
int launch_nhwc_to_nchw(const float* y, float* out, int B, int C, int H, int W, cudaStream_t s) {
  launch_k(nhwc_to_nchw_kernel, grid_for(static_cast<long long>(B) * C * H * W, 256), 256, 0, s, y, out, B, C, H * W);
  return static_cast<int>(cudaGetLastError());
}
// gscale[0] = power of two that puts max|x| at 2^-4..2^-3, gscale[1] = its inverse (fp16 build; 1 / 1 in the bf16 build).
// Many blocks (the single-block version took 33 us for the 0.2 M gradients of a B = 256 step, on the critical path of the
// backward): block maxima meet in scratch[0] (float bits through atomicMax), the last block to arrive (ticket in scratch[1])
// publishes the scale.  scratch = gscale + 2, zeroed by the launcher.
__global__ void __launch_bounds__(256) grad_scale_kernel(const float* __restrict__ x, long long n, float* __restrict__ gscale) {
  pdl_enter();
  unsigned* scratch = reinterpret_cast<unsigned*>(gscale + 2);
  float m = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    m = fmaxf(m, fabsf(x[i]));
  m = warp_max(m);
  __shared__ float red[8];
  __shared__ bool last;
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t = fmaxf(t, red[w]);
    if (!(t == t) || t > 3.0e38f) t = __uint_as_float(0x7f800000u);     // NaN / inf anywhere: reported as inf below
    atomicMax(scratch, __float_as_uint(t));
    __threadfence();
    last = atomicAdd(scratch + 1, 1u) == gridDim.x - 1;
    if (last) {
      __threadfence();
      t = __uint_as_float(atomicExch(scratch, 0u));
      scratch[1] = 0u;
      float sc = 1.f;
#ifndef EDB_ACT_BF16
      if (t > 0.f && isfinite(t)) {
        int e;
        frexpf(t, &e);                          // t = f * 2^e, f in [0.5, 1)
        sc = ldexpf(1.f, -3 - e);               // sc * t in [2^-4, 2^-3)
      }
#endif
      gscale[0] = sc;
      gscale[1] = 1.f / sc;
    }
  }
}
int launch_grad_scale(const float* x, long long n, float* gscale, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(gscale + 2, 0, 2 * sizeof(unsigned), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  launch_k(grad_scale_kernel, grid_for(n, 256 * 4, 148), 256, 0, s, x, n, gscale);
  return static_cast<int>(cudaGetLastError());
}
__global__ void nchw_to_nhwc_h16_kernel(const float* __restrict__ x, h16* __restrict__ out,
                                         const float* __restrict__ gscale, int B, int C, int HW, int Cpad) {
  pdl_enter();
  const float gsc = gscale ? gscale[0] : 1.f;
  const long long total = static_cast<long long>(B) * HW * Cpad;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(idx % Cpad);
    const int p = static_cast<int>((idx / Cpad) % HW);
    const long long b = idx / (static_cast<long long>(Cpad) * HW);
    out[idx] = f2h(c < C ? x[(b * C + c) * HW + p] * gsc : 0.f);
  }
}
int launch_nchw_to_nhwc_h16(const float* x, h16* out, const float* gscale, int B, int C, int H, int W, int Cpad,
                            cudaStream_t s) {
  launch_k(nchw_to_nhwc_h16_kernel, grid_for(static_cast<long long>(B) * H * W * Cpad, 256), 256, 0, s, x, out, gscale, B, C,
                                                                                                  H * W, Cpad);
  return static_cast<int>(cudaGetLastError());
}

// ---- L1 loss + its gradient (ddpm_enc.py:297-301,1198-1213) -----------------------------------------------------------
__global__ void l1_loss_kernel(const float* __restrict__ e, const float* __restrict__ n, float* __restrict__ loss,
                               float* __restrict__ de, long long cnt, float inv) {
  pdl_enter();
  float a = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < cnt;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float d = e[i] - n[i];
    a += fabsf(d);
    if (de) de[i] = d > 0.f ? inv : (d < 0.f ? -inv : 0.f);
  }
  a = warp_sum(a);
  __shared__ float red[32];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) atomicAdd(loss, t * inv);
  }
}
int launch_l1_loss(const float* eps_hat, const float* noise, float* loss, float* deps, long long n, cudaStream_t s) {
  launch_k(l1_loss_kernel, grid_for(n, 256, 148 * 4), 256, 0, s, eps_hat, noise, loss, deps, n, 1.0f / static_cast<float>(n));
  return static_cast<int>(cudaGetLastError());
}

// ---- DDIM update (ddim.py:197-206), same operation order in fp32 ------------------------------------------------------
__global__ void ddim_update_kernel(const float* __restrict__ x, const float* __restrict__ e, const float* __restrict__ nz,
                                   const float* __restrict__ coef, float temperature, float* __restrict__ x_prev,
                                   float* __restrict__ pred_x0, long long n) {
  pdl_enter();
  const float a_t = coef[0], a_prev = coef[1], sigma = coef[2], somat = coef[3];
  const float sqrt_at = sqrtf(a_t), sqrt_ap = sqrtf(a_prev);
  const float dir = sqrtf(__fsub_rn(__fsub_rn(1.f, a_prev), __fmul_rn(sigma, sigma)));
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float et = e[i];
    const float p0 = __fdiv_rn(x[i] - __fmul_rn(somat, et), sqrt_at);
    const float d = __fmul_rn(dir, et);
    const float noise = nz ? __fmul_rn(__fmul_rn(sigma, nz[i]), temperature) : 0.f;
    x_prev[i] = __fadd_rn(__fadd_rn(__fmul_rn(sqrt_ap, p0), d), noise);
    if (pred_x0) pred_x0[i] = p0;
  }
}
int launch_ddim_update(const float* x, const float* e_t, const float* noise, const float* coef, float temperature,
                       float* x_prev, float* pred_x0, long long n, cudaStream_t s) {
  launch_k(ddim_update_kernel, grid_for(n, 256), 256, 0, s, x, e_t, noise, coef, temperature, x_prev, pred_x0, n);
  return static_cast<int>(cudaGetLastError());
}

// ---- DDIM loop state on the device, so that ONE captured step graph can be replayed S times (ddim.py:141-166) ----------
// cursor[0] = loop iteration i; writes ts[b] = t_table[i] (ddim.py:143), idx[0] = i, then advances the cursor.
__global__ void ddim_select_kernel(const float* __restrict__ t_table, int* __restrict__ cursor, float* __restrict__ ts,
                                   int* __restrict__ idx, int B) {
  pdl_enter();
  const int i = *cursor;
  __syncthreads();
  const float t = t_table[i];
  for (int b = threadIdx.x; b < B; b += blockDim.x) ts[b] = t;
  if (threadIdx.x == 0) {
    *idx = i;
    *cursor = i + 1;
  }
}
int launch_ddim_select(const float* t_table, int* cursor, float* ts, int* idx, int B, cudaStream_t s) {
  launch_k(ddim_select_kernel, 1, 256, 0, s, t_table, cursor, ts, idx, B);
  return static_cast<int>(cudaGetLastError());
}
__global__ void ddim_update_indexed_kernel(const float* __restrict__ x, const float* __restrict__ e,
                                           const float* __restrict__ noise_table, const float* __restrict__ coef_table,
                                           const int* __restrict__ idx, float temperature, float* __restrict__ x_prev,
                                           float* __restrict__ pred_x0, long long n) {
  pdl_enter();
  const int it = *idx;
  const float* coef = coef_table + 4 * it;
  const float* nz = noise_table ? noise_table + static_cast<long long>(it) * n : nullptr;
  const float a_t = coef[0], a_prev = coef[1], sigma = coef[2], somat = coef[3];
  const float sqrt_at = sqrtf(a_t), sqrt_ap = sqrtf(a_prev);
  const float dir = sqrtf(__fsub_rn(__fsub_rn(1.f, a_prev), __fmul_rn(sigma, sigma)));
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float et = e[i];
    const float p0 = __fdiv_rn(x[i] - __fmul_rn(somat, et), sqrt_at);
    const float d = __fmul_rn(dir, et);
    const float noise = nz ? __fmul_rn(__fmul_rn(sigma, nz[i]), temperature) : 0.f;
    x_prev[i] = __fadd_rn(__fadd_rn(__fmul_rn(sqrt_ap, p0), d), noise);
    if (pred_x0) pred_x0[i] = p0;
  }
}
int launch_ddim_update_indexed(const float* x, const float* e_t, const float* noise_table, const float* coef_table,
                               const int* idx, float temperature, float* x_prev, float* pred_x0, long long n,
                               cudaStream_t s) {
  launch_k(ddim_update_indexed_kernel, grid_for(n, 256), 256, 0, s, x, e_t, noise_table, coef_table, idx, temperature, x_prev,
                                                             pred_x0, n);
  return static_cast<int>(cudaGetLastError());
}

// ---- weight packing: fp32 reference-layout parameters -> h16 GEMM operands (and fused fp32 biases) -------------------
// One block = one unit of one table entry (PackEntry::blk0 = first block of the entry):
//   kinds 0 / 1 (weights): a tile of 32 rows x kPackCols source columns (cin-tile * taps) goes through shared memory, so
//     the fp32 reads are contiguous runs of the reference tensor and the 16-bit writes are contiguous runs of the
//     operand, for the forward (tap-major K) and the transposed layout alike;
//   other kinds (biases, stem): 256 consecutive source elements, element-wise.
static constexpr int kPackCols = 288;      // 32 input channels x 9 taps, or 288 input channels x 1 tap
__global__ void __launch_bounds__(256) pack_kernel(const PackEntry* __restrict__ tab, int n_entries,
                                                   h16* __restrict__ abf, float* __restrict__ af32) {
  pdl_enter();
  __shared__ float tile[32][kPackCols + 1];
  __shared__ int s_entry;
  if (threadIdx.x == 0) {
    int lo = 0, hi = n_entries - 1;   // last entry with blk0 <= blockIdx.x
    const int bid = static_cast<int>(blockIdx.x);
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tab[mid].blk0 <= bid) lo = mid; else hi = mid - 1;
    }
    s_entry = lo;
  }
  __syncthreads();
  const PackEntry en = tab[s_entry];
  const int tb = static_cast<int>(blockIdx.x) - en.blk0;
  if (en.kind == 0 || en.kind == 1) {
    const int ci_tile = kPackCols / en.taps;
    const int cts = (en.cin + ci_tile - 1) / ci_tile;
    const int rt = tb / cts, cit = tb - rt * cts;
    const int r0 = rt * 32, ci0 = cit * ci_tile;
    const int nr = min(32, en.rows - r0), nci = min(ci_tile, en.cin - ci0);
    const int w = nci * en.taps;                          // source columns of this tile
    const long long row_len = static_cast<long long>(en.cin) * en.taps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // warps walk rows / (channel, tap) pairs, lanes walk the contiguous index: no per-element integer division
    // (all kPackCols / 32 loads of a row are issued before the first shared-memory store: one load in flight per thread
    // left the kernel at 1.1 TB/s)
    for (int rr = warp; rr < nr; rr += 8) {
      const float* src = en.src + (r0 + rr) * row_len + static_cast<long long>(ci0) * en.taps;
      float v[kPackCols / 32];
#pragma unroll
      for (int i = 0; i < kPackCols / 32; ++i) {
        const int k = lane + 32 * i;
        v[i] = k < w ? __ldg(src + k) : 0.f;
      }
#pragma unroll
      for (int i = 0; i < kPackCols / 32; ++i) {
        const int k = lane + 32 * i;
        if (k < w) tile[rr][k] = v[i];
      }
    }
    __syncthreads();
    h16* dst = abf + en.dst_off;
    if (en.kind == 0) {         // dst[row0 + r][col0 + tap * cin_pad + ci]: ci fastest
      // two channels per lane: 4-byte stores (the kernel was issue-bound on 2-byte stores); every offset below is even
      for (int rr = warp; rr < nr; rr += 8) {
        h16* drow = dst + static_cast<long long>(en.row0 + r0 + rr) * en.dst_ld + en.col0 + ci0;
        for (int tap = 0; tap < en.taps; ++tap)
          for (int c = 2 * lane; c < nci; c += 64) {
            const float v0 = tile[rr][c * en.taps + tap];
            if (c + 1 < nci) *reinterpret_cast<h162*>(drow + tap * en.cin_pad + c) = f22h2(v0, tile[rr][(c + 1) * en.taps + tap]);
            else drow[tap * en.cin_pad + c] = f2h(v0);
          }
      }
    } else {                    // dst[row0 + ci][col0 + tap * rows_pad + r]: r fastest
      const int pairs = nci * en.taps;
      // a warp writes two (channel, tap) rows per step, 16 lanes x two output rows (one 4-byte store) each; an odd last
      // row is paired with the operand's zero padding (rows_pad >= r0 + 32)
      const int half = lane >> 4, l2 = (lane & 15) * 2;
      for (int pr = 2 * warp + half; pr < pairs; pr += 16) {
        const int c = pr / en.taps, tap = pr - c * en.taps;
        h16* drow = dst + static_cast<long long>(en.row0 + ci0 + c) * en.dst_ld + en.col0 + tap * en.rows_pad + r0;
        if (l2 < nr) *reinterpret_cast<h162*>(drow + l2) = f22h2(tile[l2][pr], l2 + 1 < nr ? tile[l2 + 1][pr] : 0.f);
      }
    }
    return;
  }
  const long long n = static_cast<long long>(en.rows) * en.cin * en.taps;
  const long long e = static_cast<long long>(tb) * 256 + threadIdx.x;
  if (e >= n) return;
  const float v = en.src[e];
  if (en.kind == 2) {
    af32[en.dst_off + en.col0 + e] = v;
  } else if (en.kind == 3) {
    af32[en.dst_off + en.col0 + e] = v + en.src2[e];
  } else {
    const int per_row = en.cin * en.taps;
    const int r = static_cast<int>(e / per_row);
    const int rem = static_cast<int>(e - static_cast<long long>(r) * per_row);
    long long row, col;
    if (en.kind == 4) {        // stem: flat reference order
      row = en.row0 + r;
      col = en.col0 + rem;
    } else if (en.kind == 5) { // stem transposed
      row = en.row0 + rem;
      col = en.col0 + r;
    } else {                   // 6: transposed, tap-major rows: dst[row0 + tap * cin_pad + ci][col0 + r]  (src [r][ci][tap])
      const int ci = rem / en.taps, tap = rem - ci * en.taps;
      row = en.row0 + static_cast<long long>(tap) * en.cin_pad + ci;
      col = en.col0 + r;
    }
    abf[en.dst_off + row * en.dst_ld + col] = f2h(v);
  }
}
int pack_entry_blocks(const PackEntry& e) {
  if (e.kind == 0 || e.kind == 1) {
    const int ci_tile = kPackCols / e.taps;
    return ((e.rows + 31) / 32) * ((e.cin + ci_tile - 1) / ci_tile);
  }
  const long long n = static_cast<long long>(e.rows) * e.cin * e.taps;
  return static_cast<int>((n + 255) / 256);
}
int launch_pack(const PackEntry* table_dev, int n_entries, int n_blocks, h16* arena_h16, float* arena_f32,
                cudaStream_t s) {
  if (n_blocks <= 0) return 0;
  launch_k(pack_kernel, n_blocks, 256, 0, s, table_dev, n_entries, arena_h16, arena_f32);
  return static_cast<int>(cudaGetLastError());
}

// ---- fused AdamW (+EMA) over flat fp32 buffers (torch.optim.AdamW semantics: decoupled decay) --------------------------
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, float* __restrict__ ema, long long n, float lr, float b1, float b2,
                             float eps, float wd, float bc1, float bc2, float ema_decay) {
  pdl_enter();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float w = p[i];
    const float gi = g[i];
    w *= (1.f - lr * wd);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / sqrtf(bc2) + eps;
    w -= (lr / bc1) * (mi / denom);
    p[i] = w;
    if (ema) ema[i] -= (1.f - ema_decay) * (ema[i] - w);   // ema.py:38-42
  }
}
int launch_adamw(float* p, const float* g, float* m, float* v, float* ema, long long n, float lr, float beta1,
                 float beta2, float eps, float wd, float bc1, float bc2, float ema_decay, cudaStream_t s) {
  launch_k(adamw_kernel, grid_for(n, 256, 148 * 16), 256, 0, s, p, g, m, v, ema, n, lr, beta1, beta2, eps, wd, bc1, bc2,
                                                          ema_decay);
  return static_cast<int>(cudaGetLastError());
}

// same update with the step-dependent scalars read from device memory, so a captured graph of the optimiser step can be
// replayed: hyper = {lr, 1-beta1^t, 1-beta2^t, ema_decay}
__global__ void adamw_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, float* __restrict__ ema, long long n,
                                 const float* __restrict__ hyper, float b1, float b2, float eps, float wd,
                                 float grad_scale) {
  pdl_enter();
  const float lr = hyper[0], bc1 = hyper[1], bc2 = hyper[2], ema_decay = hyper[3];
  const float rs = rsqrtf(bc2), step = lr / bc1, keep = 1.f - lr * wd, om = 1.f - ema_decay;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float w = p[i] * keep;
    const float gi = g[i] * grad_scale;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    w -= step * (mi / (sqrtf(vi) * rs + eps));
    p[i] = w;
    if (ema) ema[i] -= om * (ema[i] - w);
  }
}
// 16-byte form of the same update (same arithmetic per element, so results are bit-identical): five float4 loads in flight
// per thread instead of five scalars -- the scalar kernel ran at 58 % of the copy bandwidth.  n4 = n / 4 vectors; the
// (< 4) tail elements are done by one thread.
__device__ __forceinline__ void adamw_one(float& w, float gi, float& mi, float& vi, float* e, float keep, float b1, float b2,
                                          float step, float rs, float eps, float om, float grad_scale) {
  w *= keep;
  gi *= grad_scale;
  mi = b1 * mi + (1.f - b1) * gi;
  vi = b2 * vi + (1.f - b2) * gi * gi;
  w -= step * (mi / (sqrtf(vi) * rs + eps));
  if (e) *e -= om * (*e - w);
}
__global__ void __launch_bounds__(256) adamw_dev_vec_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                            float* __restrict__ m, float* __restrict__ v,
                                                            float* __restrict__ ema, long long n,
                                                            const float* __restrict__ hyper, float b1, float b2, float eps,
                                                            float wd, float grad_scale) {
  pdl_enter();
  const float lr = hyper[0], bc1 = hyper[1], bc2 = hyper[2], ema_decay = hyper[3];
  const float rs = rsqrtf(bc2), step = lr / bc1, keep = 1.f - lr * wd, om = 1.f - ema_decay;
  const long long n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  float4* e4 = reinterpret_cast<float4*>(ema);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 w = p4[i], gi = g4[i], mi = m4[i], vi = v4[i], ei = ema ? e4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    adamw_one(w.x, gi.x, mi.x, vi.x, ema ? &ei.x : nullptr, keep, b1, b2, step, rs, eps, om, grad_scale);
    adamw_one(w.y, gi.y, mi.y, vi.y, ema ? &ei.y : nullptr, keep, b1, b2, step, rs, eps, om, grad_scale);
    adamw_one(w.z, gi.z, mi.z, vi.z, ema ? &ei.z : nullptr, keep, b1, b2, step, rs, eps, om, grad_scale);
    adamw_one(w.w, gi.w, mi.w, vi.w, ema ? &ei.w : nullptr, keep, b1, b2, step, rs, eps, om, grad_scale);
    p4[i] = w;
    m4[i] = mi;
    v4[i] = vi;
    if (ema) e4[i] = ei;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long i = n4 << 2; i < n; ++i) {
      float w = p[i], mi = m[i], vi = v[i];
      adamw_one(w, g[i], mi, vi, ema ? ema + i : nullptr, keep, b1, b2, step, rs, eps, om, grad_scale);
      p[i] = w;
      m[i] = mi;
      v[i] = vi;
    }
}
// Step counter and the step-dependent scalars stay on the device (one thread): a captured optimiser graph advances them
// itself, so a host that runs many replays ahead of the GPU can neither tear nor pre-date them.
__global__ void adamw_hyper_advance_kernel(int* __restrict__ step, const float* __restrict__ lr, float* __restrict__ hyper,
                                           float b1, float b2, float ema_decay_max) {
  pdl_enter();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int t = step[0] + 1;
  step[0] = t;
  hyper[0] = lr[0];
  hyper[1] = static_cast<float>(1.0 - pow(static_cast<double>(b1), static_cast<double>(t)));
  hyper[2] = static_cast<float>(1.0 - pow(static_cast<double>(b2), static_cast<double>(t)));
  const float warm = (1.f + t) / (10.f + t);          // LitEma warm-up, ema.py:31-33 (num_updates counted from 1)
  hyper[3] = warm < ema_decay_max ? warm : ema_decay_max;
}
int launch_adamw_hyper_advance(int* step, const float* lr, float* hyper, float beta1, float beta2, float ema_decay_max,
                               cudaStream_t s) {
  launch_k(adamw_hyper_advance_kernel, 1, 32, 0, s, step, lr, hyper, beta1, beta2, ema_decay_max);
  return static_cast<int>(cudaGetLastError());
}

int launch_adamw_dev(float* p, const float* g, float* m, float* v, float* ema, long long n, const float* hyper,
                     float beta1, float beta2, float eps, float wd, float grad_scale, cudaStream_t s) {
  const uintptr_t all = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                        reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(ema);
  if ((all & 15) == 0 && n >= 4)
    launch_k(adamw_dev_vec_kernel, grid_for(n >> 2, 256, 148 * 8), 256, 0, s, p, g, m, v, ema, n, hyper, beta1, beta2, eps, wd,
                                                                        grad_scale);
  else
    launch_k(adamw_dev_kernel, grid_for(n, 256, 148 * 16), 256, 0, s, p, g, m, v, ema, n, hyper, beta1, beta2, eps, wd,
                                                                grad_scale);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
