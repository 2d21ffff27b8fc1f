// Kernels of the first-stage VQ-f4 encoder that the UNet path does not have (ldm/modules/diffusionmodules/model.py):
//   * Downsample with conv (:63-79): pad right / bottom by one, 3x3 stride 2 -> explicit tap-major im2col + plain GEMM;
//   * AttnBlock (:153-204): ONE head of width C over the H*W tokens of a sample, softmax(q k^T C^-1/2) v;
//   * quant_conv (autoencoder.py:49,315) folded into the NHWC -> NCHW conversion of the 3-channel encoder output.
#include "common.cuh"
#include "kernels.cuh"

namespace edb {
namespace {

#ifdef EDB_ACT_BF16
#define EDB_MMA_TV "bf16"
#else
#define EDB_MMA_TV "f16"
#endif

__device__ __forceinline__ void vmma_k16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." EDB_MMA_TV "." EDB_MMA_TV ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t vpack2(float a, float b) {
  const h162 v = f22h2(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float vquad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float vquad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// cols[(b, oy, ox)][tap * C64 + c] = x[b][2 oy + ky][2 ox + kx][c]  (0 beyond the bottom / right edge and for c >= C)
// thread = (output pixel, kernel row, 8-channel chunk): the row's three taps are three independent 16-byte loads in flight
__global__ void __launch_bounds__(256) vq_down_im2col_kernel(const h16* __restrict__ x, h16* __restrict__ cols, int B, int H,
                                                             int C, int C64) {
  pdl_enter();
  const int Ho = H / 2, cv = C64 / 8;
  const long long total = static_cast<long long>(B) * Ho * Ho * 3 * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int ky = static_cast<int>(r % 3);
    r /= 3;                                      // output pixel (b, oy, ox)
    const int ox = static_cast<int>(r % Ho);
    const long long bo = r / Ho;
    const int oy = static_cast<int>(bo % Ho);
    const long long b = bo / Ho;
    const int y = 2 * oy + ky, x0 = 2 * ox;
    const bool ok = y < H && c8 * 8 < C;
    const h16* src = x + ((b * H + y) * H + x0) * C + c8 * 8;
    uint4 v[3];
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      v[kx] = make_uint4(0u, 0u, 0u, 0u);
      if (ok && x0 + kx < H) v[kx] = *reinterpret_cast<const uint4*>(src + static_cast<long long>(kx) * C);
    }
    h16* dst = cols + (r * 9 + ky * 3) * C64 + c8 * 8;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) *reinterpret_cast<uint4*>(dst + static_cast<long long>(kx) * C64) = v[kx];
  }
}

// qkv [B*N, 3D] (q | k | v of the single head), out [B*N, D].  CTA = (sample, block of 16 * warps queries); the sample's K and
// V rows sit in shared memory (pitch D + 8: conflict-free fragment reads); a warp owns 16 queries: Q fragments straight from
// global memory, two passes over the keys (row max; then exp / sum / P V with P as the A fragment, V through ldmatrix.trans).
template <int D>
__global__ void __launch_bounds__(256) vq_attn_kernel(const h16* __restrict__ qkv, h16* __restrict__ out, int N, float scale_log2) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  constexpr int P = D + 8;
  h16* Ks = reinterpret_cast<h16*>(smraw);     // [N][P]
  h16* Vs = Ks + N * P;                        // [N][P]
  const int b = blockIdx.y;
  const long long rowb = static_cast<long long>(b) * N;
  for (int idx = threadIdx.x; idx < N * (D / 8); idx += blockDim.x) {
    const int j = idx / (D / 8), c = idx - j * (D / 8);
    const h16* src = qkv + (rowb + j) * 3 * D + c * 8;
    *reinterpret_cast<uint4*>(Ks + j * P + c * 8) = *reinterpret_cast<const uint4*>(src + D);
    *reinterpret_cast<uint4*>(Vs + j * P + c * 8) = *reinterpret_cast<const uint4*>(src + 2 * D);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int q0 = (blockIdx.x * (blockDim.x >> 5) + warp) * 16;
  uint32_t qa[D / 4];
  if (q0 < N) {
    const h16* p0 = qkv + (rowb + q0 + g) * 3 * D + 2 * t;
    const h16* p1 = p0 + 8LL * 3 * D;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      qa[4 * ks] = __ldg(reinterpret_cast<const unsigned int*>(p0 + 16 * ks));
      qa[4 * ks + 1] = __ldg(reinterpret_cast<const unsigned int*>(p1 + 16 * ks));
      qa[4 * ks + 2] = __ldg(reinterpret_cast<const unsigned int*>(p0 + 16 * ks + 8));
      qa[4 * ks + 3] = __ldg(reinterpret_cast<const unsigned int*>(p1 + 16 * ks + 8));
    }
  }
  __syncthreads();
  if (q0 >= N) return;
  auto scores8 = [&](float* c, int n0) {      // c[16 x 8] = Q K^T for keys n0 .. n0+7
    c[0] = c[1] = c[2] = c[3] = 0.f;
    const h16* kp = Ks + (n0 + g) * P + 2 * t;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks)
      vmma_k16(c, qa + 4 * ks, *reinterpret_cast<const uint32_t*>(kp + 16 * ks), *reinterpret_cast<const uint32_t*>(kp + 16 * ks + 8));
  };
  float m0 = -INFINITY, m1 = -INFINITY;
  for (int n0 = 0; n0 < N; n0 += 8) {
    float c[4];
    scores8(c, n0);
    m0 = fmaxf(m0, fmaxf(c[0], c[1]));
    m1 = fmaxf(m1, fmaxf(c[2], c[3]));
  }
  m0 = vquad_max(m0) * scale_log2;
  m1 = vquad_max(m1) * scale_log2;
  float l0 = 0.f, l1 = 0.f;
  float o[D / 8][4];
#pragma unroll
  for (int nd = 0; nd < D / 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f;
  // ldmatrix row addresses of this lane: matrices 0 / 1 = keys +0..7 / +8..15 at column d0, matrices 2 / 3 = the same at d0 + 8
  const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lcol = (lane >> 4) * 8;
  for (int kc = 0; kc < N; kc += 16) {
    float c[2][4];
    scores8(c[0], kc);
    scores8(c[1], kc + 8);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      c[i][0] = ex2_approx(c[i][0] * scale_log2 - m0);
      c[i][1] = ex2_approx(c[i][1] * scale_log2 - m0);
      c[i][2] = ex2_approx(c[i][2] * scale_log2 - m1);
      c[i][3] = ex2_approx(c[i][3] * scale_log2 - m1);
      l0 += c[i][0] + c[i][1];
      l1 += c[i][2] + c[i][3];
    }
    const uint32_t pa[4] = {vpack2(c[0][0], c[0][1]), vpack2(c[0][2], c[0][3]), vpack2(c[1][0], c[1][1]), vpack2(c[1][2], c[1][3])};
#pragma unroll
    for (int nd = 0; nd < D / 8; nd += 2) {
      uint32_t b0, b1, b2, b3;      // B fragments (k = key, n = d) of two adjacent 8-wide d tiles: transposed 8x8 blocks of V
      const uint32_t addr = smem_u32(Vs + (kc + lrow) * P + nd * 8 + lcol);
      asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                   : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3)
                   : "r"(addr));
      vmma_k16(o[nd], pa, b0, b1);
      vmma_k16(o[nd + 1], pa, b2, b3);
    }
  }
  const float i0 = 1.f / vquad_sum(l0), i1 = 1.f / vquad_sum(l1);
  h16* o0 = out + (rowb + q0 + g) * D + 2 * t;
  h16* o1 = o0 + 8LL * D;
#pragma unroll
  for (int nd = 0; nd < D / 8; ++nd) {
    *reinterpret_cast<uint32_t*>(o0 + 8 * nd) = vpack2(o[nd][0] * i0, o[nd][1] * i0);
    *reinterpret_cast<uint32_t*>(o1 + 8 * nd) = vpack2(o[nd][2] * i1, o[nd][3] * i1);
  }
}

// z[b][e][pix] = bq[e] + sum_c Wq[e][c] * h[(b, pix)][c]     (h fp32 NHWC [M, zc], z fp32 NCHW)
__global__ void vq_tail_kernel(const float* __restrict__ h, const float* __restrict__ wq, const float* __restrict__ bq,
                               float* __restrict__ z, int B, int HW, int zc, int ed) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * ed * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int pix = static_cast<int>(i % HW);
    const long long r = i / HW;
    const int e = static_cast<int>(r % ed), b = static_cast<int>(r / ed);
    float a = bq[e];
    const float* hp = h + (static_cast<long long>(b) * HW + pix) * zc;
    for (int c = 0; c < zc; ++c) a = fmaf(wq[e * zc + c], hp[c], a);
    z[i] = a;
  }
}

// nearest-neighbour x2 (Upsample.forward, model.py:53-54): y[b][2y+i][2x+j][c] = x[b][y][x][c]
__global__ void vq_upsample2x_kernel(const h16* __restrict__ x, h16* __restrict__ y, int B, int H, int C) {
  pdl_enter();
  const int cv = C / 8, Ho = 2 * H;
  const long long total = static_cast<long long>(B) * Ho * Ho * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long r = i / cv;
    const int ox = static_cast<int>(r % Ho);
    r /= Ho;
    const int oy = static_cast<int>(r % Ho), b = static_cast<int>(r / Ho);
    *reinterpret_cast<uint4*>(y + i * 8) =
        *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(b) * H + (oy >> 1)) * H + (ox >> 1)) * C + c8 * 8);
  }
}

// Front of VQModelInterface.decode (autoencoder.py:340-366), one thread per latent pixel, NCHW fp32 in and out:
//   q = quantize ? nearest code of h (taming VectorQuantizer2 forward value: argmin_k |h|^2 + |e_k|^2 - 2 h.e_k, first
//       minimum wins like torch.argmin) : h
//   out[c] = b[c] + sum_e W[c][e] q[e] + sum_n W[c][ed + n] s[b][n]      (the concatenated disentangled channels are constant
//       over the grid; s == NULL is the reference's zeros branch)
__global__ void vq_post_quant_kernel(const float* __restrict__ h, const float* __restrict__ sdis,
                                     const float* __restrict__ codebook, int n_embed, const float* __restrict__ w,
                                     const float* __restrict__ bias, float* __restrict__ out, int B, int HW, int ed, int dd,
                                     int zc) {
  pdl_enter();
  const long long total = static_cast<long long>(B) * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int pix = static_cast<int>(i % HW), b = static_cast<int>(i / HW);
    float q[8];
    float zz = 0.f;
    for (int e = 0; e < ed; ++e) {
      q[e] = h[(static_cast<long long>(b) * ed + e) * HW + pix];
      zz += q[e] * q[e];
    }
    if (codebook) {
      float best = INFINITY;
      int bi = 0;
      for (int k = 0; k < n_embed; ++k) {
        float ee = 0.f, ze = 0.f;
        for (int e = 0; e < ed; ++e) {
          const float c = codebook[k * ed + e];
          ee += c * c;
          ze += q[e] * c;
        }
        const float d = zz + ee - 2.f * ze;
        if (d < best) {
          best = d;
          bi = k;
        }
      }
      for (int e = 0; e < ed; ++e) q[e] = codebook[bi * ed + e];
    }
    for (int c = 0; c < zc; ++c) {
      float a = bias[c];
      const float* wr = w + c * (ed + dd);
      for (int e = 0; e < ed; ++e) a = fmaf(wr[e], q[e], a);
      if (sdis)
        for (int n = 0; n < dd; ++n) a = fmaf(wr[ed + n], sdis[static_cast<long long>(b) * dd + n], a);
      out[(static_cast<long long>(b) * zc + c) * HW + pix] = a;
    }
  }
}

inline int vq_grid(long long n, int threads) {
  const long long b = (n + threads - 1) / threads;
  return static_cast<int>(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

}  // namespace

int launch_vq_down_im2col(const h16* x, h16* cols, int B, int H, int C, cudaStream_t s) {
  const int C64 = (C + 63) / 64 * 64;
  if ((H & 1) || (C & 7)) return static_cast<int>(cudaErrorInvalidValue);
  const long long total = static_cast<long long>(B) * (H / 2) * (H / 2) * 3 * (C64 / 8);
  launch_k(vq_down_im2col_kernel, vq_grid(total, 256), 256, 0, s, x, cols, B, H, C, C64);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
static int vq_attn_t(const h16* qkv, h16* out, int B, int N, cudaStream_t s) {
  const size_t smem = static_cast<size_t>(2) * N * (D + 8) * sizeof(h16);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(vq_attn_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int warps = N / 16 < 8 ? N / 16 : 8;
  launch_k(vq_attn_kernel<D>, dim3((N / 16 + warps - 1) / warps, B), warps * 32, smem, s, qkv, out, N,
           rsqrtf(static_cast<float>(D)) * 1.4426950408889634f);
  return static_cast<int>(cudaGetLastError());
}
int launch_vq_attn(const h16* qkv, h16* out, int B, int N, int D, cudaStream_t s) {
  if (N < 16 || (N & 15) || N > 256) return static_cast<int>(cudaErrorInvalidValue);
  switch (D) {
    case 64: return vq_attn_t<64>(qkv, out, B, N, s);
    case 128: return vq_attn_t<128>(qkv, out, B, N, s);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

int launch_vq_tail(const float* h, const float* wq, const float* bq, float* z, int B, int HW, int zc, int ed, cudaStream_t s) {
  launch_k(vq_tail_kernel, vq_grid(static_cast<long long>(B) * ed * HW, 256), 256, 0, s, h, wq, bq, z, B, HW, zc, ed);
  return static_cast<int>(cudaGetLastError());
}

int launch_vq_upsample2x(const h16* x, h16* y, int B, int H, int C, cudaStream_t s) {
  if (C & 7) return static_cast<int>(cudaErrorInvalidValue);
  launch_k(vq_upsample2x_kernel, vq_grid(static_cast<long long>(B) * 4 * H * H * (C / 8), 256), 256, 0, s, x, y, B, H, C);
  return static_cast<int>(cudaGetLastError());
}

int launch_vq_post_quant(const float* h, const float* sdis, const float* codebook, int n_embed, const float* w,
                         const float* bias, float* out, int B, int HW, int ed, int dd, int zc, cudaStream_t s) {
  if (ed > 8) return static_cast<int>(cudaErrorInvalidValue);
  launch_k(vq_post_quant_kernel, vq_grid(static_cast<long long>(B) * HW, 128), 128, 0, s, h, sdis, codebook, n_embed, w, bias,
           out, B, HW, ed, dd, zc);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
