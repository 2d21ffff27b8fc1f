// GroupNorm (+FiLM +SiLU +2x resample) and LayerNorm, forward and backward.  HBM-bound kernels: one read of the
// input slab from HBM (the second pass hits L1/L2), one write of the result.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.cuh"

namespace edb {

// ---------------------------------------------------------------------------------------------------------------
// GroupNorm: one CTA per sample; thread = (8-channel chunk, pixel lane)
// ---------------------------------------------------------------------------------------------------------------
struct GnThread {
  int C, nv, HW, chunk, prow, rp, c0;
  const h16* src;  // sample base of the source holding this thread's chunk
  int ld, cofs;
};

__device__ __forceinline__ GnThread gn_thread(const GnParams& p, int b) {
  GnThread t;
  t.C = p.C0 + p.C1;
  t.nv = t.C >> 3;
  t.HW = p.H * p.W;
  t.chunk = threadIdx.x % t.nv;
  t.prow = threadIdx.x / t.nv;
  t.rp = blockDim.x / t.nv;
  t.c0 = t.chunk * 8;
  if (t.c0 < p.C0) {
    t.src = p.x0 + static_cast<long long>(b) * t.HW * p.C0;
    t.ld = p.C0;
    t.cofs = t.c0;
  } else {
    t.src = p.x1 + static_cast<long long>(b) * t.HW * p.C1;
    t.ld = p.C1;
    t.cofs = t.c0 - p.C0;
  }
  return t;
}

// Per-channel block reduction of two 8-channel register vectors (one per thread) without shared-memory atomics:
// lanes of a warp that hold the same chunk are combined by shuffles (when nv divides 32), each remaining holder
// writes one row of `part` ([rows][2][C]), and the rows are summed per channel into out0 / out1.
// Must be called by every thread of the block (contains __syncthreads).
__device__ __forceinline__ void gn_channel_reduce(const GnThread& t, bool active, float* s, float* q, float* part,
                                                  float* out0, float* out1) {
  const int C = t.C;
  int rows;
  if ((32 % t.nv) == 0) {   // blockDim.x == 512: full warps, lane % nv == chunk
    for (int o = t.nv; o < 32; o <<= 1) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
        q[k] += __shfl_xor_sync(0xffffffffu, q[k], o);
      }
    }
    rows = blockDim.x >> 5;
    if ((threadIdx.x & 31) < t.nv) {
      float* pr = part + (threadIdx.x >> 5) * 2 * C;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        pr[t.c0 + k] = s[k];
        pr[C + t.c0 + k] = q[k];
      }
    }
  } else {
    rows = t.rp;
    if (active) {
      float* pr = part + t.prow * 2 * C;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        pr[t.c0 + k] = s[k];
        pr[C + t.c0 + k] = q[k];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    float a = 0.f;
    for (int r = 0; r < rows; ++r) a += part[r * 2 * C + c];
    if (c < C) out0[c] = a; else out1[c - C] = a;
  }
  __syncthreads();
}

__device__ __forceinline__ void load8(const h16* p, float* f) { unpack8(*reinterpret_cast<const uint4*>(p), f); }
__device__ __forceinline__ void store8(h16* p, const float* f) { *reinterpret_cast<uint4*>(p) = pack8(f); }

// Parameter vectors (and the FiLM row / saved statistics of sample b) do not depend on the previous kernel and are HBM-cold
// at every launch: their lines go to L2 before the programmatic-launch wait.
__device__ __forceinline__ void gn_prefetch_params(const GnParams& p, int b, bool with_stats) {
  const int C = p.C0 + p.C1;
  const int i = threadIdx.x * 32;            // one 128-byte line of floats per thread
  if (i < C) {
    l2_prefetch_line(p.gamma + i);
    l2_prefetch_line(p.beta + i);
  }
  if (p.film && i < 2 * C) l2_prefetch_line(p.film + static_cast<long long>(b) * p.ldfilm + p.film_off + i);
  if (with_stats && i < 64) l2_prefetch_line(p.stats + b * 64 + i);
}

__global__ void __launch_bounds__(512) gn_fwd_kernel(const GnParams p) {
  pdl_launch_dependents();
  gn_prefetch_params(p, blockIdx.x, false);
  pdl_wait();
  extern __shared__ float sm[];
  const int b = blockIdx.x;
  const GnThread t = gn_thread(p, b);
  const int C = t.C;
  float* chsum = sm;
  float* chsq = sm + C;
  float* cA = sm + 2 * C;
  float* cB = sm + 3 * C;
  float* fS = sm + 4 * C;     // FiLM 1 + scale
  float* fT = sm + 5 * C;     // FiLM shift
  float* gstat = sm + 6 * C;  // [64]
  float* part = gstat + 64;   // [rows][2][C] partial sums, rows <= 32
  const bool active = t.prow < t.rp;  // threads beyond nv*rp idle (blockDim is a multiple of nv, so all active)
  // affine / FiLM parameters are fetched now, under the statistics pass, instead of after it (one dependent global
  // round trip less on the CTA's critical path); the barriers inside gn_channel_reduce publish them
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    cA[c] = p.gamma[c];
    cB[c] = p.beta[c];
    if (p.film) {
      const float* fr = p.film + static_cast<long long>(b) * p.ldfilm + p.film_off;
      fS[c] = 1.f + fr[c];
      fT[c] = fr[C + c];
    }
  }
  {
    float s[8], q[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] = q[k] = 0.f;
    if (active) {
      // 4 independent 16-byte loads in flight per thread (the trip count is a run-time value: no compiler unrolling)
      for (int pix0 = t.prow; pix0 < t.HW; pix0 += 4 * t.rp) {
        uint4 raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int pix = pix0 + u * t.rp;
          if (pix < t.HW) raw[u] = *reinterpret_cast<const uint4*>(t.src + static_cast<long long>(pix) * t.ld + t.cofs);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (pix0 + u * t.rp < t.HW) {
            float f[8];
            unpack8(raw[u], f);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              s[k] += f[k];
              q[k] += f[k] * f[k];
            }
          }
        }
      }
    }
    gn_channel_reduce(t, active, s, q, part, chsum, chsq);
  }
  const int cpg = C / 32;
  if (threadIdx.x < 32) {
    const int g = threadIdx.x;
    float S = 0.f, Q = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) {
      S += chsum[c];
      Q += chsq[c];
    }
    const float inv_n = 1.0f / (static_cast<float>(cpg) * t.HW);
    const float mean = S * inv_n;
    const float var = fmaxf(Q * inv_n - mean * mean, 0.f);
    const float rstd = rsqrtf(var + p.eps);
    gstat[g] = mean;
    gstat[32 + g] = rstd;
    p.stats[(b * 32 + g) * 2] = mean;
    p.stats[(b * 32 + g) * 2 + 1] = rstd;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    float ga = cA[c] * gstat[32 + g];
    float be = cB[c] - gstat[g] * ga;
    if (p.film) {
      const float sc = fS[c];
      ga *= sc;
      be = be * sc + fT[c];
    }
    cA[c] = ga;
    cB[c] = be;
  }
  __syncthreads();
  if (!active) return;
  float a[8], bb[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = cA[t.c0 + k];
    bb[k] = cB[t.c0 + k];
  }
  const int W = p.W, H = p.H;
  if (p.resample == RESAMPLE_NONE) {
    h16* dst = p.act + static_cast<long long>(b) * t.HW * C + t.c0;
    for (int pix0 = t.prow; pix0 < t.HW; pix0 += 4 * t.rp) {
      uint4 raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int pix = pix0 + u * t.rp;
        if (pix < t.HW) raw[u] = *reinterpret_cast<const uint4*>(t.src + static_cast<long long>(pix) * t.ld + t.cofs);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int pix = pix0 + u * t.rp;
        if (pix < t.HW) {
          float f[8];
          unpack8(raw[u], f);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            float y = a[k] * f[k] + bb[k];
            f[k] = p.silu ? silu_f(y) : y;
          }
          store8(dst + static_cast<long long>(pix) * C, f);
        }
      }
    }
  } else if (p.resample == RESAMPLE_DOWN) {
    const int Wo = W / 2, HWo = (H / 2) * Wo;
    h16* dst = p.act + static_cast<long long>(b) * HWo * C + t.c0;
    h16* dxs = p.xs ? p.xs + static_cast<long long>(b) * HWo * C + t.c0 : nullptr;
    for (int op = t.prow; op < HWo; op += t.rp) {
      const int oy = op / Wo, ox = op - oy * Wo;
      float ya[8], xa[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) ya[k] = xa[k] = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int pix = (2 * oy + (i >> 1)) * W + 2 * ox + (i & 1);
        float f[8];
        load8(t.src + static_cast<long long>(pix) * t.ld + t.cofs, f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float y = a[k] * f[k] + bb[k];
          ya[k] += p.silu ? silu_f(y) : y;
          xa[k] += f[k];
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        ya[k] *= 0.25f;
        xa[k] *= 0.25f;
      }
      store8(dst + static_cast<long long>(op) * C, ya);
      if (dxs) store8(dxs + static_cast<long long>(op) * C, xa);
    }
  } else {  // RESAMPLE_UP
    const int Wo = W * 2, HWo = 4 * t.HW;
    h16* dst = p.act + static_cast<long long>(b) * HWo * C + t.c0;
    h16* dxs = p.xs ? p.xs + static_cast<long long>(b) * HWo * C + t.c0 : nullptr;
    for (int pix = t.prow; pix < t.HW; pix += t.rp) {
      const int y = pix / W, x = pix - y * W;
      float f[8], ya[8];
      load8(t.src + static_cast<long long>(pix) * t.ld + t.cofs, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float v = a[k] * f[k] + bb[k];
        ya[k] = p.silu ? silu_f(v) : v;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const long long op = static_cast<long long>(2 * y + (i >> 1)) * Wo + 2 * x + (i & 1);
        store8(dst + op * C, ya);
        if (dxs) store8(dxs + op * C, f);
      }
    }
  }
}

// gradient arriving at input pixel (y, x) of a tensor that was resampled AFTER this point
__device__ __forceinline__ void gather_resampled(const h16* g, int C, int resample, int y, int x, int W, float* d) {
  if (resample == RESAMPLE_NONE) {
    load8(g + static_cast<long long>(y * W + x) * C, d);
  } else if (resample == RESAMPLE_DOWN) {
    load8(g + static_cast<long long>((y >> 1) * (W >> 1) + (x >> 1)) * C, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) d[k] *= 0.25f;
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) d[k] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float f[8];
      load8(g + (static_cast<long long>(2 * y + (i >> 1)) * (2 * W) + 2 * x + (i & 1)) * C, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) d[k] += f[k];
    }
  }
}

__global__ void __launch_bounds__(512) gn_bwd_kernel(const GnParams p) {
  pdl_launch_dependents();
  // the sample's forward input was saved by the forward: pull it into L2 while the previous backward kernel drains
  if (threadIdx.x == 0) {
    const long long hw = static_cast<long long>(p.H) * p.W;
    l2_prefetch_bulk(p.x0 + blockIdx.x * hw * p.C0, static_cast<uint32_t>(hw * p.C0 * 2));
    if (p.x1) l2_prefetch_bulk(p.x1 + blockIdx.x * hw * p.C1, static_cast<uint32_t>(hw * p.C1 * 2));
  }
  gn_prefetch_params(p, blockIdx.x, true);
  pdl_wait();
  extern __shared__ float sm[];
  const int b = blockIdx.x;
  const GnThread t = gn_thread(p, b);
  const int C = t.C;
  float* cA = sm;           // u = cA*x + cB   (pre-activation)
  float* cB = sm + C;
  float* cM = sm + 2 * C;   // mean per channel
  float* cR = sm + 3 * C;   // rstd per channel
  float* S1 = sm + 4 * C;   // sum dyp          -> later k1
  float* S2 = sm + 5 * C;   // sum dyp * xhat   -> later k2
  float* K3 = sm + 6 * C;
  float* gAB = sm + 7 * C;  // [64] group sums
  float* part = gAB + 64;   // [rows][2][C] partial sums, rows <= 32
  uint4* stash = (p.stash_off && p.resample == RESAMPLE_NONE) ? reinterpret_cast<uint4*>(sm + p.stash_off) : nullptr;
  const int cpg = C / 32;
  const int W = p.W;
  int HWo = t.HW;
  if (p.resample == RESAMPLE_DOWN) HWo = t.HW / 4;
  if (p.resample == RESAMPLE_UP) HWo = t.HW * 4;
  const h16* dact = p.dact + static_cast<long long>(b) * HWo * C + t.c0;
  // first batch of (x, dact) rows: issued before the parameter staging below so both round trips overlap
  uint4 rx[4], rd[4];
  auto load_batch = [&](int pix0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pix = pix0 + u * t.rp;
      if (pix < t.HW) {
        rx[u] = *reinterpret_cast<const uint4*>(t.src + static_cast<long long>(pix) * t.ld + t.cofs);
        rd[u] = *reinterpret_cast<const uint4*>(dact + static_cast<long long>(pix) * C);
      }
    }
  };
  if (p.resample == RESAMPLE_NONE) load_batch(t.prow);
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const float mean = p.stats[(b * 32 + g) * 2], rstd = p.stats[(b * 32 + g) * 2 + 1];
    float ga = p.gamma[c] * rstd;
    float be = p.beta[c] - mean * ga;
    if (p.film) {
      const float* fr = p.film + static_cast<long long>(b) * p.ldfilm + p.film_off;
      const float sc = 1.f + fr[c];
      ga *= sc;
      be = be * sc + fr[C + c];
    }
    cA[c] = ga;
    cB[c] = be;
    cM[c] = mean;
    cR[c] = rstd;
  }
  __syncthreads();
  float a[8], bb[8], mu[8], rs[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = cA[t.c0 + k];
    bb[k] = cB[t.c0 + k];
    mu[k] = cM[t.c0 + k];
    rs[k] = cR[t.c0 + k];
  }
  {
    float s1[8], s2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
    if (p.resample == RESAMPLE_NONE) {
      // 4 x (x, dact) loads in flight per thread; the first batch is already on its way
      int vi = 0;      // running vector index of this thread (the second pass walks the same pixels in the same order)
      for (int pix0 = t.prow; pix0 < t.HW; pix0 += 4 * t.rp) {
        if (pix0 != t.prow) load_batch(pix0);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (pix0 + u * t.rp < t.HW) {
            float f[8], d[8];
            unpack8(rx[u], f);
            unpack8(rd[u], d);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float dyp = p.silu ? d[k] * silu_grad_f(a[k] * f[k] + bb[k]) : d[k];
              s1[k] += dyp;
              s2[k] += dyp * (f[k] - mu[k]) * rs[k];
              d[k] = dyp;
            }
            // dy' = dy * silu'(u) is needed again by the second pass: keep it (16-bit, thread-private slot, conflict-free)
            // instead of re-reading dy and re-evaluating the SiLU gradient (35 % of the kernel's stall samples, ncu)
            if (stash) stash[(vi + u) * blockDim.x + threadIdx.x] = pack8(d);
          }
        }
        vi += 4;
      }
    } else {
      for (int pix = t.prow; pix < t.HW; pix += t.rp) {
        const int y = pix / W, x = pix - y * W;
        float f[8], d[8];
        load8(t.src + static_cast<long long>(pix) * t.ld + t.cofs, f);
        gather_resampled(dact, C, p.resample, y, x, W, d);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float dyp = p.silu ? d[k] * silu_grad_f(a[k] * f[k] + bb[k]) : d[k];
          s1[k] += dyp;
          s2[k] += dyp * (f[k] - mu[k]) * rs[k];
        }
      }
    }
    gn_channel_reduce(t, t.prow < t.rp, s1, s2, part, S1, S2);
  }
  // per-channel parameter gradients; turn S1/S2 into gsc-weighted values for the group sums
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float s1 = S1[c], s2 = S2[c];
    const float gam = p.gamma[c], bet = p.beta[c];
    float sc = 1.f;
    if (p.film) {
      sc = 1.f + p.film[static_cast<long long>(b) * p.ldfilm + p.film_off + c];
      if (p.dfilm) {
        h16* df = p.dfilm + static_cast<long long>(b) * p.lddfilm + p.film_off;
        df[c] = f2h(gam * s2 + bet * s1);   // d scale
        df[C + c] = f2h(s1);                // d shift
      }
    }
    const float ginv = p.gscale ? p.gscale[1] : 1.f;   // activation gradients carry the backward's range scale
    if (p.dgamma) atomicAdd(&p.dgamma[c], sc * s2 * ginv);
    if (p.dbeta) atomicAdd(&p.dbeta[c], sc * s1 * ginv);
    const float gsc = gam * sc;
    S1[c] = gsc * s1;
    S2[c] = gsc * s2;
    K3[c] = gsc;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int g = threadIdx.x;
    float A = 0.f, Bv = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) {
      A += S1[c];
      Bv += S2[c];
    }
    const float inv_n = 1.0f / (static_cast<float>(cpg) * t.HW);
    gAB[g] = A * inv_n;
    gAB[32 + g] = Bv * inv_n;
  }
  __syncthreads();
  float k1[8], k2[8], k3[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = t.c0 + k, g = c / cpg;
    k1[k] = rs[k] * K3[c];
    k2[k] = rs[k] * gAB[g];
    k3[k] = rs[k] * gAB[32 + g];
  }
  h16* dxp;
  int acc;
  if (t.c0 < p.C0) {
    dxp = p.dx0 + static_cast<long long>(b) * t.HW * p.C0 + t.cofs;
    acc = p.acc0;
  } else {
    dxp = p.dx1 + static_cast<long long>(b) * t.HW * p.C1 + t.cofs;
    acc = p.acc1;
  }
  const h16* dxs = p.dxs ? p.dxs + static_cast<long long>(b) * HWo * C + t.c0 : nullptr;
  if (p.resample == RESAMPLE_NONE) {
    // 2 pixels per step: up to 8 independent 16-byte loads in flight (x, dact, dxs, previous dx)
    int vi = 0;
    for (int pix0 = t.prow; pix0 < t.HW; pix0 += 2 * t.rp, vi += 2) {
      uint4 rx[2], rd[2], re[2], ra[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int pix = pix0 + u * t.rp;
        if (pix < t.HW) {
          rx[u] = *reinterpret_cast<const uint4*>(t.src + static_cast<long long>(pix) * t.ld + t.cofs);
          if (stash) rd[u] = stash[(vi + u) * blockDim.x + threadIdx.x];
          else rd[u] = *reinterpret_cast<const uint4*>(dact + static_cast<long long>(pix) * C);
          if (dxs) re[u] = *reinterpret_cast<const uint4*>(dxs + static_cast<long long>(pix) * C);
          if (acc) ra[u] = *reinterpret_cast<const uint4*>(dxp + static_cast<long long>(pix) * t.ld);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int pix = pix0 + u * t.rp;
        if (pix < t.HW) {
          float f[8], d[8], o[8];
          unpack8(rx[u], f);
          unpack8(rd[u], d);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float dyp = (p.silu && !stash) ? d[k] * silu_grad_f(a[k] * f[k] + bb[k]) : d[k];
            const float xh = (f[k] - mu[k]) * rs[k];
            o[k] = k1[k] * dyp - k2[k] - xh * k3[k];
          }
          if (dxs) {
            float e[8];
            unpack8(re[u], e);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] += e[k];
          }
          if (acc) {
            float e[8];
            unpack8(ra[u], e);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] += e[k];
          }
          store8(dxp + static_cast<long long>(pix) * t.ld, o);
        }
      }
    }
    return;
  }
  for (int pix = t.prow; pix < t.HW; pix += t.rp) {
    const int y = pix / W, x = pix - y * W;
    float f[8], d[8], o[8];
    load8(t.src + static_cast<long long>(pix) * t.ld + t.cofs, f);
    gather_resampled(dact, C, p.resample, y, x, W, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float dyp = p.silu ? d[k] * silu_grad_f(a[k] * f[k] + bb[k]) : d[k];
      const float xh = (f[k] - mu[k]) * rs[k];
      o[k] = k1[k] * dyp - k2[k] - xh * k3[k];
    }
    if (dxs) {
      float e[8];
      gather_resampled(dxs, C, p.resample, y, x, W, e);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] += e[k];
    }
    h16* q = dxp + static_cast<long long>(pix) * t.ld;
    if (acc) {
      float e[8];
      load8(q, e);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] += e[k];
    }
    store8(q, o);
  }
}

static int gn_threads(int C, int max_threads = 512) {
  const int nv = C / 8;
  int rp = max_threads / nv;
  if (rp < 1) rp = 1;
  return nv * rp;
}

static size_t gn_part_floats(int C) {
  const int nv = C / 8;
  int rp = 512 / nv;
  if (rp < 1) rp = 1;
  const int rows = (32 % nv) == 0 ? 16 : rp;
  return static_cast<size_t>(rows) * 2 * C;
}

int launch_gn_fwd(const GnParams& p, int B, cudaStream_t s) {
  const int C = p.C0 + p.C1;
  if (C % 32 != 0 || (p.C0 % 8) != 0 || C > 4096) return static_cast<int>(cudaErrorInvalidValue);
  const size_t smem = (6 * C + 64 + gn_part_floats(C)) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(gn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  launch_k(gn_fwd_kernel, B, gn_threads(C), smem, s, p);
  return static_cast<int>(cudaGetLastError());
}

int launch_gn_bwd(const GnParams& p_in, int B, cudaStream_t s) {
  GnParams p = p_in;
  const int C = p.C0 + p.C1;
  if (C % 32 != 0 || (p.C0 % 8) != 0 || C > 4096) return static_cast<int>(cudaErrorInvalidValue);
  size_t smem = (7 * C + 64 + gn_part_floats(C)) * sizeof(float);
  // dy' stash of the un-resampled SiLU path: one 16-byte slot per (thread, vector of the thread); used when the sample's
  // 16-bit slab fits next to the reduction scratch with two CTAs per SM (up to 96 KB: 16x16x192 is the largest level-0 slab)
  p.stash_off = 0;
  static const bool no_stash = getenv("ENCDIFF_GN_NO_STASH") != nullptr;   // A/B switch (parity experiments)
  if (p.resample == RESAMPLE_NONE && p.silu && !no_stash) {
    const int threads = gn_threads(C, 256), nv = C / 8, rp = threads / nv;
    const size_t vecs = static_cast<size_t>((p.H * p.W + rp - 1) / rp + 3) * threads;   // + the 4-deep batch rounding
    const size_t off = (smem + 15) & ~size_t(15);
    if (vecs * 16 <= 96 * 1024 && off + vecs * 16 <= 110 * 1024) {
      p.stash_off = static_cast<int>(off / sizeof(float));
      smem = off + vecs * 16;
    }
  }
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(gn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  // 256-thread CTAs: at 128 registers a 512-thread CTA owns a whole SM and the B = 256 samples would need two waves
  launch_k(gn_bwd_kernel, B, gn_threads(C, 256), smem, s, p);
  return static_cast<int>(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------------------------
// LayerNorm: LPR lanes per row (8/16/32), each lane owns up to CPL 8-channel chunks
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float group_sum(float v, int lpr) {
  for (int o = lpr >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int CPL>
__global__ void __launch_bounds__(256) ln_fwd_kernel(const h16* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, float eps, h16* __restrict__ y,
                                                     float* __restrict__ stats, int M, int C, int lpr) {
  pdl_launch_dependents();
  if (static_cast<int>(threadIdx.x) * 32 < C) {      // gamma / beta: HBM-cold parameters, independent of the previous kernel
    l2_prefetch_line(gamma + threadIdx.x * 32);
    l2_prefetch_line(beta + threadIdx.x * 32);
  }
  pdl_wait();
  const int nchunks = C >> 3;
  const int rows_per_block = blockDim.x / lpr;
  const int lane_in_row = threadIdx.x % lpr;
  const long long row = static_cast<long long>(blockIdx.x) * rows_per_block + threadIdx.x / lpr;
  const bool ok = row < M;
  float v[CPL][8];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = lane_in_row + j * lpr;
    if (ok && ch < nchunks) {
      unpack8(*reinterpret_cast<const uint4*>(x + row * C + ch * 8), v[j]);
#pragma unroll
      for (int k = 0; k < 8; ++k) s += v[j][k];
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[j][k] = 0.f;
    }
  }
  const float inv_c = 1.0f / static_cast<float>(C);
  const float mean = group_sum(s, lpr) * inv_c;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = lane_in_row + j * lpr;
    if (ch < nchunks) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float d = v[j][k] - mean;
        q += d * d;
      }
    }
  }
  const float rstd = rsqrtf(group_sum(q, lpr) * inv_c + eps);
  if (!ok) return;
  if (lane_in_row == 0) {
    stats[row * 2] = mean;
    stats[row * 2 + 1] = rstd;
  }
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = lane_in_row + j * lpr;
    if (ch < nchunks) {
      float o[8], ga[8], be[8];
      *reinterpret_cast<float4*>(ga) = __ldg(reinterpret_cast<const float4*>(gamma + ch * 8));
      *reinterpret_cast<float4*>(ga + 4) = __ldg(reinterpret_cast<const float4*>(gamma + ch * 8 + 4));
      *reinterpret_cast<float4*>(be) = __ldg(reinterpret_cast<const float4*>(beta + ch * 8));
      *reinterpret_cast<float4*>(be + 4) = __ldg(reinterpret_cast<const float4*>(beta + ch * 8 + 4));
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = fmaf((v[j][k] - mean) * rstd, ga[k], be[k]);
      *reinterpret_cast<uint4*>(y + row * C + ch * 8) = pack8(o);
    }
  }
}

// Persistent over rows; the next row's loads are issued before the current row is reduced (two rows in flight per
// thread).  Parameter gradients: per-thread register sums -> one shared-memory row per row slot (plain stores) -> column
// sums -> one global atomic per channel and CTA.  (The first version combined the row slots with shared-memory float
// atomics: ncu showed 44 % of the stall cycles on that CAS loop and the kernel at 1.2 TB/s.)
template <int CPL>
__global__ void __launch_bounds__(256, CPL == 1 ? 3 : 1) ln_bwd_kernel(const h16* __restrict__ x, const float* __restrict__ stats,
                                                     const float* __restrict__ gamma, const h16* __restrict__ dy,
                                                     const h16* __restrict__ dres, h16* __restrict__ dx,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                     const float* __restrict__ gscale, int M, int C, int lpr) {
  pdl_launch_dependents();
  if (static_cast<int>(threadIdx.x) >= 32 && (static_cast<int>(threadIdx.x) - 32) * 32 < C) l2_prefetch_line(gamma + (threadIdx.x - 32) * 32);
  {   // x and its statistics were saved by the forward: the block's first rows go to L2 before the wait
    // (thread k takes the block's k-th sweep of the persistent row loop)
    const long long rpb = blockDim.x / lpr;
    const long long r0 = (static_cast<long long>(blockIdx.x) + static_cast<long long>(threadIdx.x) * gridDim.x) * rpb;
    if (threadIdx.x < 16 && r0 < M) {
      const long long nr = min(rpb, M - r0);
      l2_prefetch_bulk(x + r0 * C, static_cast<uint32_t>(nr * C * 2));
      if (nr >= 2) l2_prefetch_bulk(stats + r0 * 2, static_cast<uint32_t>((nr * 8) & ~15LL));   // never past the array
    }
  }
  pdl_wait();
  extern __shared__ float sm[];  // [rows_per_block][2*C] partial sums of dgamma | dbeta
  const int nchunks = C >> 3;
  const int rows_per_block = blockDim.x / lpr;
  const int lane_in_row = threadIdx.x % lpr;
  const int slot = threadIdx.x / lpr;
  const float inv_c = 1.0f / static_cast<float>(C);
  float ag[CPL][8], ab[CPL][8];
#pragma unroll
  for (int j = 0; j < CPL; ++j)
#pragma unroll
    for (int k = 0; k < 8; ++k) ag[j][k] = ab[j][k] = 0.f;
  // this thread's channels are the same for every row: gamma lives in registers
  float gam[CPL][8];
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = lane_in_row + j * lpr;
    if (ch < nchunks) {
      *reinterpret_cast<float4*>(&gam[j][0]) = __ldg(reinterpret_cast<const float4*>(gamma + ch * 8));
      *reinterpret_cast<float4*>(&gam[j][4]) = __ldg(reinterpret_cast<const float4*>(gamma + ch * 8 + 4));
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) gam[j][k] = 0.f;
    }
  }
  const long long stride = static_cast<long long>(gridDim.x) * rows_per_block;
  long long row = static_cast<long long>(blockIdx.x) * rows_per_block + slot;
  uint4 rx[CPL], rdy[CPL], rres[CPL];
  float2 st = make_float2(0.f, 0.f);
  auto issue = [&](long long r) {
    if (r < M) {
#pragma unroll
      for (int j = 0; j < CPL; ++j) {
        const int ch = lane_in_row + j * lpr;
        if (ch < nchunks) {
          rx[j] = *reinterpret_cast<const uint4*>(x + r * C + ch * 8);
          rdy[j] = *reinterpret_cast<const uint4*>(dy + r * C + ch * 8);
          if (dres) rres[j] = *reinterpret_cast<const uint4*>(dres + r * C + ch * 8);
        }
      }
      st = *reinterpret_cast<const float2*>(stats + r * 2);
    }
  };
  issue(row);
  for (; row - slot < M; row += stride) {     // uniform trip count over the block (row - slot is the block's first row)
    const bool ok = row < M;
    float xh[CPL][8], g[CPL][8], rr[CPL][8];
    float s1 = 0.f, s2 = 0.f;
    const float mean = st.x, rstd = st.y;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      const int ch = lane_in_row + j * lpr;
      if (ok && ch < nchunks) {
        float xv[8], dv[8];
        unpack8(rx[j], xv);
        unpack8(rdy[j], dv);
        if (dres) unpack8(rres[j], rr[j]);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[j][k] = (xv[k] - mean) * rstd;
          ag[j][k] = fmaf(dv[k], xh[j][k], ag[j][k]);
          ab[j][k] += dv[k];
          g[j][k] = dv[k] * gam[j][k];
          s1 += g[j][k];
          s2 = fmaf(g[j][k], xh[j][k], s2);
        }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) xh[j][k] = g[j][k] = 0.f;
      }
    }
    issue(row + stride);                        // next row's loads fly during the reduction and the store below
    s1 = group_sum(s1, lpr) * inv_c;
    s2 = group_sum(s2, lpr) * inv_c;
    if (ok) {
#pragma unroll
      for (int j = 0; j < CPL; ++j) {
        const int ch = lane_in_row + j * lpr;
        if (ch < nchunks) {
          float o[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = rstd * (g[j][k] - s1 - xh[j][k] * s2);
          if (dres) {
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] += rr[j][k];
          }
          *reinterpret_cast<uint4*>(dx + row * C + ch * 8) = pack8(o);
        }
      }
    }
  }
  if (dgamma == nullptr) return;
  float* mine = sm + static_cast<size_t>(slot) * 2 * C;
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = lane_in_row + j * lpr;
    if (ch < nchunks) {
      *reinterpret_cast<float4*>(mine + ch * 8) = *reinterpret_cast<float4*>(&ag[j][0]);
      *reinterpret_cast<float4*>(mine + ch * 8 + 4) = *reinterpret_cast<float4*>(&ag[j][4]);
      *reinterpret_cast<float4*>(mine + C + ch * 8) = *reinterpret_cast<float4*>(&ab[j][0]);
      *reinterpret_cast<float4*>(mine + C + ch * 8 + 4) = *reinterpret_cast<float4*>(&ab[j][4]);
    }
  }
  __syncthreads();
  const float ginv = gscale ? gscale[1] : 1.f;
  // columns [0, C) = dgamma, [C, 2C) = dbeta; when 2C < blockDim the row slots are split over blockDim / 2C thread groups
  const int cols = 2 * C;
  const int groups = blockDim.x >= cols ? blockDim.x / cols : 1;
  for (int c0 = 0; c0 < cols; c0 += blockDim.x) {
    const int c = c0 + static_cast<int>(threadIdx.x) % (groups > 1 ? cols : blockDim.x);
    const int grp = groups > 1 ? threadIdx.x / cols : 0;
    if (c < cols && grp < groups) {
      float a = 0.f;
      for (int r = grp; r < rows_per_block; r += groups) a += sm[static_cast<size_t>(r) * cols + c];
      atomicAdd(c < C ? &dgamma[c] : &dbeta[c - C], a * ginv);
    }
  }
}

static void ln_shape(int C, int* lpr, int* cpl) {
  const int nchunks = C / 8;
  int l = 8;
  if (nchunks >= 32) l = 32;
  else if (nchunks >= 16) l = 16;
  *lpr = l;
  *cpl = (nchunks + l - 1) / l;
}

int launch_ln_fwd(const h16* x, const float* gamma, const float* beta, float eps, h16* y, float* stats, int M, int C,
                  cudaStream_t s) {
  if (C % 8 != 0 || C < 64 || C > 1024) return static_cast<int>(cudaErrorInvalidValue);
  int lpr, cpl;
  ln_shape(C, &lpr, &cpl);
  const int rows_per_block = 256 / lpr;
  const int grid = (M + rows_per_block - 1) / rows_per_block;
  switch (cpl) {
    case 1: launch_k(ln_fwd_kernel<1>, grid, 256, 0, s, x, gamma, beta, eps, y, stats, M, C, lpr); break;
    case 2: launch_k(ln_fwd_kernel<2>, grid, 256, 0, s, x, gamma, beta, eps, y, stats, M, C, lpr); break;
    case 3: launch_k(ln_fwd_kernel<3>, grid, 256, 0, s, x, gamma, beta, eps, y, stats, M, C, lpr); break;
    case 4: launch_k(ln_fwd_kernel<4>, grid, 256, 0, s, x, gamma, beta, eps, y, stats, M, C, lpr); break;
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
  return static_cast<int>(cudaGetLastError());
}

int launch_ln_bwd(const h16* x, const float* stats, const float* gamma, const h16* dy, const h16* dres, h16* dx,
                  float* dgamma, float* dbeta, const float* gscale, int M, int C, cudaStream_t s) {
  if (C % 8 != 0 || C < 64 || C > 1024) return static_cast<int>(cudaErrorInvalidValue);
  int lpr, cpl;
  ln_shape(C, &lpr, &cpl);
  const int rows_per_block = 256 / lpr;
  int grid = (M + rows_per_block - 1) / rows_per_block;
  // persistent grid: 2 CTAs per SM measured best on B200 (148 / 222 / 296 / 592 / 1184 CTAs: 0.68 / 0.62 / 0.60 / 0.74 /
  // 0.80 ms per step for the 48 launches); ENCDIFF_LN_BWD_CTAS overrides for tuning
  static const int cap = getenv("ENCDIFF_LN_BWD_CTAS") ? atoi(getenv("ENCDIFF_LN_BWD_CTAS")) : 148 * 2;
  if (grid > cap) grid = cap;
  const size_t smem = static_cast<size_t>(rows_per_block) * 2 * C * sizeof(float);   // <= 64 KB (C = 1024)
  if (smem > 48 * 1024) {
    static bool configured = false;
    if (!configured) {
      cudaError_t e = cudaFuncSetAttribute(ln_bwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(ln_bwd_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
      if (e != cudaSuccess) return static_cast<int>(e);
      configured = true;
    }
  }
  switch (cpl) {
    case 1: launch_k(ln_bwd_kernel<1>, grid, 256, smem, s, x, stats, gamma, dy, dres, dx, dgamma, dbeta, gscale, M, C, lpr); break;
    case 2: launch_k(ln_bwd_kernel<2>, grid, 256, smem, s, x, stats, gamma, dy, dres, dx, dgamma, dbeta, gscale, M, C, lpr); break;
    case 3: launch_k(ln_bwd_kernel<3>, grid, 256, smem, s, x, stats, gamma, dy, dres, dx, dgamma, dbeta, gscale, M, C, lpr); break;
    case 4: launch_k(ln_bwd_kernel<4>, grid, 256, smem, s, x, stats, gamma, dy, dres, dx, dgamma, dbeta, gscale, M, C, lpr); break;
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
