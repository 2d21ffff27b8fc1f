// Concept-token MLPs of the EncDiff encoder: `Encoder4.warp` / the tail of `Encoder4.forward`
// (ldm/modules/diffusionmodules/openaimodel_enc.py:1017-1041).  The reference runs latent_unit (20) independent
// 3-layer MLPs  Linear(1,64) - ELU - Linear(64,128) - ELU - Linear(128,context_dim)  in a Python loop, one scalar code
// per unit in, one context_dim-wide token out, and concatenates the tokens: ~100 tiny launches forward, ~300 with the
// backward.  Here: ONE launch forward, ONE launch backward (all units, all samples), fp32 on the CUDA cores -- the
// whole thing is 10 k MACs per (sample, unit), so it is launch-latency work, not tensor-core work.
//
//   grid = (latent_unit, sample chunks); a CTA stages its unit's weights in shared memory once and walks its samples
//   in groups of 8.  Backward recomputes the two hidden layers (cheaper than storing them), keeps the weight-gradient
//   partial sums of its chunk in registers and adds them to the fp32 gradients with one atomic per element and CTA.
#include "common.cuh"
#include "kernels.cuh"

namespace edb {

static constexpr int kH1 = 64, kH2 = 128, kG = 8, kCT = 128;   // hidden widths (fixed by the reference), group, threads

__device__ __forceinline__ float elu_f(float z) { return z > 0.f ? z : expm1f(z); }

struct ConceptSmem {
  float w2t[kH1][kH2];       // [k][j] = W2[j][k]   (layer-2 forward: lanes walk j)
  float w2[kH2][kH1 + 1];    // [j][k]              (backward dh1: lanes walk k); padded against bank conflicts
  float w3t[kH2][32];        // [j][o] = W3[o][j]
  float w1[kH1], b1[kH1], b2[kH2], b3[32];
  float x[kG];
  float h1t[kH1][kG];        // [k][s]
  float z1t[kH1][kG];
  float h2[kG][kH2 + 1];
  float dz2t[kH2][kG];       // [j][s]
  float dy[kG][32];
  float red[kCT];
};

template <bool BWD>
__device__ __forceinline__ void concept_stage_weights(ConceptSmem& sm, const ConceptParams& p, int u, int Dc) {
  const float* W2 = p.w2 + static_cast<long long>(u) * kH2 * kH1;
  for (int i = threadIdx.x; i < kH2 * kH1; i += kCT) {
    const int j = i / kH1, k = i - j * kH1;
    const float v = W2[i];
    sm.w2t[k][j] = v;
    if (BWD) sm.w2[j][k] = v;
  }
  const float* W3 = p.w3 + static_cast<long long>(u) * Dc * kH2;
  for (int i = threadIdx.x; i < Dc * kH2; i += kCT) {
    const int o = i / kH2, j = i - o * kH2;
    sm.w3t[j][o] = W3[i];
  }
  for (int i = threadIdx.x; i < kH1; i += kCT) {
    sm.w1[i] = p.w1[u * kH1 + i];
    sm.b1[i] = p.b1[u * kH1 + i];
  }
  for (int i = threadIdx.x; i < kH2; i += kCT) sm.b2[i] = p.b2[u * kH2 + i];
  for (int i = threadIdx.x; i < Dc; i += kCT) sm.b3[i] = p.b3[u * Dc + i];
}

// hidden layers of one group of kG samples (s0 .. s0+kG-1 of this CTA's chunk; samples past `send` are zero-padded)
__device__ __forceinline__ void concept_hidden(ConceptSmem& sm, const ConceptParams& p, int u, int s0, int send,
                                               float* z2_out /* [kG] of thread j */) {
  if (threadIdx.x < kG) {
    const int s = s0 + threadIdx.x;
    sm.x[threadIdx.x] = s < send ? p.x[static_cast<long long>(s) * p.L + u] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kH1 * kG; i += kCT) {
    const int k = i / kG, s = i - k * kG;
    const float z = fmaf(sm.w1[k], sm.x[s], sm.b1[k]);
    sm.z1t[k][s] = z;
    sm.h1t[k][s] = elu_f(z);
  }
  __syncthreads();
  const int j = threadIdx.x;
  float acc[kG];
#pragma unroll
  for (int s = 0; s < kG; ++s) acc[s] = sm.b2[j];
#pragma unroll 4
  for (int k = 0; k < kH1; ++k) {
    const float w = sm.w2t[k][j];
    const float4 a = *reinterpret_cast<const float4*>(&sm.h1t[k][0]);
    const float4 b = *reinterpret_cast<const float4*>(&sm.h1t[k][4]);
    acc[0] = fmaf(w, a.x, acc[0]);
    acc[1] = fmaf(w, a.y, acc[1]);
    acc[2] = fmaf(w, a.z, acc[2]);
    acc[3] = fmaf(w, a.w, acc[3]);
    acc[4] = fmaf(w, b.x, acc[4]);
    acc[5] = fmaf(w, b.y, acc[5]);
    acc[6] = fmaf(w, b.z, acc[6]);
    acc[7] = fmaf(w, b.w, acc[7]);
  }
#pragma unroll
  for (int s = 0; s < kG; ++s) {
    z2_out[s] = acc[s];
    sm.h2[s][j] = elu_f(acc[s]);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kCT) concept_mlp_fwd_kernel(const ConceptParams p) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  ConceptSmem& sm = *reinterpret_cast<ConceptSmem*>(smraw);
  const int u = blockIdx.x, Dc = p.Dc;
  const int per = (p.B + gridDim.y - 1) / gridDim.y;
  const int sbeg = blockIdx.y * per, send = min(p.B, sbeg + per);
  concept_stage_weights<false>(sm, p, u, Dc);
  __syncthreads();
  for (int s0 = sbeg; s0 < send; s0 += kG) {
    float z2[kG];
    concept_hidden(sm, p, u, s0, send, z2);
    for (int i = threadIdx.x; i < kG * Dc; i += kCT) {
      const int s = i / Dc, o = i - s * Dc;
      float a = sm.b3[o];
#pragma unroll 8
      for (int j = 0; j < kH2; ++j) a = fmaf(sm.w3t[j][o], sm.h2[s][j], a);
      if (s0 + s < send) p.y[(static_cast<long long>(s0 + s) * p.L + u) * Dc + o] = a;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kCT) concept_mlp_bwd_kernel(const ConceptParams p) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  ConceptSmem& sm = *reinterpret_cast<ConceptSmem*>(smraw);
  const int u = blockIdx.x, Dc = p.Dc;
  const int per = (p.B + gridDim.y - 1) / gridDim.y;
  const int sbeg = blockIdx.y * per, send = min(p.B, sbeg + per);
  concept_stage_weights<true>(sm, p, u, Dc);
  __syncthreads();
  const int j = threadIdx.x;           // layer-2 / layer-3 role: hidden unit j
  const int k1 = threadIdx.x & (kH1 - 1), half = threadIdx.x >> 6;   // layer-1 role: hidden unit k1, samples 4*half..+3
  float gw2[kH1];                      // d W2[j][:]
  float gw3[32];                       // d W3[:][j]
  float gb2 = 0.f, gw1 = 0.f, gb1 = 0.f, gb3 = 0.f;
#pragma unroll
  for (int k = 0; k < kH1; ++k) gw2[k] = 0.f;
#pragma unroll
  for (int o = 0; o < 32; ++o) gw3[o] = 0.f;
  for (int s0 = sbeg; s0 < send; s0 += kG) {
    float z2[kG];
    concept_hidden(sm, p, u, s0, send, z2);
    for (int i = threadIdx.x; i < kG * Dc; i += kCT) {
      const int s = i / Dc, o = i - s * Dc;
      sm.dy[s][o] = s0 + s < send ? p.dy[(static_cast<long long>(s0 + s) * p.L + u) * Dc + o] : 0.f;
    }
    __syncthreads();
    if (threadIdx.x < Dc) {
#pragma unroll
      for (int s = 0; s < kG; ++s) gb3 += sm.dy[s][threadIdx.x];
    }
    // layer 3 backward: dh2[s][j], dW3[o][j]; then through ELU: dz2 = dh2 * (z2 > 0 ? 1 : h2 + 1)
    float dz2[kG];
#pragma unroll
    for (int s = 0; s < kG; ++s) {
      const float h2 = sm.h2[s][j];
      float a = 0.f;
#pragma unroll
      for (int o = 0; o < 32; ++o) {       // fully unrolled so that gw3 stays in registers
        if (o < Dc) {
          const float d = sm.dy[s][o];
          a = fmaf(sm.w3t[j][o], d, a);
          gw3[o] = fmaf(d, h2, gw3[o]);
        }
      }
      dz2[s] = a * (z2[s] > 0.f ? 1.f : h2 + 1.f);
      gb2 += dz2[s];
      sm.dz2t[j][s] = dz2[s];
    }
    // dW2[j][k] += sum_s dz2[s][j] * h1[s][k]
#pragma unroll
    for (int k = 0; k < kH1; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&sm.h1t[k][0]);
      const float4 b = *reinterpret_cast<const float4*>(&sm.h1t[k][4]);
      float g = gw2[k];
      g = fmaf(dz2[0], a.x, g);
      g = fmaf(dz2[1], a.y, g);
      g = fmaf(dz2[2], a.z, g);
      g = fmaf(dz2[3], a.w, g);
      g = fmaf(dz2[4], b.x, g);
      g = fmaf(dz2[5], b.y, g);
      g = fmaf(dz2[6], b.z, g);
      g = fmaf(dz2[7], b.w, g);
      gw2[k] = g;
    }
    __syncthreads();
    // layer 2 backward: dh1[s][k1] = sum_j W2[j][k1] dz2[s][j] for this thread's 4 samples; then layer 1
    float dh1[4] = {0.f, 0.f, 0.f, 0.f};
    for (int jj = 0; jj < kH2; ++jj) {
      const float w = sm.w2[jj][k1];
      const float4 d = *reinterpret_cast<const float4*>(&sm.dz2t[jj][4 * half]);
      dh1[0] = fmaf(w, d.x, dh1[0]);
      dh1[1] = fmaf(w, d.y, dh1[1]);
      dh1[2] = fmaf(w, d.z, dh1[2]);
      dh1[3] = fmaf(w, d.w, dh1[3]);
    }
    float dxp[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int s = 4 * half + i;
      const float z1 = sm.z1t[k1][s];
      const float dz1 = dh1[i] * (z1 > 0.f ? 1.f : sm.h1t[k1][s] + 1.f);
      gw1 = fmaf(dz1, sm.x[s], gw1);
      gb1 += dz1;
      dxp[i] = sm.w1[k1] * dz1;
    }
    // dx[s] = sum_k w1[k] dz1[s][k]: the 64 threads of a half are two warps
#pragma unroll
    for (int i = 0; i < 4; ++i) dxp[i] = warp_sum(dxp[i]);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) sm.red[(threadIdx.x >> 5) * 4 + i] = dxp[i];
    }
    __syncthreads();
    if (threadIdx.x < kG && p.dx) {
      const int s = threadIdx.x, h = s >> 2, i = s & 3;
      if (s0 + s < send) p.dx[static_cast<long long>(s0 + s) * p.L + u] = sm.red[(2 * h) * 4 + i] + sm.red[(2 * h + 1) * 4 + i];
    }
    __syncthreads();
  }
  // parameter gradients of this chunk
  float* dW2 = p.dw2 + static_cast<long long>(u) * kH2 * kH1 + static_cast<long long>(j) * kH1;
#pragma unroll
  for (int k = 0; k < kH1; ++k) atomicAdd(dW2 + k, gw2[k]);
#pragma unroll
  for (int o = 0; o < 32; ++o)
    if (o < Dc) atomicAdd(p.dw3 + (static_cast<long long>(u) * Dc + o) * kH2 + j, gw3[o]);
  atomicAdd(p.db2 + u * kH2 + j, gb2);
  if (threadIdx.x < Dc) atomicAdd(p.db3 + u * Dc + threadIdx.x, gb3);
  // layer-1 gradients: the two halves hold partial sums of the same k1
  sm.red[threadIdx.x] = gw1;
  __syncthreads();
  if (threadIdx.x < kH1) atomicAdd(p.dw1 + u * kH1 + threadIdx.x, sm.red[threadIdx.x] + sm.red[threadIdx.x + kH1]);
  __syncthreads();
  sm.red[threadIdx.x] = gb1;
  __syncthreads();
  if (threadIdx.x < kH1) atomicAdd(p.db1 + u * kH1 + threadIdx.x, sm.red[threadIdx.x] + sm.red[threadIdx.x + kH1]);
}

static int concept_chunks(int B) {
  int c = (B + 63) / 64;             // about 64 samples (8 groups) per CTA
  return c < 1 ? 1 : (c > 16 ? 16 : c);
}

int launch_concept_mlp_fwd(const ConceptParams& p, cudaStream_t s) {
  if (p.Dc < 1 || p.Dc > 32 || p.L < 1 || p.B < 1) return static_cast<int>(cudaErrorInvalidValue);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(concept_mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(ConceptSmem)));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  launch_k(concept_mlp_fwd_kernel, dim3(p.L, concept_chunks(p.B)), kCT, sizeof(ConceptSmem), s, p);
  return static_cast<int>(cudaGetLastError());
}

int launch_concept_mlp_bwd(const ConceptParams& p, cudaStream_t s) {
  if (p.Dc < 1 || p.Dc > 32 || p.L < 1 || p.B < 1) return static_cast<int>(cudaErrorInvalidValue);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(concept_mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(ConceptSmem)));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  launch_k(concept_mlp_bwd_kernel, dim3(p.L, concept_chunks(p.B)), kCT, sizeof(ConceptSmem), s, p);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
