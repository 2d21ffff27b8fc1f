// Spatial self-attention on the warp-level tensor-core path (mma.sync m16n8k8 / m16n8k16, fp32 accumulate).
// Reference: CrossAttention.forward with context=None, ldm/modules/attention.py:170-193.
//
// Shapes here are tiny per (sample, head): N in {16, 64, 256} tokens, d_head in {8, 16, 32} — far below a tcgen05 tile
// (M = 128, K >= 16 would pad d_head = 8 to 16 and need N = 256 keys staged as descriptors), so the right tool is the
// register-fragment MMA: a warp owns 16 queries (forward / dQ) or 16 keys (dK, dV); K, V (and their transposes) of the
// (sample, head) live in shared memory; scores never leave registers.  The accumulator layout of two adjacent
// m16n8 score tiles is exactly the A-fragment layout of a k16 step, so P (and dS) feed the second MMA without any
// shuffle.  Softmax is two-pass (row max first, then exp / sum / PV): the QK^T MMAs are recomputed instead of
// rescaling O, they are ~free.
#include "common.cuh"
#include "kernels.cuh"

namespace edb {

static constexpr float kLog2eM = 1.4426950408889634f;

#ifdef EDB_ACT_BF16
#define EDB_MMA_T "bf16"
#else
#define EDB_MMA_T "f16"
#endif

__device__ __forceinline__ void mma_k8(float* c, uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32." EDB_MMA_T "." EDB_MMA_T ".f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void mma_k16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." EDB_MMA_T "." EDB_MMA_T ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t lds32(const h16* p) { return *reinterpret_cast<const uint32_t*>(p); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  const h162 v = f22h2(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// A-operand fragments of rows r0..r0+15 of a row-major [rows][D] matrix (k = the D axis):
//   D == 8: one k8 step (2 registers);  D >= 16: D/16 k16 steps (4 registers each)
template <int D>
struct AFrag {
  uint32_t r[D == 8 ? 2 : D / 4];
  __device__ __forceinline__ void load(const h16* X, int r0, int g, int t) {
    const h16* p0 = X + (r0 + g) * D + 2 * t;
    const h16* p1 = p0 + 8 * D;
    if (D == 8) {
      r[0] = lds32(p0);
      r[1] = lds32(p1);
    } else {
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
        r[4 * ks] = lds32(p0 + 16 * ks);
        r[4 * ks + 1] = lds32(p1 + 16 * ks);
        r[4 * ks + 2] = lds32(p0 + 16 * ks + 8);
        r[4 * ks + 3] = lds32(p1 + 16 * ks + 8);
      }
    }
  }
};
// c[16 x 8] += A(rows) * Y(rows n0..n0+7)^T, Y row-major [rows][D] (B[k = d][n = row])
template <int D>
__device__ __forceinline__ void mma_rowsT(float* c, const AFrag<D>& a, const h16* Y, int n0, int g, int t) {
  const h16* p = Y + (n0 + g) * D + 2 * t;
  if (D == 8) {
    mma_k8(c, a.r[0], a.r[1], lds32(p));
  } else {
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) mma_k16(c, a.r + 4 * ks, lds32(p + 16 * ks), lds32(p + 16 * ks + 8));
  }
}

// ---------------------------------------------------------------------------------------------------------------
// forward.  CTA = (sample, group of G heads); warp = (head slot, 16-query tile).
// qkv [B*N, 3C] (q | k | v, head h at columns h*D), out [B*N, C], lse [B*heads*N] (log2 domain).
// ---------------------------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(512) self_attn_fwd_mma_kernel(const h16* __restrict__ qkv, h16* __restrict__ out,
                                                                float* __restrict__ lse, int N, int heads, int G,
                                                                float scale_log2) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  const int NP = N + 8;
  h16* Ks = reinterpret_cast<h16*>(smraw);   // [G][N][D]
  h16* Vt = Ks + G * N * D;                  // [G][D][NP]
  h16* Qs = Vt + G * D * NP;                 // [G][N][D]
  const int C = heads * D;
  const int groups = heads / G;
  const int b = blockIdx.x / groups, hg = blockIdx.x - b * groups;
  constexpr int DC = D / 8;
  for (int idx = threadIdx.x; idx < G * N * DC; idx += blockDim.x) {
    const int slot = idx / (N * DC);
    const int rem = idx - slot * N * DC;
    const int j = rem / DC, dc = rem - j * DC;
    const int h = hg * G + slot;
    const h16* base = qkv + (static_cast<long long>(b) * N + j) * 3 * C + h * D + dc * 8;
    *reinterpret_cast<uint4*>(Qs + (slot * N + j) * D + dc * 8) = *reinterpret_cast<const uint4*>(base);
    *reinterpret_cast<uint4*>(Ks + (slot * N + j) * D + dc * 8) = *reinterpret_cast<const uint4*>(base + C);
    const uint4 v = *reinterpret_cast<const uint4*>(base + 2 * C);
    const h16* ve = reinterpret_cast<const h16*>(&v);
#pragma unroll
    for (int k = 0; k < 8; ++k) Vt[(slot * D + dc * 8 + k) * NP + j] = ve[k];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int qtiles = N >> 4;
  const int slot = warp / qtiles, q0 = (warp - slot * qtiles) << 4;
  if (slot >= G) return;
  const int h = hg * G + slot;
  const h16* Kh = Ks + slot * N * D;
  const h16* Vh = Vt + slot * D * NP;
  AFrag<D> qa;
  qa.load(Qs + slot * N * D, q0, g, t);
  // pass 1: row maxima of the raw scores (rows g and g+8 of this tile)
  float m0 = -INFINITY, m1 = -INFINITY;
  for (int kt = 0; kt < N; kt += 8) {
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    mma_rowsT<D>(c, qa, Kh, kt, g, t);
    m0 = fmaxf(m0, fmaxf(c[0], c[1]));
    m1 = fmaxf(m1, fmaxf(c[2], c[3]));
  }
  m0 = quad_max(m0) * scale_log2;
  m1 = quad_max(m1) * scale_log2;
  // pass 2: p = 2^(s*scale_log2 - m), l = sum p, O = P V
  float l0 = 0.f, l1 = 0.f;
  float o[DC][4];
#pragma unroll
  for (int nt = 0; nt < DC; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
  for (int kc = 0; kc < N; kc += 16) {
    float c[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    mma_rowsT<D>(c[0], qa, Kh, kc, g, t);
    mma_rowsT<D>(c[1], qa, Kh, kc + 8, g, t);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      c[i][0] = exp2f(c[i][0] * scale_log2 - m0);
      c[i][1] = exp2f(c[i][1] * scale_log2 - m0);
      c[i][2] = exp2f(c[i][2] * scale_log2 - m1);
      c[i][3] = exp2f(c[i][3] * scale_log2 - m1);
      l0 += c[i][0] + c[i][1];
      l1 += c[i][2] + c[i][3];
    }
    uint32_t pa[4] = {pack2(c[0][0], c[0][1]), pack2(c[0][2], c[0][3]), pack2(c[1][0], c[1][1]), pack2(c[1][2], c[1][3])};
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) {
      const h16* vp = Vh + (nt * 8 + g) * NP + kc + 2 * t;
      mma_k16(o[nt], pa, lds32(vp), lds32(vp + 8));
    }
  }
  l0 = quad_sum(l0);
  l1 = quad_sum(l1);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  const long long row0 = static_cast<long long>(b) * N + q0 + g;
#pragma unroll
  for (int nt = 0; nt < DC; ++nt) {
    *reinterpret_cast<uint32_t*>(out + row0 * C + h * D + nt * 8 + 2 * t) = pack2(o[nt][0] * i0, o[nt][1] * i0);
    *reinterpret_cast<uint32_t*>(out + (row0 + 8) * C + h * D + nt * 8 + 2 * t) = pack2(o[nt][2] * i1, o[nt][3] * i1);
  }
  if (t == 0) {
    float* lp = lse + (static_cast<long long>(b) * heads + h) * N + q0 + g;
    lp[0] = m0 + log2f(l0);
    lp[8] = m1 + log2f(l1);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// backward.  Same CTA / warp mapping; every warp first produces dQ for its 16 queries, then dK and dV for its 16 keys
// (transposed-role MMAs: S^T = K Q^T, so P^T and dS^T come out in A-fragment layout for dV = P^T dO, dK = dS^T Q).
// dO is normalised by a per-CTA power of two f (max|f dO| in [0.5, 1)) before it becomes an MMA operand, so that the
// fp16 dS fragments stay in the normal range whatever the loss scale; results are multiplied back by 1/f.
// dqkv [B*N, 3C] = dq | dk | dv.
// ---------------------------------------------------------------------------------------------------------------
// B fragment of a transposed [D][NP] matrix (B[k = row index][n = d]): rows k0+2t.., column d0+g
__device__ __forceinline__ void mma_T(float* c, const uint32_t* a, const h16* Yt, int NP, int d0, int k0, int g, int t) {
  const h16* p = Yt + (d0 + g) * NP + k0 + 2 * t;
  mma_k16(c, a, lds32(p), lds32(p + 8));
}

template <int D>
__global__ void __launch_bounds__(512) self_attn_bwd_mma_kernel(const h16* __restrict__ qkv, const h16* __restrict__ out,
                                                                const h16* __restrict__ dout,
                                                                const float* __restrict__ lse, h16* __restrict__ dqkv,
                                                                int N, int heads, int G, float scale) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  const int NP = N + 8;
  h16* Qs = reinterpret_cast<h16*>(smraw);   // [G][N][D]
  h16* Ks = Qs + G * N * D;
  h16* Vs = Ks + G * N * D;
  h16* Os = Vs + G * N * D;                  // f * dO
  h16* Qt = Os + G * N * D;                  // [G][D][NP]
  h16* Kt = Qt + G * D * NP;
  h16* Ot = Kt + G * D * NP;                 // (f * dO)^T
  float* Ls = reinterpret_cast<float*>(Ot + G * D * NP);   // [G][N] lse (log2 domain)
  float* Ds = Ls + G * N;                                  // [G][N] delta = rowsum(f dO * O)
  __shared__ float s_red[16];
  __shared__ float s_f;
  const int C = heads * D;
  const int groups = heads / G;
  const int b = blockIdx.x / groups, hg = blockIdx.x - b * groups;
  constexpr int DC = D / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // ---- CTA-wide max |dO| -> power-of-two normaliser ------------------------------------------------------------
  float amax = 0.f;
  for (int idx = threadIdx.x; idx < G * N * DC; idx += blockDim.x) {
    const int slot = idx / (N * DC);
    const int rem = idx - slot * N * DC;
    const int j = rem / DC, dc = rem - j * DC;
    float f[8];
    unpack8(*reinterpret_cast<const uint4*>(dout + (static_cast<long long>(b) * N + j) * C + (hg * G + slot) * D + dc * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(f[k]));
  }
  amax = warp_max(amax);
  if (lane == 0) s_red[warp] = amax;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) m = fmaxf(m, s_red[w]);
    float fsc = 1.f;
    if (m > 0.f && isfinite(m)) {
      int e;
      frexpf(m, &e);
      fsc = ldexpf(1.f, -e);
    }
    s_f = fsc;
  }
  __syncthreads();
  const float fsc = s_f, finv = 1.f / s_f;
  // ---- stage Q, K, V, f*dO (row-major and the transposes the second MMAs need), lse, delta ---------------------
  for (int idx = threadIdx.x; idx < G * N * DC; idx += blockDim.x) {
    const int slot = idx / (N * DC);
    const int rem = idx - slot * N * DC;
    const int j = rem / DC, dc = rem - j * DC;
    const int h = hg * G + slot;
    const long long row = static_cast<long long>(b) * N + j;
    const h16* base = qkv + row * 3 * C + h * D + dc * 8;
    const uint4 q4 = *reinterpret_cast<const uint4*>(base);
    const uint4 k4 = *reinterpret_cast<const uint4*>(base + C);
    const int o = (slot * N + j) * D + dc * 8;
    *reinterpret_cast<uint4*>(Qs + o) = q4;
    *reinterpret_cast<uint4*>(Ks + o) = k4;
    *reinterpret_cast<uint4*>(Vs + o) = *reinterpret_cast<const uint4*>(base + 2 * C);
    float d8[8];
    unpack8(*reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8), d8);
#pragma unroll
    for (int k = 0; k < 8; ++k) d8[k] *= fsc;
    const uint4 o4 = pack8(d8);
    *reinterpret_cast<uint4*>(Os + o) = o4;
    const h16* qe = reinterpret_cast<const h16*>(&q4);
    const h16* ke = reinterpret_cast<const h16*>(&k4);
    const h16* oe = reinterpret_cast<const h16*>(&o4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int tt = (slot * D + dc * 8 + k) * NP + j;
      Qt[tt] = qe[k];
      Kt[tt] = ke[k];
      Ot[tt] = oe[k];
    }
  }
  for (int idx = threadIdx.x; idx < G * N; idx += blockDim.x) {
    const int slot = idx / N, j = idx - slot * N;
    const int h = hg * G + slot;
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
    Ds[idx] = dl * fsc;
    Ls[idx] = lse[(static_cast<long long>(b) * heads + h) * N + j];
  }
  __syncthreads();
  const int g = lane >> 2, t = lane & 3;
  const int qtiles = N >> 4;
  const int slot = warp / qtiles, r0 = (warp - slot * qtiles) << 4;
  if (slot >= G) return;
  const int h = hg * G + slot;
  const h16* Qh = Qs + slot * N * D;
  const h16* Kh = Ks + slot * N * D;
  const h16* Vh = Vs + slot * N * D;
  const h16* Oh = Os + slot * N * D;
  const h16* Qth = Qt + slot * D * NP;
  const h16* Kth = Kt + slot * D * NP;
  const h16* Oth = Ot + slot * D * NP;
  const float* Lh = Ls + slot * N;
  const float* Dh = Ds + slot * N;
  const float scale_log2 = scale * kLog2eM;
  const long long grow = static_cast<long long>(b) * N + r0 + g;
  {  // ---- dQ for queries r0 .. r0+15 ----
    AFrag<D> qa, oa;
    qa.load(Qh, r0, g, t);
    oa.load(Oh, r0, g, t);
    const float ls0 = Lh[r0 + g], ls1 = Lh[r0 + g + 8], de0 = Dh[r0 + g], de1 = Dh[r0 + g + 8];
    float dq[DC][4];
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) dq[nt][0] = dq[nt][1] = dq[nt][2] = dq[nt][3] = 0.f;
    for (int kc = 0; kc < N; kc += 16) {
      float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      float dp[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        mma_rowsT<D>(s[i], qa, Kh, kc + 8 * i, g, t);
        mma_rowsT<D>(dp[i], oa, Vh, kc + 8 * i, g, t);
        s[i][0] = exp2f(s[i][0] * scale_log2 - ls0) * (dp[i][0] - de0) * scale;
        s[i][1] = exp2f(s[i][1] * scale_log2 - ls0) * (dp[i][1] - de0) * scale;
        s[i][2] = exp2f(s[i][2] * scale_log2 - ls1) * (dp[i][2] - de1) * scale;
        s[i][3] = exp2f(s[i][3] * scale_log2 - ls1) * (dp[i][3] - de1) * scale;
      }
      uint32_t da[4] = {pack2(s[0][0], s[0][1]), pack2(s[0][2], s[0][3]), pack2(s[1][0], s[1][1]), pack2(s[1][2], s[1][3])};
#pragma unroll
      for (int nt = 0; nt < DC; ++nt) mma_T(dq[nt], da, Kth, NP, nt * 8, kc, g, t);
    }
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) {
      *reinterpret_cast<uint32_t*>(dqkv + grow * 3 * C + h * D + nt * 8 + 2 * t) = pack2(dq[nt][0] * finv, dq[nt][1] * finv);
      *reinterpret_cast<uint32_t*>(dqkv + (grow + 8) * 3 * C + h * D + nt * 8 + 2 * t) = pack2(dq[nt][2] * finv, dq[nt][3] * finv);
    }
  }
  {  // ---- dK, dV for keys r0 .. r0+15 ----
    AFrag<D> ka, va;
    ka.load(Kh, r0, g, t);
    va.load(Vh, r0, g, t);
    float dk[DC][4], dv[DC][4];
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) {
      dk[nt][0] = dk[nt][1] = dk[nt][2] = dk[nt][3] = 0.f;
      dv[nt][0] = dv[nt][1] = dv[nt][2] = dv[nt][3] = 0.f;
    }
    for (int qc = 0; qc < N; qc += 16) {
      float st[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      float dpt[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      float ds[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        mma_rowsT<D>(st[i], ka, Qh, qc + 8 * i, g, t);     // S^T[key][query]
        mma_rowsT<D>(dpt[i], va, Oh, qc + 8 * i, g, t);    // dP^T[key][query]
        const float2 lq = *reinterpret_cast<const float2*>(Lh + qc + 8 * i + 2 * t);   // columns 2t, 2t+1 = queries
        const float2 dq2 = *reinterpret_cast<const float2*>(Dh + qc + 8 * i + 2 * t);
        st[i][0] = exp2f(st[i][0] * scale_log2 - lq.x);
        st[i][1] = exp2f(st[i][1] * scale_log2 - lq.y);
        st[i][2] = exp2f(st[i][2] * scale_log2 - lq.x);
        st[i][3] = exp2f(st[i][3] * scale_log2 - lq.y);
        ds[i][0] = st[i][0] * (dpt[i][0] - dq2.x) * scale;
        ds[i][1] = st[i][1] * (dpt[i][1] - dq2.y) * scale;
        ds[i][2] = st[i][2] * (dpt[i][2] - dq2.x) * scale;
        ds[i][3] = st[i][3] * (dpt[i][3] - dq2.y) * scale;
      }
      uint32_t pa[4] = {pack2(st[0][0], st[0][1]), pack2(st[0][2], st[0][3]), pack2(st[1][0], st[1][1]), pack2(st[1][2], st[1][3])};
      uint32_t da[4] = {pack2(ds[0][0], ds[0][1]), pack2(ds[0][2], ds[0][3]), pack2(ds[1][0], ds[1][1]), pack2(ds[1][2], ds[1][3])};
#pragma unroll
      for (int nt = 0; nt < DC; ++nt) {
        mma_T(dv[nt], pa, Oth, NP, nt * 8, qc, g, t);
        mma_T(dk[nt], da, Qth, NP, nt * 8, qc, g, t);
      }
    }
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) {
      h16* p0 = dqkv + grow * 3 * C + C + h * D + nt * 8 + 2 * t;
      h16* p1 = dqkv + (grow + 8) * 3 * C + C + h * D + nt * 8 + 2 * t;
      *reinterpret_cast<uint32_t*>(p0) = pack2(dk[nt][0] * finv, dk[nt][1] * finv);
      *reinterpret_cast<uint32_t*>(p1) = pack2(dk[nt][2] * finv, dk[nt][3] * finv);
      *reinterpret_cast<uint32_t*>(p0 + C) = pack2(dv[nt][0] * finv, dv[nt][1] * finv);
      *reinterpret_cast<uint32_t*>(p1 + C) = pack2(dv[nt][2] * finv, dv[nt][3] * finv);
    }
  }
}

static int mma_group(int N, int heads) {
  const int qtiles = N / 16;
  int G = 16 / qtiles;
  if (G < 1) G = 1;
  if (G > heads) G = heads;
  while (heads % G) --G;
  return G;
}

template <int D>
static int fwd_mma_t(const h16* qkv, h16* out, float* lse, int B, int N, int heads, cudaStream_t st) {
  const int G = mma_group(N, heads);
  const size_t smem = static_cast<size_t>(G) * (2 * N * D + D * (N + 8)) * sizeof(h16);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(self_attn_fwd_mma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int threads = G * (N / 16) * 32;
  launch_k(self_attn_fwd_mma_kernel<D>, B * (heads / G), threads, smem, st, qkv, out, lse, N, heads, G,
                                                                      rsqrtf(static_cast<float>(D)) * kLog2eM);
  return static_cast<int>(cudaGetLastError());
}

// returns -1 when the shape is not handled here (caller falls back to the CUDA-core kernel)
int launch_self_attn_fwd_mma(const h16* qkv, h16* out, float* lse, int B, int N, int heads, int D, cudaStream_t s) {
  if (N < 16 || (N & 15) || N > 256) return -1;
  switch (D) {
    case 8: return fwd_mma_t<8>(qkv, out, lse, B, N, heads, s);
    case 16: return fwd_mma_t<16>(qkv, out, lse, B, N, heads, s);
    case 32: return fwd_mma_t<32>(qkv, out, lse, B, N, heads, s);
    default: return -1;
  }
}

template <int D>
static int bwd_mma_t(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                     int heads, cudaStream_t st) {
  const int G = mma_group(N, heads);
  const size_t smem = static_cast<size_t>(G) * ((4 * N * D + 3 * D * (N + 8)) * sizeof(h16) + 2 * N * sizeof(float));
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(self_attn_bwd_mma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int threads = G * (N / 16) * 32;
  launch_k(self_attn_bwd_mma_kernel<D>, B * (heads / G), threads, smem, st, qkv, out, dout, lse, dqkv, N, heads, G,
                                                                      rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

int launch_self_attn_bwd_mma(const h16* qkv, const h16* out, const h16* dout, const float* lse, h16* dqkv, int B, int N,
                             int heads, int D, cudaStream_t s) {
  if (N < 16 || (N & 15) || N > 256) return -1;
  switch (D) {
    case 8: return bwd_mma_t<8>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 16: return bwd_mma_t<16>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    case 32: return bwd_mma_t<32>(qkv, out, dout, lse, dqkv, B, N, heads, s);
    default: return -1;
  }
}

}  // namespace edb
