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
  // The row sums l ride on the tensor pipe: one extra MMA per 16-key block against an all-ones B fragment (every
  // column of `ls` is the row sum of the SAME 16-bit P that feeds P V, so normalisation and numerator agree) instead of
  // 8 FADDs + a quad reduction -- the loop is instruction-issue bound (ncu: 60 % issue-active, 37 instructions per block).
  float ls[4] = {0.f, 0.f, 0.f, 0.f};
  const uint32_t ones = pack2(1.f, 1.f);
  float o[DC][4];
#pragma unroll
  for (int nt = 0; nt < DC; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
  for (int kc = 0; kc < N; kc += 16) {
    float c[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    mma_rowsT<D>(c[0], qa, Kh, kc, g, t);
    mma_rowsT<D>(c[1], qa, Kh, kc + 8, g, t);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      c[i][0] = ex2_approx(c[i][0] * scale_log2 - m0);
      c[i][1] = ex2_approx(c[i][1] * scale_log2 - m0);
      c[i][2] = ex2_approx(c[i][2] * scale_log2 - m1);
      c[i][3] = ex2_approx(c[i][3] * scale_log2 - m1);
    }
    uint32_t pa[4] = {pack2(c[0][0], c[0][1]), pack2(c[0][2], c[0][3]), pack2(c[1][0], c[1][1]), pack2(c[1][2], c[1][3])};
    mma_k16(ls, pa, ones, ones);
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) {
      const h16* vp = Vh + (nt * 8 + g) * NP + kc + 2 * t;
      mma_k16(o[nt], pa, lds32(vp), lds32(vp + 8));
    }
  }
  const float l0 = ls[0], l1 = ls[2];
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
  pdl_launch_dependents();
  // q | k | v, O and the log-sum-exp of this sample were saved by the forward; only dO comes from the previous kernel:
  // the sample's first head group pulls them into L2 before the wait
  if (threadIdx.x == 0 && blockIdx.x % (heads / G) == 0) {
    const long long b0 = blockIdx.x / (heads / G);
    const int Cc = heads * D;
    l2_prefetch_bulk(qkv + b0 * N * 3 * Cc, static_cast<uint32_t>(N * 3 * Cc * 2));
    l2_prefetch_bulk(out + b0 * N * Cc, static_cast<uint32_t>(N * Cc * 2));
  }
  pdl_wait();
  extern __shared__ __align__(16) uint8_t smraw[];
  const int NP = N + 8;
  h16* Qs = reinterpret_cast<h16*>(smraw);   // [G][N][D]
  h16* Ks = Qs + G * N * D;
  h16* Vs = Ks + G * N * D;
  h16* Os = Vs + G * N * D;                  // dO
  h16* Qt = Os + G * N * D;                  // [G][D][NP]
  h16* Kt = Qt + G * D * NP;
  h16* Ot = Kt + G * D * NP;                 // dO^T
  float* Ls = reinterpret_cast<float*>(Ot + G * D * NP);   // [G][N] lse (log2 domain)
  float* Ds = Ls + G * N;                                  // [G][N] delta = rowsum(dO * O)
  __shared__ float s_red[16];
  const int C = heads * D;
  const int groups = heads / G;
  const int b = blockIdx.x / groups, hg = blockIdx.x - b * groups;
  constexpr int DC = D / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float amax = 0.f;      // CTA-wide max |dO| (collected while staging) -> power-of-two scale of the fp16 dS fragments
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
    const uint4 o4 = *reinterpret_cast<const uint4*>(dout + row * C + h * D + dc * 8);
    {
      float d8[8];
      unpack8(o4, d8);
#pragma unroll
      for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(d8[k]));
    }
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
    Ds[idx] = dl;
    Ls[idx] = lse[(static_cast<long long>(b) * heads + h) * N + j];
  }
  amax = warp_max(amax);
  if (lane == 0) s_red[warp] = amax;
  __syncthreads();
  float fsc = 1.f;
  {
    float m = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) m = fmaxf(m, s_red[w]);
    if (m > 0.f && isfinite(m)) {
      int e;
      frexpf(m, &e);
      fsc = ldexpf(1.f, -e);
    }
  }
  const float finv = 1.f / fsc, sfs = scale * fsc;   // dS is carried multiplied by fsc, dq / dk are multiplied back
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
    // sfs = scale * fsc is folded into the exponent (P * sfs = 2^(x - lse + log2 sfs)): one multiply less per score
    const float lsf = log2f(sfs);
    const float ls0 = Lh[r0 + g] - lsf, ls1 = Lh[r0 + g + 8] - lsf, de0 = Dh[r0 + g], de1 = Dh[r0 + g + 8];
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
        s[i][0] = ex2_approx(s[i][0] * scale_log2 - ls0) * (dp[i][0] - de0);
        s[i][1] = ex2_approx(s[i][1] * scale_log2 - ls0) * (dp[i][1] - de0);
        s[i][2] = ex2_approx(s[i][2] * scale_log2 - ls1) * (dp[i][2] - de1);
        s[i][3] = ex2_approx(s[i][3] * scale_log2 - ls1) * (dp[i][3] - de1);
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
        float2 dq2 = *reinterpret_cast<const float2*>(Dh + qc + 8 * i + 2 * t);
        dq2.x *= -sfs;       // ds = P * fma(dP, sfs, -delta * sfs)
        dq2.y *= -sfs;
        st[i][0] = ex2_approx(st[i][0] * scale_log2 - lq.x);
        st[i][1] = ex2_approx(st[i][1] * scale_log2 - lq.y);
        st[i][2] = ex2_approx(st[i][2] * scale_log2 - lq.x);
        st[i][3] = ex2_approx(st[i][3] * scale_log2 - lq.y);
        ds[i][0] = st[i][0] * fmaf(dpt[i][0], sfs, dq2.x);
        ds[i][1] = st[i][1] * fmaf(dpt[i][1], sfs, dq2.y);
        ds[i][2] = st[i][2] * fmaf(dpt[i][2], sfs, dq2.x);
        ds[i][3] = st[i][3] * fmaf(dpt[i][3], sfs, dq2.y);
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
      *reinterpret_cast<uint32_t*>(p0 + C) = pack2(dv[nt][0], dv[nt][1]);
      *reinterpret_cast<uint32_t*>(p1 + C) = pack2(dv[nt][2], dv[nt][3]);
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


// ===============================================================================================================
// Cross-attention against T <= 32 concept tokens on the same fragment path (attention.py:170-193 with a context).
// K | V come already projected (fp32 [B, T, 2C], launch_cross_kv_project); they are converted to h16 and padded to
// NK = 32 keys in shared memory, padded keys are masked out of the softmax.  A warp owns 16 queries and sees the whole
// score row (32 keys = four m16n8 tiles) in registers, so softmax is single pass and needs no log-sum-exp tensor.
// CTA = (sample, G heads), warp = (head slot, 16-query tile), as in the self-attention kernels.
// ===============================================================================================================
static constexpr int kNK = 32;          // padded key count
static constexpr int kNKP = kNK + 8;    // pitch of the transposed [D][keys] arrays

template <int D>
__device__ __forceinline__ void stage_kv(const float* __restrict__ kvb, int T, int C, int hg, int G, h16* Ks, h16* Vs,
                                         h16* Kt, h16* Vt) {
  // kvb: this sample's [T][2C];  Ks/Vs [G][kNK][D] row-major, Kt/Vt [G][D][kNKP] (any of them may be null).
  // The G heads of the CTA are G*D contiguous floats per key: float4 loads, 4 in flight per thread, then scatter.
  const int w4 = G * D / 4;                    // float4 per key and matrix
  const int total = kNK * w4;
  for (int i0 = threadIdx.x; i0 < total; i0 += 4 * blockDim.x) {
    float4 kq[4], vq[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      kq[u] = vq[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < total) {
        const int j = i / w4, c4 = i - j * w4;
        if (j < T) {
          const float* p = kvb + static_cast<long long>(j) * 2 * C + hg * G * D + c4 * 4;
          kq[u] = *reinterpret_cast<const float4*>(p);
          vq[u] = *reinterpret_cast<const float4*>(p + C);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      if (i < total) {
        const int j = i / w4, c = (i - j * w4) * 4;
        const int slot = c / D, d = c - slot * D;
        const float kf[4] = {kq[u].x, kq[u].y, kq[u].z, kq[u].w};
        const float vf[4] = {vq[u].x, vq[u].y, vq[u].z, vq[u].w};
        if (Ks) {
          *reinterpret_cast<h162*>(Ks + (slot * kNK + j) * D + d) = f22h2(kf[0], kf[1]);
          *reinterpret_cast<h162*>(Ks + (slot * kNK + j) * D + d + 2) = f22h2(kf[2], kf[3]);
        }
        if (Vs) {
          *reinterpret_cast<h162*>(Vs + (slot * kNK + j) * D + d) = f22h2(vf[0], vf[1]);
          *reinterpret_cast<h162*>(Vs + (slot * kNK + j) * D + d + 2) = f22h2(vf[2], vf[3]);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (Kt) Kt[(slot * D + d + k) * kNKP + j] = f2h(kf[k]);
          if (Vt) Vt[(slot * D + d + k) * kNKP + j] = f2h(vf[k]);
        }
      }
    }
  }
}

template <int D>
__global__ void __launch_bounds__(512) cross_attn_fwd_mma_kernel(const h16* __restrict__ q, const float* __restrict__ kv,
                                                                 h16* __restrict__ out, int N, int heads, int G, int T,
                                                                 float scale_log2) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  const int NT = (N + 15) & ~15;             // query rows padded to whole 16-row tiles (zero rows, never stored)
  h16* Qs = reinterpret_cast<h16*>(smraw);   // [G][NT][D]
  h16* Ks = Qs + G * NT * D;                 // [G][kNK][D]
  h16* Vt = Ks + G * kNK * D;                // [G][D][kNKP]
  const int C = heads * D;
  const int groups = heads / G;
  const int b = blockIdx.x / groups, hg = blockIdx.x - b * groups;
  constexpr int DC = D / 8;
  for (int idx = threadIdx.x; idx < G * NT * DC; idx += blockDim.x) {
    const int slot = idx / (NT * DC);
    const int rem = idx - slot * NT * DC;
    const int j = rem / DC, dc = rem - j * DC;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (j < N) v = *reinterpret_cast<const uint4*>(q + (static_cast<long long>(b) * N + j) * C + (hg * G + slot) * D + dc * 8);
    *reinterpret_cast<uint4*>(Qs + (slot * NT + j) * D + dc * 8) = v;
  }
  stage_kv<D>(kv + static_cast<long long>(b) * T * 2 * C, T, C, hg, G, Ks, nullptr, nullptr, Vt);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int qtiles = NT >> 4;
  const int slot = warp / qtiles, q0 = (warp - slot * qtiles) << 4;
  if (slot >= G) return;
  const int h = hg * G + slot;
  const h16* Kh = Ks + slot * kNK * D;
  const h16* Vh = Vt + slot * D * kNKP;
  AFrag<D> qa;
  qa.load(Qs + slot * NT * D, q0, g, t);
  float s[4][4];
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
    mma_rowsT<D>(s[nt], qa, Kh, nt * 8, g, t);
    const int key = nt * 8 + 2 * t;
    if (key >= T) s[nt][0] = s[nt][2] = -INFINITY;
    if (key + 1 >= T) s[nt][1] = s[nt][3] = -INFINITY;
    m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
    m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
  }
  m0 = quad_max(m0) * scale_log2;
  m1 = quad_max(m1) * scale_log2;
  float l0 = 0.f, l1 = 0.f;
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    s[nt][0] = ex2_approx(s[nt][0] * scale_log2 - m0);
    s[nt][1] = ex2_approx(s[nt][1] * scale_log2 - m0);
    s[nt][2] = ex2_approx(s[nt][2] * scale_log2 - m1);
    s[nt][3] = ex2_approx(s[nt][3] * scale_log2 - m1);
    l0 += s[nt][0] + s[nt][1];
    l1 += s[nt][2] + s[nt][3];
  }
  const float i0 = 1.f / quad_sum(l0), i1 = 1.f / quad_sum(l1);
  float o[DC][4];
#pragma unroll
  for (int nt = 0; nt < DC; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
  for (int kc = 0; kc < 2; ++kc) {
    uint32_t pa[4] = {pack2(s[2 * kc][0], s[2 * kc][1]), pack2(s[2 * kc][2], s[2 * kc][3]),
                      pack2(s[2 * kc + 1][0], s[2 * kc + 1][1]), pack2(s[2 * kc + 1][2], s[2 * kc + 1][3])};
#pragma unroll
    for (int nt = 0; nt < DC; ++nt) mma_T(o[nt], pa, Vh, kNKP, nt * 8, kc * 16, g, t);
  }
  const long long row0 = static_cast<long long>(b) * N + q0 + g;
#pragma unroll
  for (int nt = 0; nt < DC; ++nt) {
    if (q0 + g < N)
      *reinterpret_cast<uint32_t*>(out + row0 * C + h * D + nt * 8 + 2 * t) = pack2(o[nt][0] * i0, o[nt][1] * i0);
    if (q0 + g + 8 < N)
      *reinterpret_cast<uint32_t*>(out + (row0 + 8) * C + h * D + nt * 8 + 2 * t) = pack2(o[nt][2] * i1, o[nt][3] * i1);
  }
}

// backward: dq [B*N, C] h16 and dkv [B, T, 2C] fp32 (dK | dV, complete sums: the CTA sees every query of its heads, so
// no atomics in global memory and no zero-fill).  dO is normalised by a per-CTA power of two as in the self-attention
// backward; dq carries the caller's gradient scale, dkv likewise (the projection kernel applies 1/gscale).
template <int D>
__global__ void __launch_bounds__(512) cross_attn_bwd_mma_kernel(const h16* __restrict__ q, const float* __restrict__ kv,
                                                                 const h16* __restrict__ dout, h16* __restrict__ dq,
                                                                 float* __restrict__ dkv, int N, int heads, int G, int T,
                                                                 float scale) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smraw[];
  const int NT = (N + 15) & ~15;             // query rows padded to whole tiles: zero rows contribute nothing
  const int NP = NT + 8;
  h16* Qs = reinterpret_cast<h16*>(smraw);   // [G][NT][D]
  h16* Os = Qs + G * NT * D;                 // dO
  h16* Qt = Os + G * NT * D;                 // [G][D][NP]
  h16* Ot = Qt + G * D * NP;
  h16* Ks = Ot + G * D * NP;                 // [G][kNK][D]
  h16* Vs = Ks + G * kNK * D;
  h16* Kt = Vs + G * kNK * D;                // [G][D][kNKP]
  float* acc = reinterpret_cast<float*>(Kt + G * D * kNKP);   // [warps][kNK][2D] per-warp dK | dV partial sums
  float* qst = acc + (blockDim.x >> 5) * kNK * 2 * D;         // [warps][16][3] row max, 1/l, delta
  __shared__ float s_red[16];
  const int C = heads * D;
  const int groups = heads / G;
  const int b = blockIdx.x / groups, hg = blockIdx.x - b * groups;
  constexpr int DC = D / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float amax = 0.f;      // CTA-wide max |dO|, collected while staging
  for (int idx = threadIdx.x; idx < G * NT * DC; idx += blockDim.x) {
    const int slot = idx / (NT * DC);
    const int rem = idx - slot * NT * DC;
    const int j = rem / DC, dc = rem - j * DC;
    const long long goff = (static_cast<long long>(b) * N + j) * C + (hg * G + slot) * D + dc * 8;
    uint4 q4 = make_uint4(0u, 0u, 0u, 0u), o4 = q4;
    if (j < N) {
      q4 = *reinterpret_cast<const uint4*>(q + goff);
      o4 = *reinterpret_cast<const uint4*>(dout + goff);
      float d8[8];
      unpack8(o4, d8);
#pragma unroll
      for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(d8[k]));
    }
    const int o = (slot * NT + j) * D + dc * 8;
    *reinterpret_cast<uint4*>(Qs + o) = q4;
    *reinterpret_cast<uint4*>(Os + o) = o4;
    const h16* qe = reinterpret_cast<const h16*>(&q4);
    const h16* oe = reinterpret_cast<const h16*>(&o4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int tt = (slot * D + dc * 8 + k) * NP + j;
      Qt[tt] = qe[k];
      Ot[tt] = oe[k];
    }
  }
  stage_kv<D>(kv + static_cast<long long>(b) * T * 2 * C, T, C, hg, G, Ks, Vs, Kt, nullptr);
  amax = warp_max(amax);
  if (lane == 0) s_red[warp] = amax;
  __syncthreads();
  float fsc = 1.f;
  {
    float m = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) m = fmaxf(m, s_red[w]);
    if (m > 0.f && isfinite(m)) {
      int e;
      frexpf(m, &e);
      fsc = ldexpf(1.f, -e);
    }
  }
  const float finv = 1.f / fsc, sfs = scale * fsc;   // dS carried multiplied by fsc; dq / dK multiplied back
  const int g = lane >> 2, t = lane & 3;
  const int qtiles = NT >> 4;
  const int slot = warp / qtiles, q0 = (warp - slot * qtiles) << 4;
  if (slot < G) {
    const int h = hg * G + slot;
    const h16* Qh = Qs + slot * NT * D;
    const h16* Oh = Os + slot * NT * D;
    const h16* Qth = Qt + slot * D * NP;
    const h16* Oth = Ot + slot * D * NP;
    const h16* Kh = Ks + slot * kNK * D;
    const h16* Vh = Vs + slot * kNK * D;
    const h16* Kth = Kt + slot * D * kNKP;
    float* st = qst + warp * 48;
    const float scale_log2 = scale * kLog2eM;
    {  // ---- dQ for queries q0 .. q0+15 (rows g and g+8 of the tile) ----
      AFrag<D> qa, oa;
      qa.load(Qh, q0, g, t);
      oa.load(Oh, q0, g, t);
      float s[4][4], dp[4][4];
      float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
        mma_rowsT<D>(s[nt], qa, Kh, nt * 8, g, t);
        mma_rowsT<D>(dp[nt], oa, Vh, nt * 8, g, t);
        const int key = nt * 8 + 2 * t;
        if (key >= T) s[nt][0] = s[nt][2] = -INFINITY;
        if (key + 1 >= T) s[nt][1] = s[nt][3] = -INFINITY;
        m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
        m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
      }
      m0 = quad_max(m0) * scale_log2;
      m1 = quad_max(m1) * scale_log2;
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        s[nt][0] = ex2_approx(s[nt][0] * scale_log2 - m0);
        s[nt][1] = ex2_approx(s[nt][1] * scale_log2 - m0);
        s[nt][2] = ex2_approx(s[nt][2] * scale_log2 - m1);
        s[nt][3] = ex2_approx(s[nt][3] * scale_log2 - m1);
        l0 += s[nt][0] + s[nt][1];
        l1 += s[nt][2] + s[nt][3];
      }
      const float i0 = 1.f / quad_sum(l0), i1 = 1.f / quad_sum(l1);
      float de0 = 0.f, de1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        s[nt][0] *= i0;
        s[nt][1] *= i0;
        s[nt][2] *= i1;
        s[nt][3] *= i1;
        de0 += s[nt][0] * dp[nt][0] + s[nt][1] * dp[nt][1];
        de1 += s[nt][2] * dp[nt][2] + s[nt][3] * dp[nt][3];
      }
      de0 = quad_sum(de0);
      de1 = quad_sum(de1);
      if (t == 0) {
        st[g * 3] = m0;
        st[g * 3 + 1] = i0;
        st[g * 3 + 2] = de0;
        st[(g + 8) * 3] = m1;
        st[(g + 8) * 3 + 1] = i1;
        st[(g + 8) * 3 + 2] = de1;
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        s[nt][0] *= (dp[nt][0] - de0) * sfs;
        s[nt][1] *= (dp[nt][1] - de0) * sfs;
        s[nt][2] *= (dp[nt][2] - de1) * sfs;
        s[nt][3] *= (dp[nt][3] - de1) * sfs;
      }
      float dqa[DC][4];
#pragma unroll
      for (int nt = 0; nt < DC; ++nt) dqa[nt][0] = dqa[nt][1] = dqa[nt][2] = dqa[nt][3] = 0.f;
#pragma unroll
      for (int kc = 0; kc < 2; ++kc) {
        uint32_t da[4] = {pack2(s[2 * kc][0], s[2 * kc][1]), pack2(s[2 * kc][2], s[2 * kc][3]),
                          pack2(s[2 * kc + 1][0], s[2 * kc + 1][1]), pack2(s[2 * kc + 1][2], s[2 * kc + 1][3])};
#pragma unroll
        for (int nt = 0; nt < DC; ++nt) mma_T(dqa[nt], da, Kth, kNKP, nt * 8, kc * 16, g, t);
      }
      const long long grow = static_cast<long long>(b) * N + q0 + g;
#pragma unroll
      for (int nt = 0; nt < DC; ++nt) {
        if (q0 + g < N)
          *reinterpret_cast<uint32_t*>(dq + grow * C + h * D + nt * 8 + 2 * t) = pack2(dqa[nt][0] * finv, dqa[nt][1] * finv);
        if (q0 + g + 8 < N)
          *reinterpret_cast<uint32_t*>(dq + (grow + 8) * C + h * D + nt * 8 + 2 * t) = pack2(dqa[nt][2] * finv, dqa[nt][3] * finv);
      }
    }
    __syncwarp();
    // ---- this warp's 16 queries' contribution to dK, dV: transposed roles (keys are the MMA rows) ----
    float* accs = acc + warp * kNK * 2 * D;     // written by this warp only: plain stores, summed over the head's warps below
#pragma unroll 1
    for (int kt = 0; kt < kNK; kt += 16) {
      AFrag<D> ka, va;
      ka.load(Kh, kt, g, t);
      va.load(Vh, kt, g, t);
      float pt[2][4], ds[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        float dpt[4] = {0.f, 0.f, 0.f, 0.f};
        pt[i][0] = pt[i][1] = pt[i][2] = pt[i][3] = 0.f;
        mma_rowsT<D>(pt[i], ka, Qh, q0 + 8 * i, g, t);      // S^T[key kt+g / +8][query q0 + 8i + 2t / +1]
        mma_rowsT<D>(dpt, va, Oh, q0 + 8 * i, g, t);
        const float* sa = st + (8 * i + 2 * t) * 3;
        const float ma = sa[0], ia = sa[1], dea = sa[2], mb = sa[3], ib = sa[4], deb = sa[5];
        const bool k0ok = kt + g < T, k1ok = kt + g + 8 < T;
        pt[i][0] = k0ok ? ex2_approx(pt[i][0] * scale_log2 - ma) * ia : 0.f;
        pt[i][1] = k0ok ? ex2_approx(pt[i][1] * scale_log2 - mb) * ib : 0.f;
        pt[i][2] = k1ok ? ex2_approx(pt[i][2] * scale_log2 - ma) * ia : 0.f;
        pt[i][3] = k1ok ? ex2_approx(pt[i][3] * scale_log2 - mb) * ib : 0.f;
        ds[i][0] = pt[i][0] * (dpt[0] - dea) * sfs;
        ds[i][1] = pt[i][1] * (dpt[1] - deb) * sfs;
        ds[i][2] = pt[i][2] * (dpt[2] - dea) * sfs;
        ds[i][3] = pt[i][3] * (dpt[3] - deb) * sfs;
      }
      uint32_t pa[4] = {pack2(pt[0][0], pt[0][1]), pack2(pt[0][2], pt[0][3]), pack2(pt[1][0], pt[1][1]), pack2(pt[1][2], pt[1][3])};
      uint32_t da[4] = {pack2(ds[0][0], ds[0][1]), pack2(ds[0][2], ds[0][3]), pack2(ds[1][0], ds[1][1]), pack2(ds[1][2], ds[1][3])};
#pragma unroll
      for (int nt = 0; nt < DC; ++nt) {
        float dv[4] = {0.f, 0.f, 0.f, 0.f}, dk[4] = {0.f, 0.f, 0.f, 0.f};
        mma_T(dv, pa, Oth, NP, nt * 8, q0, g, t);
        mma_T(dk, da, Qth, NP, nt * 8, q0, g, t);
        float* r0 = accs + (kt + g) * 2 * D + nt * 8 + 2 * t;
        float* r1 = r0 + 8 * 2 * D;
        *reinterpret_cast<float2*>(r0) = make_float2(dk[0], dk[1]);
        *reinterpret_cast<float2*>(r1) = make_float2(dk[2], dk[3]);
        *reinterpret_cast<float2*>(r0 + D) = make_float2(dv[0], dv[1]);
        *reinterpret_cast<float2*>(r1 + D) = make_float2(dv[2], dv[3]);
      }
    }
  }
  __syncthreads();
  float* dst = dkv + static_cast<long long>(b) * T * 2 * C;
  for (int idx = threadIdx.x; idx < G * T * 2 * D; idx += blockDim.x) {
    const int sl = idx / (T * 2 * D);
    const int rem = idx - sl * T * 2 * D;
    const int j = rem / (2 * D), c = rem - j * 2 * D;
    const int hh = hg * G + sl;
    float v = 0.f;
    for (int w = 0; w < qtiles; ++w) v += acc[((sl * qtiles + w) * kNK + j) * 2 * D + c];
    if (c < D) v *= finv;      // dK came through the scaled dS; dV = P^T dO did not
    if (c < D) dst[static_cast<long long>(j) * 2 * C + hh * D + c] = v;
    else dst[static_cast<long long>(j) * 2 * C + C + hh * D + (c - D)] = v;
  }
}

template <int D>
static int cross_fwd_mma_t(const h16* q, const float* kv, h16* out, int B, int N, int heads, int T, cudaStream_t st) {
  const int NT = (N + 15) & ~15;
  const int G = mma_group(NT, heads);
  const size_t smem = static_cast<size_t>(G) * (NT * D + kNK * D + D * kNKP) * sizeof(h16);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_fwd_mma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  const int threads = G * (NT / 16) * 32;
  launch_k(cross_attn_fwd_mma_kernel<D>, B * (heads / G), threads, smem, st, q, kv, out, N, heads, G, T,
           rsqrtf(static_cast<float>(D)) * kLog2eM);
  return static_cast<int>(cudaGetLastError());
}

template <int D>
static int cross_bwd_mma_t(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads,
                           int T, cudaStream_t st) {
  const int NT = (N + 15) & ~15;
  const int G = mma_group(NT, heads);
  const int warps = G * (NT / 16);
  const size_t smem = static_cast<size_t>(G) * (2 * NT * D + 2 * D * (NT + 8) + 2 * kNK * D + D * kNKP) * sizeof(h16) +
                      (static_cast<size_t>(warps) * kNK * 2 * D + warps * 48) * sizeof(float);
  static size_t configured = 48 * 1024;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(cross_attn_bwd_mma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = smem;
  }
  launch_k(cross_attn_bwd_mma_kernel<D>, B * (heads / G), warps * 32, smem, st, q, kv, dout, dq, dkv, N, heads, G, T,
           rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

int launch_cross_attn_fwd_mma(const h16* q, const float* kv, h16* out, int B, int N, int heads, int D, int T, cudaStream_t s) {
  if (!kv || N < 1 || N > 256 || T < 1 || T > kNK) return -1;
  switch (D) {
    case 8: return cross_fwd_mma_t<8>(q, kv, out, B, N, heads, T, s);
    case 16: return cross_fwd_mma_t<16>(q, kv, out, B, N, heads, T, s);
    case 32: return cross_fwd_mma_t<32>(q, kv, out, B, N, heads, T, s);
    default: return -1;
  }
}

int launch_cross_attn_bwd_mma(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads,
                              int D, int T, cudaStream_t s) {
  if (!kv || N < 1 || N > 256 || T < 1 || T > kNK) return -1;
  switch (D) {
    case 8: return cross_bwd_mma_t<8>(q, kv, dout, dq, dkv, B, N, heads, T, s);
    case 16: return cross_bwd_mma_t<16>(q, kv, dout, dq, dkv, B, N, heads, T, s);
    case 32: return cross_bwd_mma_t<32>(q, kv, dout, dq, dkv, B, N, heads, T, s);
    default: return -1;
  }
}

}  // namespace edb
