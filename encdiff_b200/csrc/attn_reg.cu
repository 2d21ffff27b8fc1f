// Cross-attention against T <= 32 concept tokens, register-resident (no shared-memory staging, no CTA barrier in the
// forward): reference CrossAttention.forward with a context, ldm/modules/attention.py:170-193, and its autograd.
//
// Why a second implementation next to attn_mma.cu: ncu (profiles/r02_attn_ncu.txt) showed the staged kernels to be
// instruction-issue bound on bookkeeping, not on the attention math -- 1 196 warp-instructions per (head, 16-query tile)
// in the backward, of which ~170 are the math: index arithmetic with integer divisions, transposed 2-byte staging
// stores, per-warp partial sums reduced through shared memory behind three CTA barriers.  Here a warp owns one
// (sample, head[, token split]) and walks its 16-query tiles with everything loop-invariant hoisted into registers:
//   * Q / dO tiles are loaded from global memory straight into mma.sync A-fragments (all tiles of the warp up front:
//     every load of the warp is in flight at once);
//   * the head's K / V (T x D, fp32 in `kv`) are converted once into the B-fragments of S = Q K^T, dP = dO V^T,
//     O = P V and dQ = dS K;
//   * dK / dV accumulate in MMA accumulators across the warp's tiles (tokens are the k axis of dV = P^T dO,
//     dK = dS^T Q), so no cross-warp reduction exists unless the tokens of a head are split over two warps;
//   * the transposed operands (P^T, dS^T as A-fragments; dO^T, Q^T as B-fragments) are made with movmatrix from the
//     registers that already hold P, dS, dO, Q: the exponentials are computed once and nothing is recomputed.
// Padded keys (>= T) are masked out of the softmax; padded query rows (>= N, only when N is not a multiple of 16) are
// loaded as zeros and never stored.
#include "common.cuh"
#include "kernels.cuh"

namespace edb {
namespace {

constexpr float kLog2eR = 1.4426950408889634f;

#ifdef EDB_ACT_BF16
#define EDB_MMA_TR "bf16"
#else
#define EDB_MMA_TR "f16"
#endif

__device__ __forceinline__ void rmma_k8(float* c, uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32." EDB_MMA_TR "." EDB_MMA_TR ".f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void rmma_k16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." EDB_MMA_TR "." EDB_MMA_TR ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// transpose of an 8x8 16-bit matrix held one register per lane in fragment layout (row = lane / 4, columns 2 (lane % 4), +1)
__device__ __forceinline__ uint32_t movm_t(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
__device__ __forceinline__ uint32_t rpack2(float a, float b) {
  const h162 v = f22h2(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t ldg32(const h16* p) { return __ldg(reinterpret_cast<const unsigned int*>(p)); }
__device__ __forceinline__ float rquad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float rquad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

template <int D>
struct Frag {
  static constexpr int NR = D == 8 ? 2 : D / 4;    // A-fragment registers of a [16 x D] tile (k = the D axis)
  static constexpr int NB = D == 8 ? 1 : D / 8;    // B-fragment registers of one 8-key n-tile over the whole D axis
  static constexpr int ND = D / 8;                 // 8-wide n-tiles along D
};

// A-fragment of rows row0 + g and row0 + g + 8 of X [rows][C], columns col0 .. col0 + D - 1 (same order as AFrag<D> of
// attn_mma.cu); rows that are not ok are zeros
template <int D>
__device__ __forceinline__ void load_afrag(uint32_t* r, const h16* __restrict__ X, long long row0, int C, int col0, int g,
                                           int t, bool ok0, bool ok1) {
  const h16* p0 = X + (row0 + g) * C + col0 + 2 * t;
  const h16* p1 = p0 + 8 * C;
  if (D == 8) {
    r[0] = ok0 ? ldg32(p0) : 0u;
    r[1] = ok1 ? ldg32(p1) : 0u;
  } else {
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      r[4 * ks] = ok0 ? ldg32(p0 + 16 * ks) : 0u;
      r[4 * ks + 1] = ok1 ? ldg32(p1 + 16 * ks) : 0u;
      r[4 * ks + 2] = ok0 ? ldg32(p0 + 16 * ks + 8) : 0u;
      r[4 * ks + 3] = ok1 ? ldg32(p1 + 16 * ks + 8) : 0u;
    }
  }
}

// B-fragments of  X Y^T  (k = d, n = key) for the 8 keys nt*8 .. nt*8+7 of Y = the head's [T][D] slice at `y`
// (fp32, row pitch ld): lane (g, t) holds Y[nt*8+g][2t, 2t+1] (+8 for the upper k half of a k16 step)
template <int D>
__device__ __forceinline__ void load_bfrag_rows(uint32_t* r, const float* __restrict__ y, int ld, int nt, int T, int g, int t) {
  const int key = nt * 8 + g;
  const bool ok = key < T;
  const float* p = y + static_cast<long long>(key) * ld + 2 * t;
  if (D == 8) {
    const float2 v = ok ? __ldg(reinterpret_cast<const float2*>(p)) : make_float2(0.f, 0.f);
    r[0] = rpack2(v.x, v.y);
  } else {
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      const float2 v0 = ok ? __ldg(reinterpret_cast<const float2*>(p + 16 * ks)) : make_float2(0.f, 0.f);
      const float2 v1 = ok ? __ldg(reinterpret_cast<const float2*>(p + 16 * ks + 8)) : make_float2(0.f, 0.f);
      r[2 * ks] = rpack2(v0.x, v0.y);
      r[2 * ks + 1] = rpack2(v1.x, v1.y);
    }
  }
}

// B-fragments of  P Y  (k = key, n = d): r[kt][nd][j] = (Y[16kt + 2t + 8j][8nd + g], Y[16kt + 2t + 8j + 1][8nd + g])
template <int D>
__device__ __forceinline__ void load_bfrag_cols(uint32_t (*r)[D / 8][2], const float* __restrict__ y, int ld, int T, int g, int t) {
#pragma unroll
  for (int kt = 0; kt < 2; ++kt)
#pragma unroll
    for (int nd = 0; nd < D / 8; ++nd)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int key = 16 * kt + 2 * t + 8 * j;
        const float* p = y + static_cast<long long>(key) * ld + 8 * nd + g;
        const float v0 = key < T ? __ldg(p) : 0.f;
        const float v1 = key + 1 < T ? __ldg(p + ld) : 0.f;
        r[kt][nd][j] = rpack2(v0, v1);
      }
}

// S[16 x 8*NKT] = A(16 x D) * Yb^T for the NKT key n-tiles
template <int D, int NKT>
__device__ __forceinline__ void scores(float (*s)[4], const uint32_t* a, const uint32_t (*yb)[Frag<D>::NB]) {
#pragma unroll
  for (int nt = 0; nt < NKT; ++nt) {
    s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
    if (D == 8) {
      rmma_k8(s[nt], a[0], a[1], yb[nt][0]);
    } else {
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) rmma_k16(s[nt], a + 4 * ks, yb[nt][2 * ks], yb[nt][2 * ks + 1]);
    }
  }
}

// softmax over the 8*NKT (masked) keys of rows g and g+8: s <- exp2(s * scale_log2 - max); returns 1 / rowsum
template <int NKT>
__device__ __forceinline__ void softmax_rows(float (*s)[4], int T, int t, float scale_log2, float& inv0, float& inv1) {
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int nt = 0; nt < NKT; ++nt) {
    if (nt * 8 + 8 > T) {       // only the last tile(s) can hold padded keys
      const int key = nt * 8 + 2 * t;
      if (key >= T) s[nt][0] = s[nt][2] = -INFINITY;
      if (key + 1 >= T) s[nt][1] = s[nt][3] = -INFINITY;
    }
    m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
    m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
  }
  m0 = rquad_max(m0) * scale_log2;
  m1 = rquad_max(m1) * scale_log2;
  float l0 = 0.f, l1 = 0.f;
#pragma unroll
  for (int nt = 0; nt < NKT; ++nt) {
    s[nt][0] = ex2_approx(s[nt][0] * scale_log2 - m0);
    s[nt][1] = ex2_approx(s[nt][1] * scale_log2 - m0);
    s[nt][2] = ex2_approx(s[nt][2] * scale_log2 - m1);
    s[nt][3] = ex2_approx(s[nt][3] * scale_log2 - m1);
    l0 += s[nt][0] + s[nt][1];
    l1 += s[nt][2] + s[nt][3];
  }
  inv0 = 1.f / rquad_sum(l0);
  inv1 = 1.f / rquad_sum(l1);
}

// A-fragments (k = 16 keys per kt) of a [16 x 32] row-major accumulator set (missing n-tiles are zeros)
template <int NKT>
__device__ __forceinline__ void pack_afrags(uint32_t (*pa)[4], const float (*p)[4]) {
#pragma unroll
  for (int kt = 0; kt < 2; ++kt) {
    pa[kt][0] = 2 * kt < NKT ? rpack2(p[2 * kt][0], p[2 * kt][1]) : 0u;
    pa[kt][1] = 2 * kt < NKT ? rpack2(p[2 * kt][2], p[2 * kt][3]) : 0u;
    pa[kt][2] = 2 * kt + 1 < NKT ? rpack2(p[2 * kt + 1][0], p[2 * kt + 1][1]) : 0u;
    pa[kt][3] = 2 * kt + 1 < NKT ? rpack2(p[2 * kt + 1][2], p[2 * kt + 1][3]) : 0u;
  }
}

struct WarpItem {
  int b, h, sp;
};
template <int SPLIT>
__device__ __forceinline__ WarpItem warp_item(int w, int heads) {
  WarpItem it;
  it.sp = w % SPLIT;
  const int bh = w / SPLIT;
  it.h = bh % heads;
  it.b = bh / heads;
  return it;
}

// ---------------------------------------------------------------------------------------------------------------
// forward: out = softmax(Q K^T * D^-1/2) V.   q, out [B*N, C];  kv fp32 [B, T, 2C] (K | V, head h at columns h*D).
// warp = (sample, head, split); 4 warps per CTA.
// ---------------------------------------------------------------------------------------------------------------
template <int D, int TILES, int SPLIT, int NKT>
__global__ void __launch_bounds__(128) cross_attn_fwd_reg_kernel(const h16* __restrict__ q, const float* __restrict__ kv,
                                                                 h16* __restrict__ out, int N, int heads, int T,
                                                                 int total_warps, float scale_log2) {
  pdl_enter();
  using F = Frag<D>;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int w = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (w >= total_warps) return;
  const WarpItem it = warp_item<SPLIT>(w, heads);
  const int C = heads * D;
  const long long rowb = static_cast<long long>(it.b) * N;
  uint32_t qf[TILES][F::NR];
#pragma unroll
  for (int tl = 0; tl < TILES; ++tl) {
    const int n0 = (it.sp * TILES + tl) * 16;
    load_afrag<D>(qf[tl], q, rowb + n0, C, it.h * D, g, t, n0 + g < N, n0 + g + 8 < N);
  }
  const float* kvb = kv + static_cast<long long>(it.b) * T * 2 * C + it.h * D;
  uint32_t kb[NKT][F::NB];
#pragma unroll
  for (int nt = 0; nt < NKT; ++nt) load_bfrag_rows<D>(kb[nt], kvb, 2 * C, nt, T, g, t);
  uint32_t vk[2][F::ND][2];
  load_bfrag_cols<D>(vk, kvb + C, 2 * C, T, g, t);
#pragma unroll
  for (int tl = 0; tl < TILES; ++tl) {
    const int n0 = (it.sp * TILES + tl) * 16;
    if (n0 >= N) break;
    float s[NKT][4];
    scores<D, NKT>(s, qf[tl], kb);
    float inv0, inv1;
    softmax_rows<NKT>(s, T, t, scale_log2, inv0, inv1);
    uint32_t pa[2][4];
    pack_afrags<NKT>(pa, s);
    float o[F::ND][4];
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd) {
      o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f;
#pragma unroll
      for (int kt = 0; kt < 2; ++kt)
        if (2 * kt < NKT) rmma_k16(o[nd], pa[kt], vk[kt][nd][0], vk[kt][nd][1]);
    }
    h16* o0 = out + (rowb + n0 + g) * C + it.h * D + 2 * t;
    h16* o1 = o0 + 8 * C;
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd) {
      if (n0 + g < N) *reinterpret_cast<uint32_t*>(o0 + 8 * nd) = rpack2(o[nd][0] * inv0, o[nd][1] * inv0);
      if (n0 + g + 8 < N) *reinterpret_cast<uint32_t*>(o1 + 8 * nd) = rpack2(o[nd][2] * inv1, o[nd][3] * inv1);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// backward: dq [B*N, C] and dkv fp32 [B, T, 2C] (dK | dV, plain stores: every (sample, head) slice has one owner).
// dS is carried multiplied by a per-warp power of two f (max |f dO| in [0.5, 1)) so that its fp16 fragments stay in the
// normal range whatever the loss scale; dq and dK are multiplied back by 1/f (and by the softmax scale) at the end.
// ---------------------------------------------------------------------------------------------------------------
template <int D, int TILES, int SPLIT, int NKT>
__global__ void __launch_bounds__(128) cross_attn_bwd_reg_kernel(const h16* __restrict__ q, const float* __restrict__ kv,
                                                                 const h16* __restrict__ dout, h16* __restrict__ dq,
                                                                 float* __restrict__ dkv, int N, int heads, int T,
                                                                 int total_warps, float scale) {
  pdl_enter();
  using F = Frag<D>;
  constexpr int NACC = 2 * F::ND * 4;                 // accumulator floats of dK (and of dV) per lane
  __shared__ float red[SPLIT > 1 ? 4 * 2 * NACC * 32 : 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int w = blockIdx.x * 4 + warp;                // total_warps is a multiple of 4 (launcher): no partial CTA
  if (w >= total_warps) return;
  const WarpItem it = warp_item<SPLIT>(w, heads);
  const int C = heads * D;
  const long long rowb = static_cast<long long>(it.b) * N;
  uint32_t qf[TILES][F::NR], of[TILES][F::NR];
#pragma unroll
  for (int tl = 0; tl < TILES; ++tl) {
    const int n0 = (it.sp * TILES + tl) * 16;
    load_afrag<D>(qf[tl], q, rowb + n0, C, it.h * D, g, t, n0 + g < N, n0 + g + 8 < N);
    load_afrag<D>(of[tl], dout, rowb + n0, C, it.h * D, g, t, n0 + g < N, n0 + g + 8 < N);
  }
  const float* kvb = kv + static_cast<long long>(it.b) * T * 2 * C + it.h * D;
  uint32_t kb[NKT][F::NB], vb[NKT][F::NB];
#pragma unroll
  for (int nt = 0; nt < NKT; ++nt) {
    load_bfrag_rows<D>(kb[nt], kvb, 2 * C, nt, T, g, t);
    load_bfrag_rows<D>(vb[nt], kvb + C, 2 * C, nt, T, g, t);
  }
  uint32_t kk[2][F::ND][2];
  load_bfrag_cols<D>(kk, kvb, 2 * C, T, g, t);
  // power-of-two range scale from the warp's max |dO|
  float amax = 0.f;
#pragma unroll
  for (int tl = 0; tl < TILES; ++tl)
#pragma unroll
    for (int i = 0; i < F::NR; ++i) {
      const float2 v = h22f2(*reinterpret_cast<const h162*>(&of[tl][i]));
      amax = fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y)));
    }
  amax = warp_max(amax);
  float fsc = 1.f;
  if (amax > 0.f && isfinite(amax)) {
    int e;
    frexpf(amax, &e);
    fsc = ldexpf(1.f, -e);
  }
  const float back = scale / fsc;                     // dq, dK = (dS f)(..) * scale / f
  const float scale_log2 = scale * kLog2eR;
  float dk[2][F::ND][4], dv[2][F::ND][4];
#pragma unroll
  for (int kt = 0; kt < 2; ++kt)
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd)
#pragma unroll
      for (int i = 0; i < 4; ++i) dk[kt][nd][i] = dv[kt][nd][i] = 0.f;
#pragma unroll
  for (int tl = 0; tl < TILES; ++tl) {
    const int n0 = (it.sp * TILES + tl) * 16;
    if (n0 >= N) break;
    float p[NKT][4], dp[NKT][4];
    scores<D, NKT>(p, qf[tl], kb);
    scores<D, NKT>(dp, of[tl], vb);
    float inv0, inv1;
    softmax_rows<NKT>(p, T, t, scale_log2, inv0, inv1);
    float de0 = 0.f, de1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NKT; ++nt) {
      p[nt][0] *= inv0;
      p[nt][1] *= inv0;
      p[nt][2] *= inv1;
      p[nt][3] *= inv1;
      de0 += p[nt][0] * dp[nt][0] + p[nt][1] * dp[nt][1];
      de1 += p[nt][2] * dp[nt][2] + p[nt][3] * dp[nt][3];
    }
    de0 = rquad_sum(de0);
    de1 = rquad_sum(de1);
    uint32_t pa[2][4], da[2][4];
    pack_afrags<NKT>(pa, p);
#pragma unroll
    for (int nt = 0; nt < NKT; ++nt) {     // dp <- dS * f
      dp[nt][0] = p[nt][0] * (dp[nt][0] - de0) * fsc;
      dp[nt][1] = p[nt][1] * (dp[nt][1] - de0) * fsc;
      dp[nt][2] = p[nt][2] * (dp[nt][2] - de1) * fsc;
      dp[nt][3] = p[nt][3] * (dp[nt][3] - de1) * fsc;
    }
    pack_afrags<NKT>(da, dp);
    // dQ = dS K
    float dqa[F::ND][4];
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd) {
      dqa[nd][0] = dqa[nd][1] = dqa[nd][2] = dqa[nd][3] = 0.f;
#pragma unroll
      for (int kt = 0; kt < 2; ++kt)
        if (2 * kt < NKT) rmma_k16(dqa[nd], da[kt], kk[kt][nd][0], kk[kt][nd][1]);
    }
    h16* q0 = dq + (rowb + n0 + g) * C + it.h * D + 2 * t;
    h16* q1 = q0 + 8 * C;
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd) {
      if (n0 + g < N) *reinterpret_cast<uint32_t*>(q0 + 8 * nd) = rpack2(dqa[nd][0] * back, dqa[nd][1] * back);
      if (n0 + g + 8 < N) *reinterpret_cast<uint32_t*>(q1 + 8 * nd) = rpack2(dqa[nd][2] * back, dqa[nd][3] * back);
    }
    // dV += P^T dO, dK += dS^T Q: keys are the MMA rows, the tile's 16 queries the k axis.
    // B-fragments (k = query, n = d) are the transposes of the 8x8 blocks of the dO / Q A-fragments.
    uint32_t ot[F::ND][2], qt[F::ND][2];
    if (D == 8) {
      ot[0][0] = movm_t(of[tl][0]);
      ot[0][1] = movm_t(of[tl][1]);
      qt[0][0] = movm_t(qf[tl][0]);
      qt[0][1] = movm_t(qf[tl][1]);
    } else {
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
        ot[2 * ks][0] = movm_t(of[tl][4 * ks]);
        ot[2 * ks][1] = movm_t(of[tl][4 * ks + 1]);
        ot[2 * ks + 1][0] = movm_t(of[tl][4 * ks + 2]);
        ot[2 * ks + 1][1] = movm_t(of[tl][4 * ks + 3]);
        qt[2 * ks][0] = movm_t(qf[tl][4 * ks]);
        qt[2 * ks][1] = movm_t(qf[tl][4 * ks + 1]);
        qt[2 * ks + 1][0] = movm_t(qf[tl][4 * ks + 2]);
        qt[2 * ks + 1][1] = movm_t(qf[tl][4 * ks + 3]);
      }
    }
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) {
      if (2 * kt >= NKT) continue;
      // A-fragment of X^T (rows = keys 16kt.., k = queries): {T(x[0]), T(x[2]), T(x[1]), T(x[3])}
      uint32_t pt[4], dt[4];
      pt[0] = movm_t(pa[kt][0]);
      pt[2] = movm_t(pa[kt][1]);
      dt[0] = movm_t(da[kt][0]);
      dt[2] = movm_t(da[kt][1]);
      if (2 * kt + 1 < NKT) {
        pt[1] = movm_t(pa[kt][2]);
        pt[3] = movm_t(pa[kt][3]);
        dt[1] = movm_t(da[kt][2]);
        dt[3] = movm_t(da[kt][3]);
      } else {
        pt[1] = pt[3] = dt[1] = dt[3] = 0u;
      }
#pragma unroll
      for (int nd = 0; nd < F::ND; ++nd) {
        rmma_k16(dv[kt][nd], pt, ot[nd][0], ot[nd][1]);
        rmma_k16(dk[kt][nd], dt, qt[nd][0], qt[nd][1]);
      }
    }
  }
  if (SPLIT > 1) {      // the splits of a (sample, head) are consecutive warps of this CTA
    float* mine = red + warp * 2 * NACC * 32;
    if (it.sp != 0) {
#pragma unroll
      for (int kt = 0; kt < 2; ++kt)
#pragma unroll
        for (int nd = 0; nd < F::ND; ++nd)
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int idx = (kt * F::ND + nd) * 4 + i;
            mine[idx * 32 + lane] = dk[kt][nd][i] * back;
            mine[(NACC + idx) * 32 + lane] = dv[kt][nd][i];
          }
    }
    __syncthreads();
    if (it.sp != 0) return;
#pragma unroll
    for (int kt = 0; kt < 2; ++kt)
#pragma unroll
      for (int nd = 0; nd < F::ND; ++nd)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int idx = (kt * F::ND + nd) * 4 + i;
          float ak = dk[kt][nd][i] * back, av = dv[kt][nd][i];
#pragma unroll
          for (int o = 1; o < SPLIT; ++o) {
            ak += red[((warp + o) * 2 * NACC + idx) * 32 + lane];
            av += red[((warp + o) * 2 * NACC + NACC + idx) * 32 + lane];
          }
          dk[kt][nd][i] = ak;
          dv[kt][nd][i] = av;
        }
  } else {
#pragma unroll
    for (int kt = 0; kt < 2; ++kt)
#pragma unroll
      for (int nd = 0; nd < F::ND; ++nd)
#pragma unroll
        for (int i = 0; i < 4; ++i) dk[kt][nd][i] *= back;
  }
  float* dst = dkv + static_cast<long long>(it.b) * T * 2 * C + it.h * D + 2 * t;
#pragma unroll
  for (int kt = 0; kt < 2; ++kt)
#pragma unroll
    for (int nd = 0; nd < F::ND; ++nd) {
      const int k0 = 16 * kt + g, k1 = k0 + 8;
      if (k0 < T) {
        float* r = dst + static_cast<long long>(k0) * 2 * C + 8 * nd;
        *reinterpret_cast<float2*>(r) = make_float2(dk[kt][nd][0], dk[kt][nd][1]);
        *reinterpret_cast<float2*>(r + C) = make_float2(dv[kt][nd][0], dv[kt][nd][1]);
      }
      if (k1 < T) {
        float* r = dst + static_cast<long long>(k1) * 2 * C + 8 * nd;
        *reinterpret_cast<float2*>(r) = make_float2(dk[kt][nd][2], dk[kt][nd][3]);
        *reinterpret_cast<float2*>(r + C) = make_float2(dv[kt][nd][2], dv[kt][nd][3]);
      }
    }
}

template <int D, int TILES, int SPLIT, int NKT>
int fwd_reg_t(const h16* q, const float* kv, h16* out, int B, int N, int heads, int T, cudaStream_t st) {
  const int warps = B * heads * SPLIT;
  launch_k(cross_attn_fwd_reg_kernel<D, TILES, SPLIT, NKT>, warps / 4, 128, 0, st, q, kv, out, N, heads, T, warps,
           rsqrtf(static_cast<float>(D)) * kLog2eR);
  return static_cast<int>(cudaGetLastError());
}
template <int D, int TILES, int SPLIT, int NKT>
int bwd_reg_t(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads, int T,
              cudaStream_t st) {
  const int warps = B * heads * SPLIT;
  launch_k(cross_attn_bwd_reg_kernel<D, TILES, SPLIT, NKT>, warps / 4, 128, 0, st, q, kv, dout, dq, dkv, N, heads, T, warps,
           rsqrtf(static_cast<float>(D)));
  return static_cast<int>(cudaGetLastError());
}

// (TILES, SPLIT) by the number of 16-query tiles of a sample: 1 and 4 tiles run on one warp per (sample, head); 16 tiles
// (N = 256) are split over two warps with 8 preloaded tiles each (d_head 8 and 16 only: register budget).
#define EDB_CROSS_REG_DISPATCH(FN, D, ...)                                                          \
  do {                                                                                              \
    if (tiles == 1) return nkt == 3 ? FN<D, 1, 1, 3>(__VA_ARGS__) : FN<D, 1, 1, 4>(__VA_ARGS__);    \
    if (tiles == 4) return nkt == 3 ? FN<D, 4, 1, 3>(__VA_ARGS__) : FN<D, 4, 1, 4>(__VA_ARGS__);    \
  } while (0)
#define EDB_CROSS_REG_DISPATCH16(FN, D, ...)                                                        \
  do {                                                                                              \
    if (tiles == 16) return nkt == 3 ? FN<D, 8, 2, 3>(__VA_ARGS__) : FN<D, 8, 2, 4>(__VA_ARGS__);   \
  } while (0)

}  // namespace

// returns -1 when the shape has no register-resident variant (the caller keeps the staged kernels of attn_mma.cu)
int launch_cross_attn_fwd_reg(const h16* q, const float* kv, h16* out, int B, int N, int heads, int D, int T, cudaStream_t s) {
  if (!kv || N < 1 || T < 1 || T > 32 || (heads & 3) || B < 1) return -1;
  const int nkt = T <= 24 ? 3 : 4, tiles = (N + 15) / 16;
  switch (D) {
    case 8:
      EDB_CROSS_REG_DISPATCH(fwd_reg_t, 8, q, kv, out, B, N, heads, T, s);
      EDB_CROSS_REG_DISPATCH16(fwd_reg_t, 8, q, kv, out, B, N, heads, T, s);
      return -1;
    case 16:
      EDB_CROSS_REG_DISPATCH(fwd_reg_t, 16, q, kv, out, B, N, heads, T, s);
      EDB_CROSS_REG_DISPATCH16(fwd_reg_t, 16, q, kv, out, B, N, heads, T, s);
      return -1;
    case 32:
      EDB_CROSS_REG_DISPATCH(fwd_reg_t, 32, q, kv, out, B, N, heads, T, s);
      return -1;
    default: return -1;
  }
}

int launch_cross_attn_bwd_reg(const h16* q, const float* kv, const h16* dout, h16* dq, float* dkv, int B, int N, int heads,
                              int D, int T, cudaStream_t s) {
  if (!kv || N < 1 || T < 1 || T > 32 || (heads & 3) || B < 1) return -1;
  const int nkt = T <= 24 ? 3 : 4, tiles = (N + 15) / 16;
  switch (D) {
    case 8:
      EDB_CROSS_REG_DISPATCH(bwd_reg_t, 8, q, kv, dout, dq, dkv, B, N, heads, T, s);
      EDB_CROSS_REG_DISPATCH16(bwd_reg_t, 8, q, kv, dout, dq, dkv, B, N, heads, T, s);
      return -1;
    case 16:
      EDB_CROSS_REG_DISPATCH(bwd_reg_t, 16, q, kv, dout, dq, dkv, B, N, heads, T, s);
      EDB_CROSS_REG_DISPATCH16(bwd_reg_t, 16, q, kv, dout, dq, dkv, B, N, heads, T, s);
      return -1;
    case 32:
      EDB_CROSS_REG_DISPATCH(bwd_reg_t, 32, q, kv, dout, dq, dkv, B, N, heads, T, s);
      return -1;
    default: return -1;
  }
}

}  // namespace edb
