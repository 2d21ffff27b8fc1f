// tcgen05 / TMEM / TMA GEMM kernels of the EncDiff B200 denoiser (sm_100a).  See gemm.cuh for the math.
//
// Warp roles (192 threads): warp 0 = TMA producer (one elected lane), warp 1 = MMA issuer (one elected lane),
// warps 2..5 = epilogue (TMEM lane quarter = warp % 4); warp 2 also owns the TMEM allocation.
#include "common.cuh"
#include "gemm.cuh"

namespace edb {

static constexpr int kTileM = 128;   // pixels per tile (UMMA M)
static constexpr int kChunk = 64;    // h16 elements per 128-byte swizzle row
static constexpr int kABytes = kTileM * kChunk * 2;  // 16 KB: one [128 x 64] h16 box

__device__ __forceinline__ void tile_origin(const TileGeom& g, int tile, int& b0, int& y0) {
  if (g.tiles_per_img > 1) {
    b0 = tile / g.tiles_per_img;
    y0 = (tile - b0 * g.tiles_per_img) * g.box_h;
  } else {
    b0 = tile * g.box_b;
    y0 = 0;
  }
}

// =============================================================================================
// sgemm
// =============================================================================================
template <int BN, int STAGES>
__global__ void __launch_bounds__(192) sgemm_kernel(const __grid_constant__ SgemmParams p) {
  pdl_launch_dependents();
  constexpr int B_BYTES = BN * kChunk * 2;
  constexpr int STAGE_BYTES = kABytes + B_BYTES;
  constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x;
  const int n_tile = blockIdx.y;

  int nkb = 0;
  for (int s = 0; s < p.nsrc; ++s) nkb += p.src_chunks[s] * p.src_taps[s];

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.nsrc; ++s) tma_prefetch_desc(&p.amap[s]);
    tma_prefetch_desc(&p.bmap);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // set-up above touched no global memory; everything below may
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int b0, y0;
      tile_origin(p.g, m_tile, b0, y0);
      int kb = 0;
      for (int s = 0; s < p.nsrc; ++s) {
        const int taps = p.src_taps[s];
        for (int t = 0; t < taps; ++t) {
          int dx = 0, dy = 0;
          if (taps == 9) {
            dy = (t / 3 - 1) * p.tap_sign;
            dx = (t % 3 - 1) * p.tap_sign;
          }
          for (int c = 0; c < p.src_chunks[s]; ++c, ++kb) {
            const int st = kb % STAGES;
            const uint32_t ph = (kb / STAGES) & 1;
            mbar_wait(&empty_bar[st], ph ^ 1);
            uint8_t* sa = smem + st * STAGE_BYTES;
            mbar_arrive_expect_tx(&full_bar[st], STAGE_BYTES);
            tma_load_4d(sa, &p.amap[s], &full_bar[st], c * kChunk, dx, y0 + dy, b0);
            tma_load_2d(sa + kABytes, &p.bmap, &full_bar[st], kb * kChunk, n_tile * BN);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BN < 16 ? 16 : BN, false, false);
      for (int kb = 0; kb < nkb; ++kb) {
        const int st = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(&full_bar[st], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
        const uint64_t da = make_sdesc(sa, 16, 1024);
        const uint64_t db = make_sdesc(sa + kABytes, 16, 1024);
#pragma unroll
        for (int k = 0; k < kChunk / 16; ++k) {
          // advance 16 h16 = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field
          umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
        }
        umma_commit(&empty_bar[st]);
      }
      umma_commit(tmem_full);
    }
  } else {
    const int q = warp & 3;
    const long long row = static_cast<long long>(m_tile) * kTileM + q * 32 + lane;
    const bool row_ok = row < p.M;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      const int n0 = n_tile * BN + c0;
      if (n0 >= p.N) break;  // warp-uniform
      float v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (!row_ok) continue;
      const int nv = min(32, p.N - n0);
      if (p.bias) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (i < nv) v[i] += __ldg(p.bias + n0 + i);
      }
      if (p.act == ACT_SILU) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = silu_f(v[i]);
      }
      if (p.res) {
        const h16* r = reinterpret_cast<const h16*>(p.res) + row * p.ldres + n0;
        if (nv == 32 && (p.ldres & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4*>(r + 8 * j), f);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * j + i] += f[i];
          }
        } else {
          for (int i = 0; i < nv; ++i) v[i] += h2f(r[i]);
        }
      }
      void* dst = p.out0;
      int ld = p.ld0, col = n0;
      if (n0 >= p.split_col) {
        dst = p.out1;
        ld = p.ld1;
        col = n0 - p.split_col;
      }
      if (p.out_fp32) {
        float* o = reinterpret_cast<float*>(dst) + row * ld + col;
        if (nv == 32 && (ld & 3) == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(o + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        } else {
          for (int i = 0; i < nv; ++i) o[i] = v[i];
        }
      } else {
        h16* o = reinterpret_cast<h16*>(dst) + row * ld + col;
        if (nv == 32 && (ld & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(o + 8 * j) = pack8(v + 8 * j);
        } else {
          for (int i = 0; i < nv; ++i) o[i] = f2h(v[i]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <int BN, int STAGES>
static int launch_sgemm_t(const SgemmParams& p, cudaStream_t stream) {
  constexpr int smem = STAGES * (kABytes + BN * kChunk * 2) + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(sgemm_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  dim3 grid((p.M + kTileM - 1) / kTileM, (p.N + BN - 1) / BN, 1);
  launch_k(sgemm_kernel<BN, STAGES>, grid, 192, smem, stream, p);
  return static_cast<int>(cudaGetLastError());
}

int sgemm_pick_bn(int N) {
  if (N <= 32) return 32;
  if (N <= 64) return 64;
  return 128;
}

int launch_sgemm(const SgemmParams& p, int bn, cudaStream_t stream) {
  switch (bn) {
    case 32: return launch_sgemm_t<32, 4>(p, stream);
    case 64: return launch_sgemm_t<64, 4>(p, stream);
    case 128: return launch_sgemm_t<128, 3>(p, stream);
    case 256: return launch_sgemm_t<256, 4>(p, stream);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

// =============================================================================================
// wgrad
// =============================================================================================
template <int BN, int STAGES>
__global__ void __launch_bounds__(192) wgrad_kernel(const __grid_constant__ WgradParams p) {
  pdl_launch_dependents();
  constexpr int P_BYTES = 2 * kABytes;            // two 64-channel boxes of dY -> UMMA M = 128
  constexpr int Q_BYTES = (BN / 64) * kABytes;    // BN/64 boxes of X
  constexpr int STAGE_BYTES = P_BYTES + Q_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int p_tile = blockIdx.x;
  const int q_tiles = (p.Cq + BN - 1) / BN;
  const int tap = blockIdx.y / q_tiles;
  const int q_tile = blockIdx.y - tap * q_tiles;
  const int t_begin = blockIdx.z * p.tiles_per_split;
  const int t_end = min(p.m_tiles, t_begin + p.tiles_per_split);
  const int nsteps = t_end - t_begin;  // >= 1 by construction of the grid

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.pmap);
    tma_prefetch_desc(&p.qmap);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // set-up above touched no global memory; everything below may
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int dx = 0, dy = 0;
      if (p.taps == 9) {
        dy = tap / 3 - 1;
        dx = tap % 3 - 1;
      }
      for (int i = 0; i < nsteps; ++i) {
        const int st = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        int b0, y0;
        tile_origin(p.g, t_begin + i, b0, y0);
        mbar_wait(&empty_bar[st], ph ^ 1);
        uint8_t* sp = smem + st * STAGE_BYTES;
        mbar_arrive_expect_tx(&full_bar[st], STAGE_BYTES);
        tma_load_4d(sp, &p.pmap, &full_bar[st], p.p_off + p_tile * 128, 0, y0, b0);
        tma_load_4d(sp + kABytes, &p.pmap, &full_bar[st], p.p_off + p_tile * 128 + 64, 0, y0, b0);
#pragma unroll
        for (int j = 0; j < BN / 64; ++j)
          tma_load_4d(sp + P_BYTES + j * kABytes, &p.qmap, &full_bar[st], q_tile * BN + j * 64, dx, y0 + dy, b0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BN, true, true);
      for (int i = 0; i < nsteps; ++i) {
        const int st = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&full_bar[st], ph);
        tc_fence_after();
        const uint32_t sp = smem_u32(smem + st * STAGE_BYTES);
#pragma unroll
        for (int k = 0; k < kTileM / 16; ++k) {
          // 16 pixels (the reduction) = 16 rows of 128 B = 2048 B per MMA step
          const uint64_t da = make_sdesc(sp + k * 2048, kABytes, 1024);
          const uint64_t db = make_sdesc(sp + P_BYTES + k * 2048, kABytes, 1024);
          umma_f16(tmem_base, da, db, idesc, (i | k) ? 1u : 0u);
        }
        umma_commit(&empty_bar[st]);
      }
      umma_commit(tmem_full);
    }
  } else {
    const int q = warp & 3;
    const int pc = p_tile * 128 + q * 32 + lane;  // channel on the dY side
    const bool row_ok = pc < p.Cp;
    const float ginv = p.gscale ? __ldg(p.gscale + 1) : 1.f;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      const int qc0 = q_tile * BN + c0;
      if (qc0 >= p.Cq) break;  // warp-uniform
      float v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (!row_ok) continue;
      float* o = p.out + static_cast<long long>(pc) * p.stride_p + static_cast<long long>(tap) * p.stride_tap;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int qc = qc0 + i;
        if (qc < p.Cq) atomicAdd(o + static_cast<long long>(qc + p.q_off) * p.stride_q, v[i] * ginv);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<BN>(tmem_base);
}

template <int BN, int STAGES>
static int launch_wgrad_t(const WgradParams& p0, int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (2 * kABytes + (BN / 64) * kABytes) + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  WgradParams p = p0;
  if (splits < 1) splits = 1;
  if (splits > p.m_tiles) splits = p.m_tiles;
  p.tiles_per_split = (p.m_tiles + splits - 1) / splits;
  splits = (p.m_tiles + p.tiles_per_split - 1) / p.tiles_per_split;  // no empty CTAs
  dim3 grid((p.Cp + 127) / 128, p.taps * ((p.Cq + BN - 1) / BN), splits);
  launch_k(wgrad_kernel<BN, STAGES>, grid, 192, smem, stream, p);
  return static_cast<int>(cudaGetLastError());
}

int launch_wgrad(const WgradParams& p, int bn, int splits, cudaStream_t stream) {
  switch (bn) {
    case 64: return launch_wgrad_t<64, 4>(p, splits, stream);
    case 128: return launch_wgrad_t<128, 3>(p, splits, stream);
    case 256: return launch_wgrad_t<256, 2>(p, splits, stream);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

}  // namespace edb
