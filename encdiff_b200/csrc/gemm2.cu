// Persistent tcgen05 / TMEM / TMA implicit-GEMM kernel (sm_100a): the workhorse behind every conv3x3 / conv1x1 /
// linear forward and input-gradient whose output is 16-bit with N % 64 == 0 (see gemm.cuh for the math; gemm.cu keeps
// the one-tile-per-CTA kernel for fp32 / odd-N outputs).
//
//   grid = min(#tiles, #SMs); CTA c works on tiles c, c + grid, ... (n-tile fastest, so concurrently running CTAs
//   share the A tile in L2).
//   warp 0      TMA producer: A boxes (tap-shifted 4D loads, zero fill = conv padding) + weight boxes into a ring of
//               STAGES shared-memory stages that runs ACROSS tiles (the next tile's loads start while the current
//               tile's MMAs drain).
//   warp 1      MMA issuer: one elected lane, tcgen05.mma into one of TWO TMEM accumulator stages.
//   warps 2..5  epilogue: TMEM -> registers (tcgen05.ld) -> bias / SiLU / residual -> 16-bit -> 128B-swizzled staging
//               panel in shared memory -> TMA store.  Runs concurrently with the mainloop of the next tile.
#include <stdlib.h>

#include "common.cuh"
#include "gemm.cuh"

namespace edb {

static constexpr int kTileM2 = 128;
static constexpr int kChunk2 = 64;
static constexpr int kABytes2 = kTileM2 * kChunk2 * 2;   // 16 KB: one [128 x 64] 16-bit box
static constexpr int kPanelBytes = kTileM2 * 128;        // 16 KB: [128 rows x 64 cols] 16-bit staging panel

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ void tile_origin2(const TileGeom& g, int tile, int& b0, int& y0) {
  if (g.tiles_per_img > 1) {
    b0 = tile / g.tiles_per_img;
    y0 = (tile - b0 * g.tiles_per_img) * g.box_h;
  } else {
    b0 = tile * g.box_b;
    y0 = 0;
  }
}

// KPS = 64-wide k-blocks per pipeline stage: one full/empty barrier hand-shake and one tcgen05.commit per KPS k-blocks
// (the hand-shake costs 100-180 ns and does not overlap with the MMAs, profiles/experiments/r01_gemm_pacing.txt)
template <int BN, int STAGES, int OCC, int KPS = 1>
__global__ void __launch_bounds__(192, OCC) sgemm2_kernel(const __grid_constant__ Sgemm2Params p) {
  pdl_launch_dependents();
  constexpr int B_BYTES = BN * kChunk2 * 2;
  constexpr int SLOT_BYTES = kABytes2 + B_BYTES;        // one k-block: A box + weight box
  constexpr int STAGE_BYTES = KPS * SLOT_BYTES;
  constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  constexpr int PANELS = BN / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + STAGES * STAGE_BYTES;                      // 2 panels
  float* bias_s = reinterpret_cast<float*>(staging + 2 * kPanelBytes);  // [2][BN]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bias_s + 2 * BN);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull = empty_bar + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int total_tiles = p.m_tiles * n_tiles;

  int nkb = 0;
  for (int s = 0; s < p.nsrc; ++s) nkb += p.src_chunks[s] * p.src_taps[s];

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.nsrc; ++s) tma_prefetch_desc(&p.amap[s]);
    tma_prefetch_desc(&p.bmap);
    tma_prefetch_desc(&p.omap);
    // The weights do not depend on the previous kernel (they are packed once per optimiser step), the activations do:
    // pull the first tile's weight blocks into L2 NOW, under the previous kernel's tail, so that after pdl_wait only the
    // activation loads pay a full memory round trip (the 153 MB operand arena does not stay in L2 across a step).
    if (static_cast<int>(blockIdx.x) < total_tiles) {
      const int n_tile0 = static_cast<int>(blockIdx.x) % n_tiles;
      const int npf = nkb < 16 ? nkb : 16;      // 8 / 16 / 48 measured: 8.935 / 8.905-8.925 / 8.912 ms per step (same box pairs)
      for (int kb = 0; kb < npf; ++kb)
        asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(
                         reinterpret_cast<uint64_t>(&p.bmap)),
                     "r"(kb * kChunk2), "r"(n_tile0 * BN)
                     : "memory");
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // set-up above touched no global memory; everything below may
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // The whole warp walks the loop with warp-uniform values (ring position, tile origin, descriptors live in uniform
    // registers) and ONE elected lane issues the asynchronous instructions.  With the loop inside `if (lane == 0)` ptxas
    // had to wrap every UTMALDG / UTCHMMA in an ELECT + R2UR.BROADCAST + BRA.U.ANY uniformisation loop (~35 cycles per
    // instruction on the single issuing thread, which is what paces a k-block at the narrow N tiles of this network).
    const bool leader = elect_one();
    {
      // The producer and the MMA issuer are single threads whose loop bodies are on the critical path of every k-block
      // (profiles/experiments/r01_gemm_pacing.txt): the ring position (stage, phase) is carried incrementally -- no
      // division / modulo by STAGES -- and nothing is done per k-block that can be done per stage or per tile.
      int st = 0;
      uint32_t ph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m_tile = t / n_tiles, n_tile = t - m_tile * n_tiles;
        int b0, y0;
        tile_origin2(p.g, m_tile, b0, y0);
        int kb = 0, slot = 0, in_stage = 0;
        for (int s = 0; s < p.nsrc; ++s) {
          const int taps = p.src_taps[s];
          for (int tp = 0; tp < taps; ++tp) {
            int dx = 0, dy = 0;
            if (taps == 9) {
              dy = (tp / 3 - 1) * p.tap_sign;
              dx = (tp % 3 - 1) * p.tap_sign;
            }
            const int chunks = p.src_chunks[s];
            for (int c = 0; c < chunks; ++c, ++kb) {
              if (slot == 0) {      // first k-block of a stage: wait for the slot, announce the bytes of the whole stage
                mbar_wait(&empty_bar[st], ph ^ 1);
                in_stage = nkb - kb < KPS ? nkb - kb : KPS;
                if (leader) mbar_arrive_expect_tx(&full_bar[st], in_stage * SLOT_BYTES);
              }
              uint8_t* sa = smem + st * STAGE_BYTES + slot * SLOT_BYTES;
              if (leader) {
                tma_load_4d(sa, &p.amap[s], &full_bar[st], c * kChunk2, dx, y0 + dy, b0);
                tma_load_2d(sa + kABytes2, &p.bmap, &full_bar[st], kb * kChunk2, n_tile * BN);
              }
              if (++slot == in_stage) {
                slot = 0;
                if (++st == STAGES) {
                  st = 0;
                  ph ^= 1;
                }
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = make_idesc(BN, false, false);
      uint32_t lt = 0, ph = 0;
      int st = 0;
      // shared-memory descriptors of slot 0 of stage 0; a descriptor's address field counts 16-byte units
      const uint64_t da0 = make_sdesc(smem_u32(smem), 16, 1024);
      const uint64_t db0 = make_sdesc(smem_u32(smem) + kABytes2, 16, 1024);
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
        const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);   // the epilogue has drained this accumulator stage
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * BN;
        for (int kb = 0; kb < nkb; kb += KPS) {
          mbar_wait(&full_bar[st], ph);
          tc_fence_after();
          const int nk = nkb - kb < KPS ? nkb - kb : KPS;
          const uint64_t soff = static_cast<uint64_t>(st * (STAGE_BYTES >> 4));
#pragma unroll
          for (int j = 0; j < KPS; ++j) {
            if (j < nk) {
              const uint64_t da = da0 + soff + j * (SLOT_BYTES >> 4);
              const uint64_t db = db0 + soff + j * (SLOT_BYTES >> 4);
              if (leader) {
#pragma unroll
                for (int k = 0; k < kChunk2 / 16; ++k) umma_f16(tacc, da + 2 * k, db + 2 * k, idesc, ((kb + j) | k) ? 1u : 0u);
              }
            }
          }
          if (leader) umma_commit(&empty_bar[st]);
          if (++st == STAGES) {
            st = 0;
            ph ^= 1;
          }
        }
        if (leader) umma_commit(&tfull[acc]);
      }
    }
  } else {
    const int q = warp & 3;                      // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;                 // row of the tile owned by this thread
    const bool storer = (warp == 2 && lane == 0);
    uint32_t lt = 0, pc = 0;                     // local tile counter, running panel counter
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
      const int m_tile = t / n_tiles, n_tile = t - m_tile * n_tiles;
      const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
      const long long row = static_cast<long long>(m_tile) * kTileM2 + r;
      const bool row_ok = row < p.M;
      float* bs = bias_s + acc * BN;
      if (p.bias) {
        for (int i = threadIdx.x - 64; i < BN; i += 128) {
          const int n = n_tile * BN + i;
          bs[i] = n < p.N ? __ldg(p.bias + n) : 0.f;
        }
      }
      // the residual rows do not depend on the MMA: fetch panel 0's before waiting for the accumulator
      uint4 rres[8];
      const bool has_res = p.res && row_ok;
      const h16* res_row = reinterpret_cast<const h16*>(p.res) + row * p.ldres + n_tile * BN;
      // (N = 32, the first stage's level-0 layers, runs as half of a 64-wide tile: the weight rows beyond N are the operand's
      // zero padding, the TMA store clips the columns beyond N, and the residual row is only read up to N)
      const int res_vecs = min(8, (p.N - n_tile * BN) >> 3);
      if (has_res) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          rres[j] = j < res_vecs ? __ldg(reinterpret_cast<const uint4*>(res_row + 8 * j)) : make_uint4(0u, 0u, 0u, 0u);
      }
      mbar_wait(&tfull[acc], aph);
      tc_fence_after();
#pragma unroll 1
      for (int pn = 0; pn < PANELS; ++pn) {
        const int n0 = n_tile * BN + pn * 64;
        if (n0 >= p.N) break;                    // uniform over the CTA
        float v[64];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + pn * 64;
        tmem_ld32(taddr, v);
        tmem_ld32(taddr + 32, v + 32);
        tmem_ld_wait();
        uint8_t* slot = staging + (pc & 1) * kPanelBytes;
        if (storer) bulk_wait_read<1>();         // the store that last used this slot has finished reading it
        epi_bar();                               // slot free; bias staged
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 64; ++i) v[i] += bs[pn * 64 + i];
        }
        if (p.act == ACT_SILU) {
#pragma unroll
          for (int i = 0; i < 64; ++i) v[i] = silu_f(v[i]);
        }
        if (has_res) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float f[8];
            unpack8(rres[j], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * j + i] += f[i];
          }
          if (pn + 1 < PANELS && n0 + 64 < p.N) {   // next panel's rows: in flight during this panel's store
#pragma unroll
            for (int j = 0; j < 8; ++j) rres[j] = __ldg(reinterpret_cast<const uint4*>(res_row + (pn + 1) * 64 + 8 * j));
          }
        }
        uint8_t* rowp = slot + r * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(rowp + ((j ^ (r & 7)) << 4)) = pack8(v + 8 * j);
        fence_async_smem();
        epi_bar();
        if (storer) {
          tma_store_2d(&p.omap, slot, n0, m_tile * kTileM2);
          bulk_commit();
        }
        ++pc;
      }
      tc_fence_before();
      mbar_arrive(&tempty[acc]);
    }
    if (storer) bulk_wait_read<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <int BN, int STAGES, int OCC, int KPS = 1>
static int launch_sgemm2_t(const Sgemm2Params& p, int num_sms, cudaStream_t stream) {
  constexpr int smem = STAGES * KPS * (kABytes2 + BN * kChunk2 * 2) + 2 * kPanelBytes + 2 * BN * 4 + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(sgemm2_kernel<BN, STAGES, OCC, KPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  const int tiles = p.m_tiles * ((p.N + BN - 1) / BN);
  const int grid = tiles < OCC * num_sms ? tiles : OCC * num_sms;
  static const bool log = getenv("ENCDIFF_GEMM_LOG") != nullptr;
  if (log) {
    int kb = 0;
    for (int s = 0; s < p.nsrc; ++s) kb += p.src_chunks[s] * p.src_taps[s];
    fprintf(stderr, "GEMMLOG sgemm2 M=%d N=%d K=%d taps=%d bn=%d st=%d grid=%d tiles=%d\n", p.M, p.N, kb * 64, p.src_taps[0], BN,
            STAGES, grid, tiles);
  }
  launch_k(sgemm2_kernel<BN, STAGES, OCC, KPS>, grid, 192, smem, stream, p);
  return static_cast<int>(cudaGetLastError());
}

// N tile by a per-SM cost model.  These GEMMs are bound by the L2 -> shared-memory fill of their stages, not by the
// tensor pipe: a k-block moves 16 KB of A plus BN/4 KB of B per CTA whatever BN is, so a wider tile moves fewer bytes
// per FLOP but leaves SMs idle when it produces fewer tiles than SMs.  cost = rounds * (k-blocks * stage bytes / fill
// rate + epilogue panels); constants from the in-graph timings in profiles/ (about 90 GB/s per SM, 0.4 us per panel).
int sgemm2_pick_bn(int m_tiles, int N, int num_sms, int nkb) {
  if (N == 32) return 64;            // half-used 64-wide tile (see the epilogue); these layers are bandwidth-bound
  if (N % 64 != 0) return 0;
  int best = 64;
  float best_cost = 1e30f;
  for (int bn = 64; bn <= 256; bn *= 2) {
    if (bn > 64 && N % bn != 0) continue;
    const int tiles = m_tiles * (N / bn);
    const int rounds = (tiles + num_sms - 1) / num_sms;
    const float fill_us = static_cast<float>(nkb) * (16.f + bn / 4.f) / 90.f;
    const float cost = rounds * (fill_us + 0.4f * (bn / 64));
    if (cost < best_cost * 0.97f) {      // prefer the narrower tile on near-ties (more CTAs in flight)
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

int launch_sgemm2(const Sgemm2Params& p, int bn, int num_sms, cudaStream_t stream) {
  int nkb = 0;
  for (int s = 0; s < p.nsrc; ++s) nkb += p.src_chunks[s] * p.src_taps[s];
  // short-K GEMMs (1x1 / linear with K <= 576 = 9 k-blocks): a tile is 1-9 k-blocks, so a deep smem ring buys nothing;
  // two small CTAs per SM overlap one CTA's epilogue / fill / drain with the other's loads instead.  Threshold and the
  // k-blocks per stage of the deep variants (2) were measured on the B = 256 step: k-blocks <= 4 / 9 / 18 / 36 / all gave
  // 10.33 / 10.27 / 10.28 / 10.40 / 10.46 ms, KPS 1 / 2 / 3 gave 10.27 / 10.11 / 10.25 ms
  // (profiles/experiments/r01_gemm_kps_sweep.txt).  Constants, not environment tunables.
  const bool shallow = nkb <= 9;
  switch (bn) {
    case 64:
      if (shallow) return launch_sgemm2_t<64, 3, 2>(p, num_sms, stream);
      return launch_sgemm2_t<64, 3, 1, 2>(p, num_sms, stream);
    case 128:
      if (shallow) return launch_sgemm2_t<128, 2, 2>(p, num_sms, stream);
      return launch_sgemm2_t<128, 2, 1, 2>(p, num_sms, stream);
    case 256: return launch_sgemm2_t<256, 3, 1>(p, num_sms, stream);
    default: return static_cast<int>(cudaErrorInvalidValue);
  }
}

}  // namespace edb
