// Persistent tcgen05 weight-gradient kernel (sm_100a).  See gemm.cuh (Wgrad2Params) for the math.
//
//   work item = (split, p_tile, tap, q_tile): D[128 x 64(+16)] = sum over the split's 128-pixel tiles of
//               dY^T[128 ch x 128 pix] * [ X(shifted by tap)[128 pix x 64 ch] | 1 ]      (both operands MN-major)
//   The extra all-ones block of the B operand makes column 64 of D the per-channel sum of dY: the bias gradient costs
//   no extra pass over dY (it replaces a separate column-sum kernel per layer).
//   warp 0 TMA producer, warp 1 MMA issuer (two TMEM accumulator stages), warps 2..5 epilogue:
//   TMEM -> registers -> * 1/gscale -> 128B-swizzled fp32 panel in shared memory -> cp.reduce.async.bulk.tensor (.add)
//   into the fp32 gradient: one bulk L2 reduction per [128 x 32] panel instead of 4096 scalar atomics.
//   Conv3x3 gradients are reduced into a tap-major staging tensor [9][Cp][Cq] (the reference layout [Cp][Cq][3][3] has
//   no 16-byte contiguous run per tap, which TMA needs); wgrad_finalize_kernel transposes it into the reference layout
//   once per backward.
#include <stdlib.h>

#include "common.cuh"
#include "gemm.cuh"

namespace edb {

static constexpr int kPix = 128;                       // pixels per reduction step (UMMA K = 8 x 16)
static constexpr int kBox = kPix * 64 * 2;             // 16 KB: one [128 pix x 64 ch] 16-bit box
static constexpr int kWBN = 64;                        // X channels per work item
static constexpr int kWN = kWBN + 16;                  // MMA N: 64 weight columns + the ones block (16 columns)
static constexpr int kWStages = 3;
static constexpr int kWStageBytes = 3 * kBox;          // dY: 2 boxes (M = 128), X: 1 box
static constexpr int kFPanel = 128 * 128;              // 16 KB: [128 rows x 32 fp32] staging panel

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void w_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void w_bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void w_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void w_epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ void w_tile_origin(const TileGeom& g, int tile, int& b0, int& y0) {
  if (g.tiles_per_img > 1) {
    b0 = tile / g.tiles_per_img;
    y0 = (tile - b0 * g.tiles_per_img) * g.box_h;
  } else {
    b0 = tile * g.box_b;
    y0 = 0;
  }
}

struct WItem {
  int split, p_tile, tap, q_tile;
};
__device__ __forceinline__ WItem w_decode(const Wgrad2Params& p, int item) {
  WItem w;
  w.q_tile = item % p.q_tiles;
  item /= p.q_tiles;
  w.tap = item % p.taps;
  item /= p.taps;
  w.p_tile = item % p.p_tiles;
  w.split = item / p.p_tiles;
  return w;
}

// Problem sources.  A CTA walks the global item sequence gi = blockIdx.x, blockIdx.x + gridDim.x, ...; at() maps gi to
// (problem, item inside the problem); `cur` is the caller's cursor into the prefix table (gi only grows).
struct WSingle {
  const Wgrad2Params* p;
  int total;
  __device__ __forceinline__ const Wgrad2Params& at(int gi, int& cur, int& item) const {
    item = gi;
    return *p;
  }
};
struct WGroup {
  const Wgrad2Params* tab;   // global memory, 64-byte aligned (tensor maps are fetched from here by the TMA unit)
  const int* prefix;         // [n + 1] first global item of each problem
  int total;
  __device__ __forceinline__ const Wgrad2Params& at(int gi, int& cur, int& item) const {
    while (gi >= __ldg(prefix + cur + 1)) ++cur;
    item = gi - __ldg(prefix + cur);
    return tab[cur];
  }
};

template <class Src>
__device__ __forceinline__ void wgrad2_body(const Src& src) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ones = smem + kWStages * kWStageBytes;      // [128 pix x 64 ch] of 1.0: the shared extra N block
  uint8_t* staging = ones + kBox;                      // 2 fp32 panels
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + 2 * kFPanel);
  uint64_t* empty_bar = full_bar + kWStages;
  uint64_t* tfull = empty_bar + kWStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total = src.total;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kWStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_slot);
  {
    const h162 one2 = f22h2(1.f, 1.f);
    uint4 o4;
    o4.x = o4.y = o4.z = o4.w = *reinterpret_cast<const uint32_t*>(&one2);
    for (int i = threadIdx.x; i < kBox / 16; i += blockDim.x) reinterpret_cast<uint4*>(ones)[i] = o4;
    w_fence_async_smem();   // generic-proxy writes -> visible to the tensor core (async proxy)
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // set-up above touched no global memory; everything below may
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    const bool leader = elect_one();   // whole warp walks the loop (uniform registers), one elected lane issues: see gemm2.cu
    {
      // single-thread loop on the critical path of every reduction step (profiles/experiments/r01_gemm_pacing.txt): ring
      // position and tile origin are carried incrementally (no division per step)
      int st = 0;
      uint32_t ph = 0;
      int cur = 0;
      for (int gi = blockIdx.x; gi < total; gi += gridDim.x) {
        int item;
        const Wgrad2Params& p = src.at(gi, cur, item);
        const WItem w = w_decode(p, item);
        const TileGeom g = p.g;
        const int p_col = p.p_off + w.p_tile * 128;
        int dx = 0, dy = 0;
        if (p.taps == 9) {
          dy = w.tap / 3 - 1;
          dx = w.tap % 3 - 1;
        }
        const int t_begin = w.split * p.tiles_per_split;
        const int t_end = min(p.m_tiles, t_begin + p.tiles_per_split);
        int b0, y0;
        w_tile_origin(g, t_begin, b0, y0);
        const int q_col = w.q_tile * kWBN;
        for (int t = t_begin; t < t_end; ++t) {
          mbar_wait(&empty_bar[st], ph ^ 1);
          uint8_t* sp = smem + st * kWStageBytes;
          if (leader) {
            mbar_arrive_expect_tx(&full_bar[st], kWStageBytes);
            tma_load_4d(sp, &p.pmap, &full_bar[st], p_col, 0, y0, b0);
            tma_load_4d(sp + kBox, &p.pmap, &full_bar[st], p_col + 64, 0, y0, b0);
            tma_load_4d(sp + 2 * kBox, &p.qmap, &full_bar[st], q_col, dx, y0 + dy, b0);
          }
          if (g.tiles_per_img > 1) {          // next 128-pixel tile: further down the same image, or the next image(s)
            y0 += g.box_h;
            if (y0 >= g.H) {
              y0 = 0;
              ++b0;
            }
          } else {
            b0 += g.box_b;
          }
          if (++st == kWStages) {
            st = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = make_idesc(kWN, true, true);
      constexpr uint32_t idesc_nobias = make_idesc(kWBN, true, true);   // items that do not feed a bias gradient skip the ones block
      const uint32_t ones_addr = smem_u32(ones);
      uint32_t lt = 0, ph = 0;
      int st = 0;
      int cur = 0;
      // descriptors of stage 0, reduction step 0.  A (dY): second 64-channel block kBox away.  B (X | ones): the second N
      // block is the shared ones box, i.e. its leading-dimension offset is (ones - X box) and shrinks by one stage size
      // per stage; both descriptor fields count 16-byte units, so a stage is a constant added / subtracted.
      const uint32_t sp0 = smem_u32(smem), sq0 = sp0 + 2 * kBox;
      const uint64_t da0 = make_sdesc(sp0, kBox, 1024);
      const uint64_t db0 = make_sdesc(sq0, ones_addr - sq0, 1024);
      constexpr uint64_t kStageUnits = kWStageBytes >> 4;
      for (int gi = blockIdx.x; gi < total; gi += gridDim.x, ++lt) {
        int item;
        const Wgrad2Params& p = src.at(gi, cur, item);
        const WItem w = w_decode(p, item);
        const int t_begin = w.split * p.tiles_per_split;
        const int nsteps = min(p.m_tiles, t_begin + p.tiles_per_split) - t_begin;
        const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * 128;
        const uint32_t id = (p.bias0 && w.q_tile == 0 && w.tap == 0) ? idesc : idesc_nobias;
        for (int i = 0; i < nsteps; ++i) {
          mbar_wait(&full_bar[st], ph);
          tc_fence_after();
          const uint64_t so = static_cast<uint64_t>(st) * kStageUnits;
          const uint64_t da = da0 + so;
          const uint64_t db = db0 + so - (so << 16);
          if (leader) {
#pragma unroll
            for (int k = 0; k < kPix / 16; ++k)   // 16 pixels (the reduction) = 16 rows of 128 B = 2048 B per MMA step
              umma_f16(tacc, da + k * (2048 >> 4), db + k * (2048 >> 4), id, (i | k) ? 1u : 0u);
            umma_commit(&empty_bar[st]);
          }
          if (++st == kWStages) {
            st = 0;
            ph ^= 1;
          }
        }
        if (leader) umma_commit(&tfull[acc]);
      }
    }
  } else {
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const bool storer = (warp == 2 && lane == 0);
    uint32_t lt = 0, pc = 0;
    int cur = 0;
    for (int gi = blockIdx.x; gi < total; gi += gridDim.x, ++lt) {
      int item;
      const Wgrad2Params& p = src.at(gi, cur, item);
      const float ginv = p.gscale ? __ldg(p.gscale + 1) : 1.f;
      const WItem w = w_decode(p, item);
      const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
      const int pch = w.p_tile * 128 + r;     // dY channel of this thread's row
      mbar_wait(&tfull[acc], aph);
      tc_fence_after();
      const uint32_t trow = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 128;
#pragma unroll 1
      for (int pn = 0; pn < kWBN / 32; ++pn) {
        const int qc0 = w.q_tile * kWBN + pn * 32;
        if (qc0 >= p.Cq) break;
        float v[32];
        tmem_ld32(trow + pn * 32, v);
        tmem_ld_wait();
        uint8_t* slot = staging + (pc & 1) * kFPanel;
        if (storer) w_bulk_wait_read<1>();
        w_epi_bar();
        uint8_t* rowp = slot + r * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(rowp + ((j ^ (r & 7)) << 4)) =
              make_float4(v[4 * j] * ginv, v[4 * j + 1] * ginv, v[4 * j + 2] * ginv, v[4 * j + 3] * ginv);
        w_fence_async_smem();
        w_epi_bar();
        if (storer) {
          if (p.out3d) tma_reduce_add_3d(&p.omap, slot, qc0, w.p_tile * 128, w.tap);
          else tma_reduce_add_2d(&p.omap, slot, p.q_off + qc0, w.p_tile * 128);
          w_bulk_commit();
        }
        ++pc;
      }
      if (p.bias0 && w.q_tile == 0 && w.tap == 0) {   // column 64 = sum over pixels of dY (ones block)
        float bsum[16];
        uint32_t* u = reinterpret_cast<uint32_t*>(bsum);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
              "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
            : "r"(trow + kWBN)
            : "memory");
        tmem_ld_wait();
        if (pch < p.Cp) {
          atomicAdd(p.bias0 + pch, bsum[0] * ginv);
          if (p.bias1) atomicAdd(p.bias1 + pch, bsum[0] * ginv);
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[acc]);
    }
    if (storer) w_bulk_wait_read<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<256>(tmem_base);
}

__global__ void __launch_bounds__(192, 1) wgrad2_kernel(const __grid_constant__ Wgrad2Params p) {
  wgrad2_body(WSingle{&p, p.q_tiles * p.taps * p.p_tiles * p.splits});
}
// every weight gradient of a backward pass as ONE launch: `tab` holds the problems in the order the backward met them
__global__ void __launch_bounds__(192, 1) wgrad2_group_kernel(const Wgrad2Params* __restrict__ tab,
                                                              const int* __restrict__ prefix, int total) {
  wgrad2_body(WGroup{tab, prefix, total});
}

int launch_wgrad2(const Wgrad2Params& p0, int num_sms, cudaStream_t stream) {
  constexpr int smem = kWStages * kWStageBytes + kBox + 2 * kFPanel + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(wgrad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  Wgrad2Params p = p0;
  p.p_tiles = (p.Cp + 127) / 128;
  p.q_tiles = (p.Cq + kWBN - 1) / kWBN;
  const int base = p.p_tiles * p.q_tiles * p.taps;
  // enough items to fill the chip about twice, at least 4 pixel tiles of reduction per item
  int splits = (2 * num_sms + base - 1) / base;
  const int max_splits = (p.m_tiles + 3) / 4;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  p.tiles_per_split = (p.m_tiles + splits - 1) / splits;
  p.splits = (p.m_tiles + p.tiles_per_split - 1) / p.tiles_per_split;
  const int total = base * p.splits;
  const int grid = total < num_sms ? total : num_sms;
  static const bool log = getenv("ENCDIFF_GEMM_LOG") != nullptr;
  if (log)
    fprintf(stderr, "GEMMLOG wgrad2 Cp=%d Cq=%d taps=%d m_tiles=%d splits=%d items=%d grid=%d\n", p.Cp, p.Cq, p.taps, p.m_tiles,
            p.splits, total, grid);
  launch_k(wgrad2_kernel, grid, 192, smem, stream, p);
  return static_cast<int>(cudaGetLastError());
}

static_assert(sizeof(Wgrad2Params) % 64 == 0, "problem tables keep their tensor maps 64-byte aligned");

// Split policy of a grouped problem: items of about `steps` pixel tiles each (the launch as a whole fills the chip, so
// a problem is split only to bound the cost of one item); returns the item count
int wgrad2_prepare_group(Wgrad2Params* p, int steps) {
  p->p_tiles = (p->Cp + 127) / 128;
  p->q_tiles = (p->Cq + kWBN - 1) / kWBN;
  int splits = (p->m_tiles + steps - 1) / steps;
  if (splits < 1) splits = 1;
  p->tiles_per_split = (p->m_tiles + splits - 1) / splits;
  p->splits = (p->m_tiles + p->tiles_per_split - 1) / p->tiles_per_split;
  return p->p_tiles * p->q_tiles * p->taps * p->splits;
}

int launch_wgrad2_group(const Wgrad2Params* tab_dev, const int* prefix_dev, int total, int num_sms, cudaStream_t stream) {
  constexpr int smem = kWStages * kWStageBytes + kBox + 2 * kFPanel + 1024 + 256;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(wgrad2_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  if (total <= 0) return 0;
  launch_k(wgrad2_group_kernel, total < num_sms ? total : num_sms, 192, smem, stream, tab_dev, prefix_dev, total);
  return static_cast<int>(cudaGetLastError());
}

// grad[p][q][tap] += stage[tap][p][q]; stage = 0       (one launch per backward for every conv3x3 weight)
// The (p, q) index space of every entry is padded to a multiple of 256, so a block works on 256 consecutive (p, q) of
// ONE entry: 9 coalesced reads per thread, transpose through shared memory, 2304 contiguous floats written.
__global__ void __launch_bounds__(256) wgrad_finalize_kernel(const WgradFinalizeEntry* __restrict__ tab, int n_entries) {
  pdl_enter();
  __shared__ float tile[256 * 9];
  __shared__ int s_entry;
  const long long base = static_cast<long long>(blockIdx.x) * 256;
  if (threadIdx.x == 0) {
    int lo = 0, hi = n_entries - 1;   // last entry with (padded) prefix <= base
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tab[mid].prefix <= base) lo = mid; else hi = mid - 1;
    }
    s_entry = lo;
  }
  __syncthreads();
  const WgradFinalizeEntry en = tab[s_entry];
  const long long e0 = base - en.prefix;            // first (p, q) of this block
  const long long e = e0 + threadIdx.x;
  if (e < en.pq) {
    float* st = en.stage + e;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      tile[threadIdx.x * 9 + t] = st[t * en.pq];
      st[t * en.pq] = 0.f;
    }
  }
  __syncthreads();
  const long long n = min(static_cast<long long>(256), en.pq - e0) * 9;   // floats of this block
  float* g = en.grad + e0 * 9;
  for (int i = threadIdx.x; i < n; i += 256) g[i] += tile[i];
}
int launch_wgrad_finalize(const WgradFinalizeEntry* tab_dev, int n_entries, long long total_padded, cudaStream_t s) {
  if (n_entries <= 0) return 0;
  launch_k(wgrad_finalize_kernel, static_cast<unsigned>(total_padded / 256), 256, 0, s, tab_dev, n_entries);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace edb
