// Host-side helpers shared by the C-ABI translation units: TMA tensor-map encoding, tile geometry, error slot.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

#include "gemm.cuh"

namespace edb {

// status codes returned through the C ABI (0 = ok)
enum {
  EDB_OK = 0,
  EDB_ERR_INVALID = 1,    // bad argument / unsupported configuration
  EDB_ERR_CUDA = 2,       // a CUDA runtime / driver call failed
  EDB_ERR_NOT_BOUND = 3,  // handle used before bind / plan
};

void set_last_error(const std::string& msg);
const char* last_error();

#define EDB_CHECK_CUDA(expr)                                                                       \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      ::edb::set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                   \
      return ::edb::EDB_ERR_CUDA;                                                                  \
    }                                                                                              \
  } while (0)

#define EDB_REQUIRE(cond, msg)                     \
  do {                                             \
    if (!(cond)) {                                 \
      ::edb::set_last_error(std::string(msg));     \
      return ::edb::EDB_ERR_INVALID;               \
    }                                              \
  } while (0)

// Pixel tiling of an (H, W) grid into 128-pixel tiles.  Returns false when the grid cannot be tiled
// (H*W must be a power of two; W <= 128).
bool make_geom(int H, int W, TileGeom* g);

// 4D (C, W, H, B) h16 NHWC activation map, box (64, W, box_h, box_b), SWIZZLE_128B, zero OOB fill.
int make_act_tmap(CUtensorMap* m, const void* base, int C, int H, int W, int B, const TileGeom& g);
// 2D (K, rows) h16 packed-weight map, box (64, bn), SWIZZLE_128B.
int make_w_tmap(CUtensorMap* m, const void* base, long long K, long long rows, int bn);
// 2D (N, M) 16-bit output map with row pitch ld elements, box (64, 128), SWIZZLE_128B (TMA store side of gemm2.cu).
int make_out_tmap(CUtensorMap* m, const void* base, long long M, int N, int ld);
int device_sm_count();
// fp32 gradient destination of wgrad2.cu: 2D (cols, rows) with row pitch `cols` floats, or 3D (cols, rows, 9);
// box (32, 128[, 1]), SWIZZLE_128B.  cols must be a multiple of 4 (16-byte row pitch).
int make_grad_tmap(CUtensorMap* m, float* base, int cols, int rows, bool three_d);

}  // namespace edb
