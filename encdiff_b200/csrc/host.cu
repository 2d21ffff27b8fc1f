#include <stdlib.h>

#include "host.cuh"
#include "types.cuh"

#include <cudaTypedefs.h>

#include <mutex>

namespace edb {

#ifdef EDB_ACT_BF16
static const CUtensorMapDataType kMapType = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
#else
static const CUtensorMapDataType kMapType = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
#endif

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* last_error() { return g_last_error.c_str(); }

// cuTensorMapEncodeTiled is a driver entry point; fetch it through the runtime so the library has no link-time
// dependency on libcuda (this file is also compiled on boxes without a driver).
static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

bool make_geom(int H, int W, TileGeom* g) {
  const int hw = H * W;
  if (H <= 0 || W <= 0 || (hw & (hw - 1)) != 0 || W > 128 || (W & (W - 1)) != 0) return false;
  g->H = H;
  g->W = W;
  if (hw >= 128) {
    g->tiles_per_img = hw / 128;
    g->box_h = 128 / W;
    g->box_b = 1;
  } else {
    g->tiles_per_img = 1;
    g->box_h = H;
    g->box_b = 128 / hw;
  }
  return true;
}

int make_act_tmap(CUtensorMap* m, const void* base, int C, int H, int W, int B, const TileGeom& g) {
  auto enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
    return EDB_ERR_CUDA;
  }
  EDB_REQUIRE((C % 8) == 0, "activation channels must be a multiple of 8 for TMA");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "activation base must be 16-byte aligned");
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(B)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2,
                           static_cast<cuuint64_t>(H) * W * C * 2};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(W), static_cast<cuuint32_t>(g.box_h),
                       static_cast<cuuint32_t>(g.box_b)};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, kMapType, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled(activation) failed with CUresult " + std::to_string(static_cast<int>(r)) +
                   " C=" + std::to_string(C) + " H=" + std::to_string(H) + " W=" + std::to_string(W) +
                   " B=" + std::to_string(B));
    return EDB_ERR_CUDA;
  }
  return EDB_OK;
}

int make_w_tmap(CUtensorMap* m, const void* base, long long K, long long rows, int bn) {
  auto enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
    return EDB_ERR_CUDA;
  }
  EDB_REQUIRE((K % 64) == 0, "packed weight K must be a multiple of 64");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "packed weight base must be 16-byte aligned");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(K) * 2};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(bn)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, kMapType, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled(weights) failed with CUresult " + std::to_string(static_cast<int>(r)));
    return EDB_ERR_CUDA;
  }
  return EDB_OK;
}

int make_out_tmap(CUtensorMap* m, const void* base, long long M, int N, int ld) {
  auto enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
    return EDB_ERR_CUDA;
  }
  // N = 32: the 64-wide store box is clipped at the tensor edge (gemm2.cu, half-used tile)
  EDB_REQUIRE(((N % 64) == 0 || N == 32) && (ld % 8) == 0, "output map: N must be a multiple of 64 (or 32) and ld of 8");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "output base must be 16-byte aligned");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(N), static_cast<cuuint64_t>(M)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, kMapType, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled(output) failed with CUresult " + std::to_string(static_cast<int>(r)));
    return EDB_ERR_CUDA;
  }
  return EDB_OK;
}

int make_grad_tmap(CUtensorMap* m, float* base, int cols, int rows, bool three_d) {
  auto enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
    return EDB_ERR_CUDA;
  }
  EDB_REQUIRE((cols % 4) == 0, "gradient map: row pitch must be a multiple of 16 bytes");
  EDB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "gradient base must be 16-byte aligned");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows), 9};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols) * 4, static_cast<cuuint64_t>(cols) * rows * 4};
  cuuint32_t box[3] = {32, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, three_d ? 3 : 2, base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled(gradient) failed with CUresult " + std::to_string(static_cast<int>(r)));
    return EDB_ERR_CUDA;
  }
  return EDB_OK;
}

int device_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

bool pdl_enabled() {
  static const bool on = getenv("ENCDIFF_NO_PDL") == nullptr;
  return on;
}

bool debug_skip_kernel(const void* func) {
  const char* pat = getenv("ENCDIFF_DEBUG_SKIP_KERNELS");   // read per launch: launches happen at capture time only
  if (!pat || !*pat) return false;
  const char* name = nullptr;
  if (cudaFuncGetName(&name, func) != cudaSuccess || !name) return false;
  const char* p = pat;
  while (*p) {
    const char* e = strchr(p, ',');
    const size_t n = e ? static_cast<size_t>(e - p) : strlen(p);
    if (n > 0 && n < 200) {
      char buf[200];
      memcpy(buf, p, n);
      buf[n] = 0;
      if (strstr(name, buf)) return true;
    }
    if (!e) break;
    p = e + 1;
  }
  return false;
}

}  // namespace edb
