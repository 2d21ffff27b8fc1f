// C-ABI entry point that allocates NOTHING: the cross-attention backward operator with its fp32 dK | dV scratch in a
// file-scope device array instead of a caller buffer.  For callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid allocation).  No counterpart in the reference.
// Same arithmetic as encdiff_op_cross_attention_bwd (api_ops.cu), which it calls through the
// same launcher; only the scratch differs.
#include "../../include/encdiff_b200.h"

#include <mutex>
#include <string>

#include "common.cuh"
#include "host.cuh"
#include "kernels.cuh"

using namespace edb;

// Every GPU that loads the library holds this array until the process exits.  ENCDIFF_NOALLOC_SCRATCH_BYTES (16 MiB) is
// the stated maximum: B * T * 2 * heads * D * 4 bytes must fit (e.g. B = 256, T = 20, heads * D = 256 needs 10.5 MB).
static __device__ __align__(256) unsigned char g_noalloc_scratch[ENCDIFF_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, not a device address: the real address
// is resolved once per device (with that device current) and cached.
static constexpr int kMaxDevices = 64;
static float* g_scratch_ptr[kMaxDevices];
static std::mutex g_scratch_mu;

static int noalloc_scratch(float** out) {
  int dev = -1;
  EDB_CHECK_CUDA(cudaGetDevice(&dev));
  EDB_REQUIRE(dev >= 0 && dev < kMaxDevices, "noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    EDB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    EDB_REQUIRE(p != nullptr, "noalloc: cudaGetSymbolAddress returned a null device address");
    g_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return EDB_OK;
}

extern "C" int encdiff_op_cross_attention_bwd_noalloc(const void* q, const float* ctx, const float* wk, const float* wv,
                                                      const float* kv, const void* dout, void* dq, float* dwk,
                                                      float* dwv, float* dctx, int B, int N, int heads, int D, int T,
                                                      int Dc, void* stream) {
  EDB_REQUIRE(q && ctx && wk && wv && dout && dq && dwk && dwv && dctx, "cross_attention_bwd_noalloc: null argument");
  EDB_REQUIRE(B > 0 && N > 0 && heads > 0 && D > 0 && T > 0 && Dc > 0, "cross_attention_bwd_noalloc: bad shape");
  const size_t need = cross_attn_bwd_scratch_bytes(B, heads, D, T);
  if (need > static_cast<size_t>(ENCDIFF_NOALLOC_SCRATCH_BYTES)) {
    set_last_error("cross_attention_bwd_noalloc: needs " + std::to_string(need) + " bytes of scratch (B*T*2*heads*D*4), "
                   "the static scratch holds ENCDIFF_NOALLOC_SCRATCH_BYTES = " +
                   std::to_string(static_cast<size_t>(ENCDIFF_NOALLOC_SCRATCH_BYTES)) +
                   "; use encdiff_op_cross_attention_bwd with a caller-owned scratch instead");
    return EDB_ERR_INVALID;
  }
  float* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  // The array keeps the previous call's contents.  The launcher clears what must start at zero itself: the CUDA-core
  // path memsets the dK | dV region it accumulates into (an enqueued memset, no allocation); the tensor-core path
  // (kv != NULL) overwrites every element it later reads.
  EDB_CHECK_CUDA(static_cast<cudaError_t>(launch_cross_attn_bwd(static_cast<const h16*>(q), ctx, wk, wv,
                                                                static_cast<const h16*>(dout), static_cast<h16*>(dq), dwk,
                                                                dwv, dctx, nullptr, scratch, B, N, heads, D, T, Dc,
                                                                static_cast<cudaStream_t>(stream), false, kv)));
  return EDB_OK;
}
