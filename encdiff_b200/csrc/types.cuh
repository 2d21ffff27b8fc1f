// The 16-bit storage / tensor-core operand type of the whole plan (activations, packed weights, activation gradients).
// Default fp16: same tcgen05 kind::f16 rate and bytes as bf16 with 8x finer rounding (11 vs 8 significand bits), which
// is what brings the eps prediction inside the 1e-2 parity bar (DESIGN.md "Numerics"); gradients are range-protected
// by a per-step power-of-two scale.  -DEDB_ACT_BF16 builds the bf16 variant.  tcgen05 rejects mixed fp16 x bf16
// operands in one MMA (measured: illegal instruction), so there is exactly one type.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace edb {

enum { FMT_F16 = 0, FMT_BF16 = 1 };   // instruction-descriptor a_format / b_format values of kind::f16

#ifdef EDB_ACT_BF16
typedef __nv_bfloat16 h16;
typedef __nv_bfloat162 h162;
#define EDB_FMT ::edb::FMT_BF16
__device__ __forceinline__ float h2f(h16 v) { return __bfloat162float(v); }
__device__ __forceinline__ h16 f2h(float v) { return __float2bfloat16(v); }
__device__ __forceinline__ float2 h22f2(h162 v) { return __bfloat1622float2(v); }
__device__ __forceinline__ h162 f22h2(float a, float b) { return __floats2bfloat162_rn(a, b); }
#else
typedef __half h16;
typedef __half2 h162;
#define EDB_FMT ::edb::FMT_F16
__device__ __forceinline__ float h2f(h16 v) { return __half2float(v); }
__device__ __forceinline__ h16 f2h(float v) { return __float2half_rn(v); }
__device__ __forceinline__ float2 h22f2(h162 v) { return __half22float2(v); }
__device__ __forceinline__ h162 f22h2(float a, float b) { return __floats2half2_rn(a, b); }
#endif

}  // namespace edb
