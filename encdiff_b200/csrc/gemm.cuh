// Parameter blocks + launchers of the two tcgen05 GEMM kernels.
//
//  sgemm  : "segmented implicit GEMM".  D[M, N] = sum_s sum_tap sum_chunk  A_s(shifted by tap)[M, 64] * Wt[N, 64]^T
//           M = pixels (b, y, x) of an NHWC h16 tensor, N = output channels, K = concatenation of up to three
//           sources, each either 1 tap (1x1 conv / linear) or 9 taps (3x3 conv, pad 1, realised by shifting the
//           TMA box and letting TMA zero-fill out-of-bounds pixels).  Covers reference call sites
//           nn.Conv2d 3x3 / 1x1 and nn.Linear (openaimodel_enc.py:201-241, attention.py:37-64,152-193,231-247)
//           and, with transposed packed weights and negated shifts, their input gradients.
//  wgrad  : dW[p, (tap, q)] = sum_pixels dY[pix, p] * X(shifted by tap)[pix, q]   (both operands MN-major),
//           fp32 atomics into the reference-layout gradient tensor.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace edb {

enum { ACT_NONE = 0, ACT_SILU = 1 };

// geometry of the pixel tiling shared by both kernels: one M/K tile = 128 consecutive pixels in (b, y, x) order
struct TileGeom {
  int H, W;            // spatial size of the tensor the tile walks over
  int tiles_per_img;   // max(1, H*W/128)
  int box_h;           // rows of one image covered by a tile (HW >= 128), else H
  int box_b;           // images covered by a tile (HW < 128), else 1
};

struct SgemmParams {
  CUtensorMap amap[3];  // 4D (C, W, H, B) h16 maps of the A sources, box (64, box_w, box_h, box_b), SWIZZLE_128B
  CUtensorMap bmap;     // 2D (Ktot, Nrows) h16 map of the packed weights, box (64, BN)
  int nsrc;
  int src_chunks[3];    // 64-channel chunks per source
  int src_taps[3];      // 1 or 9
  int tap_sign;         // +1: conv forward; -1: conv input-gradient (mirrored taps)
  TileGeom g;
  int M, N;             // valid rows / columns
  const float* bias;    // [N] fp32 or null
  const void* res;      // h16 [M, ldres] added in the epilogue, or null
  int ldres;
  int act;              // ACT_*
  int out_fp32;         // 0: 16-bit out, 1: fp32 out
  void* out0;           // columns [0, split_col)
  int ld0;
  int split_col;        // >= N when there is a single destination; multiple of 32 otherwise
  void* out1;           // columns [split_col, N)
  int ld1;
};

// persistent kernel (gemm2.cu): 16-bit output with N % 64 == 0 and a dense [M, N] destination (leading dimension ldo)
struct Sgemm2Params {
  CUtensorMap amap[3];  // as SgemmParams
  CUtensorMap bmap;     // 2D (Ktot, Nrows) packed weights, box (64, BN)
  CUtensorMap omap;     // 2D (N, M) output, box (64, 128), SWIZZLE_128B (store side)
  int nsrc;
  int src_chunks[3];
  int src_taps[3];
  int tap_sign;
  TileGeom g;
  int M, N, m_tiles;
  const float* bias;    // [N] fp32 or null
  const void* res;      // 16-bit [M, ldres] or null
  int ldres;
  int act;
};

struct WgradParams {
  CUtensorMap pmap;     // dY   4D (Cp, W, H, B) box (64, ...)
  CUtensorMap qmap;     // X    4D (Cq, W, H, B) box (64, ...)
  TileGeom g;
  int taps;             // 1 or 9
  int m_tiles;          // number of 128-pixel tiles (reduction length / 128)
  int tiles_per_split;  // reduction tiles handled by one CTA (blockIdx.z)
  int Cp, Cq;           // valid channels on each side
  int p_off;            // first channel of the dY slice (dY may hold several stacked outputs, e.g. q|k|v)
  float* out;           // fp32 gradient, element (p, q, tap) at out[p*stride_p + (q+q_off)*stride_q + tap*stride_tap]
  long long stride_p, stride_q, stride_tap;
  int q_off;
  const float* gscale;  // device {scale, 1/scale} of dY (null = 1): the fp32 result is multiplied by gscale[1]
};

// persistent weight-gradient kernel (wgrad2.cu)
struct alignas(64) Wgrad2Params {
  CUtensorMap pmap;     // dY  4D (Cdy, W, H, B) box (64, ...)
  CUtensorMap qmap;     // X   4D (Cx, W, H, B) box (64, ...)
  CUtensorMap omap;     // fp32 destination: 2D (Cq_total, Cp) [linear / 1x1] or 3D (Cq, Cp, 9) tap-major staging
                        // [conv3x3]; box (32, 128[, 1]), SWIZZLE_128B
  TileGeom g;
  int taps, m_tiles, tiles_per_split, splits, p_tiles, q_tiles;   // the last four are filled by the launcher
  int Cp, Cq, p_off, q_off;
  int out3d;            // 1: omap is the 3D staging map
  float* bias0;         // += sum_pix dY[pix, p]   (null: the layer has no bias)
  float* bias1;         // second destination of the same sum (conv2 + fused skip conv share dY) or null
  const float* gscale;  // device {scale, 1/scale} of dY (null = 1)
};
struct WgradFinalizeEntry {
  float* stage;         // [9][Cp][Cq]
  float* grad;          // [Cp][Cq][9] reference layout
  long long prefix;     // first index of this entry in the block-padded (multiples of 256) global (p, q) space
  long long pq;         // Cp * Cq
};
int launch_wgrad2(const Wgrad2Params& p, int num_sms, cudaStream_t stream);
// grouped form: problems in device memory (64-byte aligned), prefix[n + 1] = first global item of each problem
int wgrad2_prepare_group(Wgrad2Params* p, int steps);
int launch_wgrad2_group(const Wgrad2Params* tab_dev, const int* prefix_dev, int total, int num_sms, cudaStream_t stream);
int launch_wgrad_finalize(const WgradFinalizeEntry* tab_dev, int n_entries, long long total_padded, cudaStream_t s);

// BN in {32(padded small N), 64, 128, 256}; returns cudaError_t as int
int launch_sgemm(const SgemmParams& p, int bn, cudaStream_t stream);
int launch_wgrad(const WgradParams& p, int bn, int splits, cudaStream_t stream);
int sgemm_pick_bn(int N);
int launch_sgemm2(const Sgemm2Params& p, int bn, int num_sms, cudaStream_t stream);
int sgemm2_pick_bn(int m_tiles, int N, int num_sms, int nkb);   // nkb = 64-wide k-blocks per tile   // 0: shape not supported by the persistent kernel

}  // namespace edb
