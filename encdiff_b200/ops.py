"""Thin Python wrappers over the operator-level C-ABI entry points (used by the parity tests and by the
stand-alone CrossAttention / SpatialTransformer modules).  Tensors are NHWC 16-bit (`_lib.act_dtype()`) on the GPU."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

ACT_NONE, ACT_SILU = 0, 1


def _require_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("encdiff_b200 operators run on CUDA tensors only (no CPU fallback)")


def pack_conv3x3(w: torch.Tensor, rows_pad: int = 128) -> torch.Tensor:
    """[Co, Ci, 3, 3] fp32 -> h16 [Co_pad, 9*Ci_pad64] with K order (tap = ky*3+kx, ci)."""
    co, ci = w.shape[:2]
    cip = (ci + 63) // 64 * 64
    cop = (co + rows_pad - 1) // rows_pad * rows_pad
    out = torch.zeros(cop, 9, cip, dtype=_lib.act_dtype(), device=w.device)
    out[:co, :, :ci] = w.permute(0, 2, 3, 1).reshape(co, 9, ci).to(_lib.act_dtype())
    return out.reshape(cop, 9 * cip).contiguous()


def pack_conv3x3_t(w: torch.Tensor, rows_pad: int = 128) -> torch.Tensor:
    """Transposed pack for the input gradient: [Ci_pad, 9*Co_pad64], element [ci, tap*Co + co] = w[co, ci, tap]."""
    co, ci = w.shape[:2]
    cop = (co + 63) // 64 * 64
    cip = (ci + rows_pad - 1) // rows_pad * rows_pad
    out = torch.zeros(cip, 9, cop, dtype=_lib.act_dtype(), device=w.device)
    out[:ci, :, :co] = w.permute(1, 2, 3, 0).reshape(ci, 9, co).to(_lib.act_dtype())
    return out.reshape(cip, 9 * cop).contiguous()


def pack_linear(w: torch.Tensor, rows_pad: int = 128) -> torch.Tensor:
    """[N, K] fp32 -> h16 [N_pad, K_pad64]."""
    n, k = w.shape
    kp = (k + 63) // 64 * 64
    np_ = (n + rows_pad - 1) // rows_pad * rows_pad
    out = torch.zeros(np_, kp, dtype=_lib.act_dtype(), device=w.device)
    out[:n, :k] = w.to(_lib.act_dtype())
    return out


def sgemm(srcs, taps, w_packed, N, *, H, W, B, tap_sign=1, bias=None, res=None, act=ACT_NONE, out_fp32=False,
          out=None, out1=None, split_col=0, bn=0):
    """srcs: list of NHWC 16-bit (`_lib.act_dtype()`) tensors [B,H,W,C_s]; returns out [B*H*W, N] (bf16 or fp32)."""
    lib = _lib.load()
    _require_cuda(*srcs, w_packed, bias, res)
    n = len(srcs)
    M = B * H * W
    if out is None:
        out = torch.empty(M, N if out1 is None else split_col, device=srcs[0].device,
                          dtype=torch.float32 if out_fp32 else _lib.act_dtype())
    a_ptrs = (C.c_void_p * n)(*[s.data_ptr() for s in srcs])
    a_ch = (C.c_int * n)(*[s.shape[-1] for s in srcs])
    a_taps = (C.c_int * n)(*taps)
    st = lib.encdiff_op_sgemm(n, a_ptrs, a_ch, a_taps, H, W, B, w_packed.data_ptr(), w_packed.shape[0], N, tap_sign,
                              _lib.ptr(bias), _lib.ptr(res), 0 if res is None else res.stride(-2), act,
                              1 if out_fp32 else 0, out.data_ptr(), out.stride(-2), split_col, _lib.ptr(out1),
                              0 if out1 is None else out1.stride(-2), bn, _lib.current_stream())
    _lib.check(st, "op_sgemm")
    return out


def wgrad(dy, x, out, *, H, W, B, taps, stride_p, stride_q, stride_tap, q_off=0, bn=0, splits=1):
    lib = _lib.load()
    _require_cuda(dy, x, out)
    st = lib.encdiff_op_wgrad(dy.data_ptr(), dy.shape[-1], x.data_ptr(), x.shape[-1], H, W, B, taps, out.data_ptr(),
                              stride_p, stride_q, stride_tap, q_off, bn, splits, _lib.current_stream())
    _lib.check(st, "op_wgrad")
    return out
