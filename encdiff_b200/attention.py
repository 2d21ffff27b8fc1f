"""Stand-alone ``CrossAttention.forward`` / ``SpatialTransformer.forward`` over the same CUDA kernels the UNet plan
launches (reference: ldm/modules/attention.py:170-193 and :250-261; SURVEY.md §8b "Sub-module surface").

They exist so that each sub-module stays individually testable against the reference; the fused, graph-captured path
(and the only differentiable one) is ``encdiff_b200.UNetModel``.  These forwards are inference-only: they raise if
autograd would need to record them.  No CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import torch

from . import _lib, ops


def _no_autograd(module, *tensors):
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors):
        raise RuntimeError(f"encdiff_b200.{type(module).__name__}.forward is inference-only; gradients flow through "
                           "encdiff_b200.UNetModel's fused backward (call under torch.no_grad())")


def _token_geom(n: int):
    """A (H, W) factorisation of the token count the GEMM's activation tensor map accepts (H*W a power of two)."""
    if n & (n - 1) or n < 1:
        raise NotImplementedError(f"encdiff_b200: token count {n} is not a power of two")
    w = min(n, 16)
    return n // w, w


def _f32(t):
    return t.detach().float().contiguous()


def _h16(t):
    return t.to(_lib.act_dtype()).contiguous()


def _linear(x2d, B, n_tok, weight, bias=None, res=None):
    """x2d h16 [B*n_tok, K] @ weight[N, K]^T (+ bias fp32) (+ res h16) -> h16 [B*n_tok, N] on the tcgen05 GEMM."""
    H, W = _token_geom(n_tok)
    N, K = weight.shape
    if K % 64 or N % 8:
        raise NotImplementedError(f"encdiff_b200: linear {K}->{N} needs K % 64 == 0 and N % 8 == 0")
    wp = ops.pack_linear(weight.detach().float(), rows_pad=256)
    return ops.sgemm([x2d.view(B, H, W, K)], [1], wp, N, H=H, W=W, B=B,
                     bias=None if bias is None else bias.detach().float().contiguous(), res=res)


def _layer_norm(x2d, ln):
    lib = _lib.load()
    M, C = x2d.shape
    y = torch.empty_like(x2d)
    stats = torch.empty(M, 2, dtype=torch.float32, device=x2d.device)
    w, b = _f32(ln.weight), _f32(ln.bias)          # named: raw pointers do not keep temporaries alive
    _lib.check(lib.encdiff_op_layer_norm(x2d.data_ptr(), w.data_ptr(), b.data_ptr(), float(ln.eps), y.data_ptr(),
                                         stats.data_ptr(), M, C, _lib.current_stream()), "op_layer_norm")
    return y


def _attention_tokens(attn, x2d, B, n_tok, context, res=None):
    """CrossAttention on h16 tokens [B*n_tok, C]; context None (self) or fp32 [B, T, Dc]; returns h16 [B*n_tok, C]
    (+ res when given: the BasicTransformerBlock residual is fused into the to_out GEMM epilogue)."""
    lib = _lib.load()
    heads, D = attn.heads, attn.dim_head
    inner = heads * D
    dev = x2d.device
    if context is None:
        wqkv = torch.cat([attn.to_q.weight, attn.to_k.weight, attn.to_v.weight], 0)
        qkv = _linear(x2d, B, n_tok, wqkv)
        o = torch.empty(B * n_tok, inner, dtype=x2d.dtype, device=dev)
        lse = torch.empty(B * heads * n_tok, dtype=torch.float32, device=dev)
        _lib.check(lib.encdiff_op_self_attention(qkv.data_ptr(), o.data_ptr(), lse.data_ptr(), B, n_tok, heads, D,
                                                 _lib.current_stream()), "op_self_attention")
    else:
        ctx = context.detach().float().contiguous()
        if ctx.dim() != 3 or ctx.shape[0] != B or ctx.shape[2] != attn.to_k.weight.shape[1]:
            raise RuntimeError(f"encdiff_b200.CrossAttention: context must be [B={B}, T, {attn.to_k.weight.shape[1]}], "
                               f"got {tuple(ctx.shape)}")
        q = _linear(x2d, B, n_tok, attn.to_q.weight)
        o = torch.empty(B * n_tok, inner, dtype=x2d.dtype, device=dev)
        wk, wv = _f32(attn.to_k.weight), _f32(attn.to_v.weight)
        kv = torch.empty(B, ctx.shape[1], 2 * inner, dtype=torch.float32, device=dev)
        tab = torch.empty(64, dtype=torch.uint8, device=dev)
        _lib.check(lib.encdiff_op_cross_kv_project(ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(), kv.data_ptr(),
                                                   tab.data_ptr(), B, inner, ctx.shape[1], ctx.shape[2],
                                                   _lib.current_stream()), "op_cross_kv_project")
        _lib.check(lib.encdiff_op_cross_attention(q.data_ptr(), ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(),
                                                  kv.data_ptr(), o.data_ptr(), B, n_tok, heads, D, ctx.shape[1], ctx.shape[2],
                                                  _lib.current_stream()), "op_cross_attention")
    return _linear(o, B, n_tok, attn.to_out[0].weight, attn.to_out[0].bias, res=res)


def cross_attention_forward(attn, x, context=None, mask=None):
    """attention.py:170-193.  x [B, N, query_dim]; context None or [B, T, context_dim]; mask unsupported (the
    reference's UNet never passes one)."""
    if mask is not None:
        raise NotImplementedError("encdiff_b200.CrossAttention: mask is not supported")
    ops._require_cuda(x, context)
    _no_autograd(attn, x, context, *attn.parameters())
    B, n_tok, C = x.shape
    y = _attention_tokens(attn, _h16(x).view(B * n_tok, C), B, n_tok, context)
    return y.view(B, n_tok, -1).to(x.dtype)


def spatial_transformer_forward(st, x, context=None):
    """attention.py:250-261 (+ BasicTransformerBlock._forward :212-215, FeedForward / GEGLU :37-64).
    x [B, C, H, W]; context [B, T, context_dim] or None (then attn2 is self-attention, as in the reference)."""
    lib = _lib.load()
    ops._require_cuda(x, context)
    _no_autograd(st, x, context, *st.parameters())
    B, C, H, W = x.shape
    n_tok = H * W
    M = B * n_tok
    dev = x.device
    x_tok = _h16(x.permute(0, 2, 3, 1)).view(M, C)                      # 'b c h w -> b (h w) c'
    h = torch.empty_like(x_tok)
    stats = torch.empty(B, 32, 2, dtype=torch.float32, device=dev)
    gw, gb = _f32(st.norm.weight), _f32(st.norm.bias)
    _lib.check(lib.encdiff_op_group_norm(x_tok.data_ptr(), gw.data_ptr(), gb.data_ptr(), float(st.norm.eps), 0,
                                         h.data_ptr(), stats.data_ptr(), B, H, W, C, _lib.current_stream()),
               "op_group_norm")
    h = _linear(h, B, n_tok, st.proj_in.weight.view(st.proj_in.weight.shape[0], C), st.proj_in.bias)
    for blk in st.transformer_blocks:
        h = _attention_tokens(blk.attn1, _layer_norm(h, blk.norm1), B, n_tok, None, res=h)
        h = _attention_tokens(blk.attn2, _layer_norm(h, blk.norm2), B, n_tok, context, res=h)
        geglu = blk.ff.net[0]
        g = _linear(_layer_norm(h, blk.norm3), B, n_tok, geglu.proj.weight, geglu.proj.bias)
        inner = g.shape[1] // 2
        a = torch.empty(M, inner, dtype=g.dtype, device=dev)
        _lib.check(lib.encdiff_op_geglu(g.data_ptr(), a.data_ptr(), M, inner, _lib.current_stream()), "op_geglu")
        h = _linear(a, B, n_tok, blk.ff.net[2].weight, blk.ff.net[2].bias, res=h)
    y = _linear(h, B, n_tok, st.proj_out.weight.view(C, -1), st.proj_out.bias, res=x_tok)
    return y.view(B, H, W, C).permute(0, 3, 1, 2).to(x.dtype).contiguous()
