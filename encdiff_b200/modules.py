"""Parameter-holding module tree with the reference's names, shapes, construction order and initialisation, so that
``state_dict`` keys, ``named_parameters()`` order, EMA names and seeded initialisation are identical to the reference
``UNetModel`` (SURVEY.md §8b "Attributes read by callers", "Gradient / parameter ownership").

The torch layers below (nn.Conv2d, nn.Linear, ...) are never called inside the UNet: they only own the fp32 master
parameters.  All UNet compute goes through the C ABI (``encdiff_b200.unet.UNetModel``).  ``SpatialTransformer`` and
``CrossAttention`` additionally expose a stand-alone ``forward`` over the same CUDA kernels
(``encdiff_b200.attention``) so they stay individually testable against the reference (attention.py:152-261).
"""
from __future__ import annotations

import torch
from torch import nn


def zero_module(module: nn.Module) -> nn.Module:
    """Zero-initialise (reference: ResBlock out conv, SpatialTransformer.proj_out, UNetModel.out[-1])."""
    for p in module.parameters():
        p.detach().zero_()
    return module


class GroupNorm32(nn.GroupNorm):
    """32-group GroupNorm evaluated in fp32 (diffusionmodules/util.py:227-244); holder only."""


def normalization(channels: int) -> nn.GroupNorm:
    return GroupNorm32(32, channels)


class TimestepEmbedSequential(nn.Sequential):
    """Container matching openaimodel_enc.py:74-88; dispatch happens inside the CUDA plan, not here."""

    def forward(self, *args, **kwargs):  # pragma: no cover - never used
        raise RuntimeError("encdiff_b200 blocks are executed by UNetModel's fused CUDA plan, not module by module")


class ResBlock(nn.Module):
    """Holder for ResBlock parameters (openaimodel_enc.py:179-241)."""

    def __init__(self, channels, emb_channels, dropout, out_channels=None, use_scale_shift_norm=True, up=False,
                 down=False):
        super().__init__()
        if dropout != 0:
            raise NotImplementedError("encdiff_b200: dropout > 0 is not supported")
        if not use_scale_shift_norm:
            raise NotImplementedError("encdiff_b200: use_scale_shift_norm=False is not supported")
        self.channels = channels
        self.emb_channels = emb_channels
        self.out_channels = out_channels or channels
        self.updown = up or down
        self.up, self.down = up, down
        self.in_layers = nn.Sequential(normalization(channels), nn.SiLU(),
                                       nn.Conv2d(channels, self.out_channels, 3, padding=1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb_channels, 2 * self.out_channels))
        self.out_layers = nn.Sequential(normalization(self.out_channels), nn.SiLU(), nn.Dropout(p=dropout),
                                        zero_module(nn.Conv2d(self.out_channels, self.out_channels, 3, padding=1)))
        if self.out_channels == channels:
            self.skip_connection = nn.Identity()
        else:
            self.skip_connection = nn.Conv2d(channels, self.out_channels, 1)

    def forward(self, x, emb):  # pragma: no cover
        raise RuntimeError("encdiff_b200.ResBlock runs only inside UNetModel's fused CUDA plan")


class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)


class FeedForward(nn.Module):
    """attention.py:47-64 with glu=True (the only variant BasicTransformerBlock builds)."""

    def __init__(self, dim, mult=4, dropout=0.):
        super().__init__()
        inner = int(dim * mult)
        self.net = nn.Sequential(GEGLU(dim, inner), nn.Dropout(dropout), nn.Linear(inner, dim))


class CrossAttention(nn.Module):
    """Same constructor / parameters as attention.py:152-168; forward(x, context=None, mask=None) runs the CUDA
    kernels (see encdiff_b200/attention.py)."""

    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0.):
        super().__init__()
        inner = dim_head * heads
        self.query_dim = query_dim
        self.context_dim = query_dim if context_dim is None else context_dim
        self.is_self = context_dim is None
        self.scale = dim_head ** -0.5
        self.heads = heads
        self.dim_head = dim_head
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(self.context_dim, inner, bias=False)
        self.to_v = nn.Linear(self.context_dim, inner, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, query_dim), nn.Dropout(dropout))

    def forward(self, x, context=None, mask=None):
        from . import attention
        return attention.cross_attention_forward(self, x, context, mask)


class BasicTransformerBlock(nn.Module):
    """attention.py:196-215; registration order attn1, ff, attn2, norm1..3."""

    def __init__(self, dim, n_heads, d_head, dropout=0., context_dim=None):
        super().__init__()
        self.attn1 = CrossAttention(query_dim=dim, heads=n_heads, dim_head=d_head, dropout=dropout)
        self.ff = FeedForward(dim, dropout=dropout)
        self.attn2 = CrossAttention(query_dim=dim, context_dim=context_dim, heads=n_heads, dim_head=d_head,
                                    dropout=dropout)
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.norm3 = nn.LayerNorm(dim)


class SpatialTransformer(nn.Module):
    """Same constructor / parameters as attention.py:218-247; forward(x, context=None) runs the CUDA kernels."""

    def __init__(self, in_channels, n_heads, d_head, depth=1, dropout=0., context_dim=None):
        super().__init__()
        if depth != 1:
            raise NotImplementedError("encdiff_b200: transformer depth != 1 is not supported")
        if dropout != 0:
            raise NotImplementedError("encdiff_b200: dropout > 0 is not supported")
        self.in_channels = in_channels
        self.n_heads, self.d_head = n_heads, d_head
        self.context_dim = context_dim
        inner = n_heads * d_head
        if inner != in_channels:
            raise NotImplementedError("encdiff_b200: SpatialTransformer needs n_heads * d_head == in_channels")
        self.norm = nn.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Conv2d(in_channels, inner, kernel_size=1, stride=1, padding=0)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, n_heads, d_head, dropout=dropout, context_dim=context_dim)
             for _ in range(depth)])
        self.proj_out = zero_module(nn.Conv2d(inner, in_channels, kernel_size=1, stride=1, padding=0))

    def forward(self, x, context=None):
        from . import attention
        return attention.spatial_transformer_forward(self, x, context)
