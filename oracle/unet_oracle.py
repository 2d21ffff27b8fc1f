"""CPU oracle for the EncDiff denoising hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A functional fp32 PyTorch restatement of the reference algorithm (no nn.Module tree, parameters come in as a flat
``{name: tensor}`` dict keyed like the reference ``state_dict``).  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs may import this package; the product (``encdiff_b200``) never
does and has no CPU fallback.

Parity pin: the reference ships no golden vectors for this path, so the oracle is pinned against
outputs of the reference itself, generated in the build container by ``oracle/make_golden.py`` (which imports
``/root/reference``) and committed under ``tests/golden/``; ``tests/test_oracle.py`` re-checks them on every run and,
when ``/root/reference`` is present, compares against the live reference modules as well.

Every function cites the reference lines it restates (paths relative to the reference tree).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

Params = Dict[str, torch.Tensor]


# ------------------------------------------------------------------------------------------------------------------
# architecture description (restates UNetModel.__init__, ldm/modules/diffusionmodules/openaimodel_enc.py:443-694)
# ------------------------------------------------------------------------------------------------------------------
@dataclass
class UNetConfig:
    image_size: int = 16
    in_channels: int = 3
    model_channels: int = 64
    out_channels: int = 3
    num_res_blocks: int = 2
    attention_resolutions: Sequence[int] = (1, 2, 4)
    latent_unit: int = 20
    channel_mult: Sequence[int] = (1, 2, 4, 4)
    num_heads: int = 8
    context_dim: int = 16
    transformer_depth: int = 1
    use_scale_shift_norm: bool = True
    resblock_updown: bool = True


SHAPES3D = UNetConfig()  # configs/latent-diffusion/shapes3d-vq-4-16-encdiff.yaml:33-48


@dataclass
class Layer:
    kind: str          # 'conv_in' | 'res' | 'st'
    prefix: str        # state_dict prefix, e.g. 'input_blocks.1.0'
    cin: int = 0
    cout: int = 0
    up: bool = False
    down: bool = False
    heads: int = 0
    d_head: int = 0


@dataclass
class Arch:
    cfg: UNetConfig
    input_blocks: List[List[Layer]] = field(default_factory=list)
    middle: List[Layer] = field(default_factory=list)
    output_blocks: List[List[Layer]] = field(default_factory=list)
    head_channels: int = 0


def build_arch(cfg: UNetConfig) -> Arch:
    """Block list in construction order (openaimodel_enc.py:516-682; the spatial-transformer, resblock_updown,
    legacy=True branch that every shipped config takes)."""
    a = Arch(cfg)
    mc = cfg.model_channels
    a.input_blocks.append([Layer('conv_in', 'input_blocks.0.0', cfg.in_channels, mc)])
    chans = [mc]
    ch, ds = mc, 1
    for level, mult in enumerate(cfg.channel_mult):
        for _ in range(cfg.num_res_blocks):
            i = len(a.input_blocks)
            blk = [Layer('res', f'input_blocks.{i}.0', ch, mult * mc)]
            ch = mult * mc
            if ds in cfg.attention_resolutions:
                blk.append(Layer('st', f'input_blocks.{i}.1', ch, ch, heads=cfg.num_heads, d_head=ch // cfg.num_heads))
            a.input_blocks.append(blk)
            chans.append(ch)
        if level != len(cfg.channel_mult) - 1:
            i = len(a.input_blocks)
            a.input_blocks.append([Layer('res', f'input_blocks.{i}.0', ch, ch, down=True)])
            chans.append(ch)
            ds *= 2
    a.middle = [Layer('res', 'middle_block.0', ch, ch),
                Layer('st', 'middle_block.1', ch, ch, heads=cfg.num_heads, d_head=ch // cfg.num_heads),
                Layer('res', 'middle_block.2', ch, ch)]
    for level, mult in list(enumerate(cfg.channel_mult))[::-1]:
        for i in range(cfg.num_res_blocks + 1):
            ich = chans.pop()
            j = len(a.output_blocks)
            blk = [Layer('res', f'output_blocks.{j}.0', ch + ich, mc * mult)]
            ch = mc * mult
            if ds in cfg.attention_resolutions:
                blk.append(Layer('st', f'output_blocks.{j}.1', ch, ch, heads=cfg.num_heads, d_head=ch // cfg.num_heads))
            if level and i == cfg.num_res_blocks:
                blk.append(Layer('res', f'output_blocks.{j}.{len(blk)}', ch, ch, up=True))
                ds //= 2
            a.output_blocks.append(blk)
    a.head_channels = ch
    return a


def _res_params(l: Layer, emb_ch: int) -> List[Tuple[str, Tuple[int, ...]]]:
    p = l.prefix
    out = [(f'{p}.in_layers.0.weight', (l.cin,)), (f'{p}.in_layers.0.bias', (l.cin,)),
           (f'{p}.in_layers.2.weight', (l.cout, l.cin, 3, 3)), (f'{p}.in_layers.2.bias', (l.cout,)),
           (f'{p}.emb_layers.1.weight', (2 * l.cout, emb_ch)), (f'{p}.emb_layers.1.bias', (2 * l.cout,)),
           (f'{p}.out_layers.0.weight', (l.cout,)), (f'{p}.out_layers.0.bias', (l.cout,)),
           (f'{p}.out_layers.3.weight', (l.cout, l.cout, 3, 3)), (f'{p}.out_layers.3.bias', (l.cout,))]
    if l.cin != l.cout:
        out += [(f'{p}.skip_connection.weight', (l.cout, l.cin, 1, 1)), (f'{p}.skip_connection.bias', (l.cout,))]
    return out


def _st_params(l: Layer, ctx_dim: int, depth: int) -> List[Tuple[str, Tuple[int, ...]]]:
    p, c = l.prefix, l.cin
    out = [(f'{p}.norm.weight', (c,)), (f'{p}.norm.bias', (c,)),
           (f'{p}.proj_in.weight', (c, c, 1, 1)), (f'{p}.proj_in.bias', (c,))]
    for d in range(depth):
        t = f'{p}.transformer_blocks.{d}'
        out += [(f'{t}.attn1.to_q.weight', (c, c)), (f'{t}.attn1.to_k.weight', (c, c)),
                (f'{t}.attn1.to_v.weight', (c, c)), (f'{t}.attn1.to_out.0.weight', (c, c)),
                (f'{t}.attn1.to_out.0.bias', (c,)),
                (f'{t}.ff.net.0.proj.weight', (8 * c, c)), (f'{t}.ff.net.0.proj.bias', (8 * c,)),
                (f'{t}.ff.net.2.weight', (c, 4 * c)), (f'{t}.ff.net.2.bias', (c,)),
                (f'{t}.attn2.to_q.weight', (c, c)), (f'{t}.attn2.to_k.weight', (c, ctx_dim)),
                (f'{t}.attn2.to_v.weight', (c, ctx_dim)), (f'{t}.attn2.to_out.0.weight', (c, c)),
                (f'{t}.attn2.to_out.0.bias', (c,)),
                (f'{t}.norm1.weight', (c,)), (f'{t}.norm1.bias', (c,)),
                (f'{t}.norm2.weight', (c,)), (f'{t}.norm2.bias', (c,)),
                (f'{t}.norm3.weight', (c,)), (f'{t}.norm3.bias', (c,))]
    out += [(f'{p}.proj_out.weight', (c, c, 1, 1)), (f'{p}.proj_out.bias', (c,))]
    return out


def param_specs(cfg: UNetConfig) -> List[Tuple[str, Tuple[int, ...]]]:
    """(name, shape) in the order of the reference ``UNetModel.named_parameters()`` (module registration order:
    time_embed, input_blocks, middle_block, output_blocks, out; BasicTransformerBlock registers attn1, ff, attn2,
    norm1-3 in that order, attention.py:199-205)."""
    a = build_arch(cfg)
    mc, emb = cfg.model_channels, 4 * cfg.model_channels
    specs: List[Tuple[str, Tuple[int, ...]]] = [
        ('time_embed.0.weight', (emb, mc)), ('time_embed.0.bias', (emb,)),
        ('time_embed.2.weight', (emb, emb)), ('time_embed.2.bias', (emb,))]

    def add(layers):
        for l in layers:
            if l.kind == 'conv_in':
                specs.extend([(f'{l.prefix}.weight', (l.cout, l.cin, 3, 3)), (f'{l.prefix}.bias', (l.cout,))])
            elif l.kind == 'res':
                specs.extend(_res_params(l, emb))
            else:
                specs.extend(_st_params(l, cfg.context_dim, cfg.transformer_depth))

    for blk in a.input_blocks:
        add(blk)
    add(a.middle)
    for blk in a.output_blocks:
        add(blk)
    specs += [('out.0.weight', (a.head_channels,)), ('out.0.bias', (a.head_channels,)),
              ('out.2.weight', (cfg.out_channels, mc, 3, 3)), ('out.2.bias', (cfg.out_channels,))]
    return specs


def make_params(cfg: UNetConfig, seed: int = 1, dtype=torch.float32) -> Params:
    """Deterministic non-trivial weights for parity runs: tensor i is drawn from its own ``Generator(seed*100003+i)``
    so both the reference module (via load_state_dict) and the CUDA path get bit-identical parameters without
    depending on module construction order.  Matrices/filters ~ N(0, 1/fan_in), norm scales ~ 1 + 0.1 N,
    biases ~ 0.1 N (the zero-initialised reference layers must be non-zero or every check is 0 == 0)."""
    out: Params = {}
    for i, (name, shape) in enumerate(param_specs(cfg)):
        g = torch.Generator().manual_seed(seed * 100003 + i)
        r = torch.randn(shape, generator=g, dtype=torch.float32)
        if len(shape) == 1:
            is_norm_scale = name.endswith('weight')
            t = 1.0 + 0.1 * r if is_norm_scale else 0.1 * r
        else:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            t = r / math.sqrt(fan_in)
        out[name] = t.to(dtype)
    return out


# ------------------------------------------------------------------------------------------------------------------
# forward (restates openaimodel_enc.py:712-748, attention.py:37-64,170-261, util.py:179-199,242-244)
# ------------------------------------------------------------------------------------------------------------------
def timestep_embedding(timesteps: torch.Tensor, dim: int, max_period: float = 10000.0) -> torch.Tensor:
    """Sinusoidal embedding, cosine half first (diffusionmodules/util.py:179-199)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half).to(timesteps.device)
    args = timesteps[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def group_norm32(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, eps: float) -> torch.Tensor:
    """GroupNorm with 32 groups evaluated in fp32 (util.py:227-244; eps 1e-5 in ResBlock/head, 1e-6 in
    SpatialTransformer, attention.py:76-77)."""
    return F.group_norm(x.float(), 32, w, b, eps).type(x.dtype)


def resblock(p: Params, l: Layer, x: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """ResBlock._forward with use_scale_shift_norm (openaimodel_enc.py:255-275)."""
    pre = l.prefix
    h = F.silu(group_norm32(x, p[f'{pre}.in_layers.0.weight'], p[f'{pre}.in_layers.0.bias'], 1e-5))
    if l.down:    # Downsample(use_conv=False) == AvgPool2d(2) on both branches (:154-156,256-261)
        h, x = F.avg_pool2d(h, 2), F.avg_pool2d(x, 2)
    elif l.up:    # Upsample(use_conv=False) == nearest x2 (:116)
        h, x = F.interpolate(h, scale_factor=2, mode='nearest'), F.interpolate(x, scale_factor=2, mode='nearest')
    h = F.conv2d(h, p[f'{pre}.in_layers.2.weight'], p[f'{pre}.in_layers.2.bias'], padding=1)
    e = F.linear(F.silu(emb), p[f'{pre}.emb_layers.1.weight'], p[f'{pre}.emb_layers.1.bias'])
    scale, shift = e[:, :l.cout, None, None], e[:, l.cout:, None, None]
    h = group_norm32(h, p[f'{pre}.out_layers.0.weight'], p[f'{pre}.out_layers.0.bias'], 1e-5) * (1 + scale) + shift
    h = F.conv2d(F.silu(h), p[f'{pre}.out_layers.3.weight'], p[f'{pre}.out_layers.3.bias'], padding=1)
    if l.cin != l.cout:
        x = F.conv2d(x, p[f'{pre}.skip_connection.weight'], p[f'{pre}.skip_connection.bias'])
    return x + h


def attention(p: Params, pre: str, x: torch.Tensor, context, heads: int) -> torch.Tensor:
    """CrossAttention.forward (attention.py:170-193); context=None is self-attention."""
    ctx = x if context is None else context
    q = F.linear(x, p[f'{pre}.to_q.weight'])
    k = F.linear(ctx, p[f'{pre}.to_k.weight'])
    v = F.linear(ctx, p[f'{pre}.to_v.weight'])
    b, n, c = q.shape
    d = c // heads

    def split(t):
        return t.reshape(b, t.shape[1], heads, d).permute(0, 2, 1, 3)

    q, k, v = split(q), split(k), split(v)
    sim = torch.matmul(q, k.transpose(-1, -2)) * (d ** -0.5)
    out = torch.matmul(sim.softmax(dim=-1), v)
    out = out.permute(0, 2, 1, 3).reshape(b, n, c)
    return F.linear(out, p[f'{pre}.to_out.0.weight'], p[f'{pre}.to_out.0.bias'])


def feed_forward(p: Params, pre: str, x: torch.Tensor) -> torch.Tensor:
    """FeedForward with GEGLU, exact-erf GELU, value half first (attention.py:37-64)."""
    y = F.linear(x, p[f'{pre}.net.0.proj.weight'], p[f'{pre}.net.0.proj.bias'])
    val, gate = y.chunk(2, dim=-1)
    return F.linear(val * F.gelu(gate), p[f'{pre}.net.2.weight'], p[f'{pre}.net.2.bias'])


def transformer_block(p: Params, pre: str, x: torch.Tensor, context: torch.Tensor, heads: int) -> torch.Tensor:
    """BasicTransformerBlock._forward (attention.py:211-215); the checkpoint wrapper changes no values."""
    def ln(t, i):
        return F.layer_norm(t, t.shape[-1:], p[f'{pre}.norm{i}.weight'], p[f'{pre}.norm{i}.bias'], 1e-5)
    x = attention(p, f'{pre}.attn1', ln(x, 1), None, heads) + x
    x = attention(p, f'{pre}.attn2', ln(x, 2), context, heads) + x
    x = feed_forward(p, f'{pre}.ff', ln(x, 3)) + x
    return x


def spatial_transformer(p: Params, l: Layer, x: torch.Tensor, context: torch.Tensor, depth: int = 1) -> torch.Tensor:
    """SpatialTransformer.forward (attention.py:250-261)."""
    pre = l.prefix
    b, c, hh, ww = x.shape
    h = F.group_norm(x, 32, p[f'{pre}.norm.weight'], p[f'{pre}.norm.bias'], 1e-6)
    h = F.conv2d(h, p[f'{pre}.proj_in.weight'], p[f'{pre}.proj_in.bias'])
    h = h.permute(0, 2, 3, 1).reshape(b, hh * ww, c)
    for d in range(depth):
        h = transformer_block(p, f'{pre}.transformer_blocks.{d}', h, context, l.heads)
    h = h.reshape(b, hh, ww, c).permute(0, 3, 1, 2)
    h = F.conv2d(h, p[f'{pre}.proj_out.weight'], p[f'{pre}.proj_out.bias'])
    return h + x


def _run_block(p: Params, cfg: UNetConfig, layers: List[Layer], h, emb, context):
    """TimestepEmbedSequential.forward type dispatch (openaimodel_enc.py:80-88)."""
    for l in layers:
        if l.kind == 'conv_in':
            h = F.conv2d(h, p[f'{l.prefix}.weight'], p[f'{l.prefix}.bias'], padding=1)
        elif l.kind == 'res':
            h = resblock(p, l, h, emb)
        else:
            h = spatial_transformer(p, l, h, context, cfg.transformer_depth)
    return h


def unet_forward(p: Params, cfg: UNetConfig, x: torch.Tensor, timesteps: torch.Tensor, context,
                 taps: Dict[str, torch.Tensor] = None) -> torch.Tensor:
    """UNetModel.forward (openaimodel_enc.py:712-748).  ``context`` is a sequence; only context[0], reshaped to
    (B, -1, context_dim), is used (:729).  ``taps`` (optional dict) receives per-block activations."""
    a = build_arch(cfg)
    b = x.shape[0]
    emb = timestep_embedding(timesteps, cfg.model_channels)
    emb = F.linear(emb, p['time_embed.0.weight'], p['time_embed.0.bias'])
    emb = F.linear(F.silu(emb), p['time_embed.2.weight'], p['time_embed.2.bias'])
    ctx = context[0].reshape(b, -1, cfg.context_dim)
    hs = []
    h = x
    for i, blk in enumerate(a.input_blocks):
        h = _run_block(p, cfg, blk, h, emb, ctx)
        hs.append(h)
        if taps is not None:
            taps[f'input_blocks.{i}'] = h
    h = _run_block(p, cfg, a.middle, h, emb, ctx)
    if taps is not None:
        taps['middle_block'] = h
    for i, blk in enumerate(a.output_blocks):
        h = torch.cat([h, hs.pop()], dim=1)   # h first, skip second (:742)
        h = _run_block(p, cfg, blk, h, emb, ctx)
        if taps is not None:
            taps[f'output_blocks.{i}'] = h
    h = F.silu(group_norm32(h, p['out.0.weight'], p['out.0.bias'], 1e-5))
    return F.conv2d(h, p['out.2.weight'], p['out.2.bias'], padding=1)
