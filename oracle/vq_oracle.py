"""CPU restatement (plain torch ops on a {name: tensor} dict) of the first-stage VQ-f4 ENCODE and DECODE paths of the reference
(decode: ``VQModelInterface.decode``, ldm/models/autoencoder.py:328-369 = nearest-code quantisation + optional concat of the
disentangled representation + post_quant_conv + Decoder, ldm/modules/diffusionmodules/model.py:462-568; see vq_decode):
encode:
``VQModelInterface.encode`` = ``quant_conv(Encoder(x))`` (ldm/models/autoencoder.py:313-316; Encoder:
ldm/modules/diffusionmodules/model.py:368-459; ResnetBlock :82-143; AttnBlock :153-204; Downsample :63-79; Normalize
:37-38; swish :32-34).  TEST INFRASTRUCTURE ONLY: imported by tests/ and bench.py's baseline legs, never by
encdiff_b200/.  Pinned to tests/golden/vq_encoder_*.npz, which oracle/make_vq_golden.py produced by running the
UNMODIFIED reference Encoder + quant_conv on the same seeded parameters (tests/test_vq_oracle.py)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Tuple

import torch
import torch.nn.functional as F


@dataclass(frozen=True)
class VQConfig:
    """ddconfig of configs/latent-diffusion/shapes3d-vq-4-16-encdiff.yaml:61-75 + embed_dim (:53)."""
    ch: int = 32
    ch_mult: Tuple[int, ...] = (1, 2, 4)
    num_res_blocks: int = 2
    in_channels: int = 3
    resolution: int = 64
    z_channels: int = 3
    embed_dim: int = 3
    double_z: bool = False


SHAPES3D_VQ = VQConfig()
TINY_VQ = VQConfig(ch=32, ch_mult=(1, 2), num_res_blocks=1, resolution=16)


def param_shapes(cfg: VQConfig):
    """(name, shape) in the registration order of the reference modules (= state_dict / named_parameters order of
    ``encoder.*`` followed by ``quant_conv.*`` as VQModel registers them, autoencoder.py:30-51)."""
    out = []

    def conv(name, co, ci, k):
        out.append((name + ".weight", (co, ci, k, k)))
        out.append((name + ".bias", (co,)))

    def norm(name, c):
        out.append((name + ".weight", (c,)))
        out.append((name + ".bias", (c,)))

    def res(name, ci, co):
        norm(name + ".norm1", ci)
        conv(name + ".conv1", co, ci, 3)
        norm(name + ".norm2", co)
        conv(name + ".conv2", co, co, 3)
        if ci != co:
            conv(name + ".nin_shortcut", co, ci, 1)

    conv("encoder.conv_in", cfg.ch, cfg.in_channels, 3)
    in_mult = (1,) + tuple(cfg.ch_mult)
    block_in = cfg.ch
    for lvl in range(len(cfg.ch_mult)):
        block_in = cfg.ch * in_mult[lvl]
        block_out = cfg.ch * cfg.ch_mult[lvl]
        for i in range(cfg.num_res_blocks):
            res(f"encoder.down.{lvl}.block.{i}", block_in, block_out)
            block_in = block_out
        if lvl != len(cfg.ch_mult) - 1:
            conv(f"encoder.down.{lvl}.downsample.conv", block_in, block_in, 3)
    res("encoder.mid.block_1", block_in, block_in)
    norm("encoder.mid.attn_1.norm", block_in)
    for n in ("q", "k", "v", "proj_out"):
        conv(f"encoder.mid.attn_1.{n}", block_in, block_in, 1)
    res("encoder.mid.block_2", block_in, block_in)
    norm("encoder.norm_out", block_in)
    zc = 2 * cfg.z_channels if cfg.double_z else cfg.z_channels
    conv("encoder.conv_out", zc, block_in, 3)
    conv("quant_conv", cfg.embed_dim, cfg.z_channels, 1)
    return out


def make_params(cfg: VQConfig, seed: int = 1) -> Dict[str, torch.Tensor]:
    """Seeded parameters of reference-like scale (conv weights ~ U(-1/sqrt(fan_in), +), norm weights around 1): the SAME
    dict is loaded into the reference modules (make_vq_golden.py), into this oracle and into encdiff_b200.VQEncoder."""
    g = torch.Generator().manual_seed(seed)
    p = {}
    for name, shape in param_shapes(cfg):
        if ".norm" in name and name.endswith(".weight"):
            p[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif ".norm" in name:
            p[name] = 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".weight"):
            fan_in = shape[1] * shape[2] * shape[3]
            p[name] = (torch.rand(shape, generator=g) * 2 - 1) * (3.0 / fan_in) ** 0.5
        else:
            p[name] = 0.05 * torch.randn(shape, generator=g)
    return p


def _norm(p, name, x):                       # Normalize, model.py:37-38
    return F.group_norm(x, 32, p[name + ".weight"], p[name + ".bias"], eps=1e-6)


def _swish(x):                               # nonlinearity, model.py:32-34
    return x * torch.sigmoid(x)


def _conv(p, name, x, stride=1, padding=1):
    return F.conv2d(x, p[name + ".weight"], p[name + ".bias"], stride=stride, padding=padding)


def resnet_block(p, name, x):                # ResnetBlock.forward, model.py:121-143 (temb is None, dropout 0)
    h = _conv(p, name + ".conv1", _swish(_norm(p, name + ".norm1", x)))
    h = _conv(p, name + ".conv2", _swish(_norm(p, name + ".norm2", h)))
    if name + ".nin_shortcut.weight" in p:
        x = _conv(p, name + ".nin_shortcut", x, padding=0)
    return x + h


def attn_block(p, name, x):                  # AttnBlock.forward, model.py:178-204
    h = _norm(p, name + ".norm", x)
    q = _conv(p, name + ".q", h, padding=0)
    k = _conv(p, name + ".k", h, padding=0)
    v = _conv(p, name + ".v", h, padding=0)
    b, c, hh, ww = q.shape
    w_ = torch.bmm(q.reshape(b, c, hh * ww).permute(0, 2, 1), k.reshape(b, c, hh * ww)) * (int(c) ** (-0.5))
    w_ = torch.softmax(w_, dim=2)
    h = torch.bmm(v.reshape(b, c, hh * ww), w_.permute(0, 2, 1)).reshape(b, c, hh, ww)
    return x + _conv(p, name + ".proj_out", h, padding=0)


def downsample(p, name, x):                  # Downsample.forward with_conv, model.py:72-79: pad right/bottom, stride 2
    return _conv(p, name + ".conv", F.pad(x, (0, 1, 0, 1), mode="constant", value=0), stride=2, padding=0)


def vq_encode(p: Dict[str, torch.Tensor], cfg: VQConfig, x: torch.Tensor) -> torch.Tensor:
    """x [B, 3, R, R] in [-1, 1] -> latent [B, embed_dim, R/2^(levels-1), ...]: Encoder.forward (model.py:437-459) then
    quant_conv (autoencoder.py:313-316)."""
    h = _conv(p, "encoder.conv_in", x)
    for lvl in range(len(cfg.ch_mult)):
        for i in range(cfg.num_res_blocks):
            h = resnet_block(p, f"encoder.down.{lvl}.block.{i}", h)
        if lvl != len(cfg.ch_mult) - 1:
            h = downsample(p, f"encoder.down.{lvl}.downsample", h)
    h = resnet_block(p, "encoder.mid.block_1", h)
    h = attn_block(p, "encoder.mid.attn_1", h)
    h = resnet_block(p, "encoder.mid.block_2", h)
    h = _conv(p, "encoder.conv_out", _swish(_norm(p, "encoder.norm_out", h)))
    return _conv(p, "quant_conv", h, padding=0)


# ---------------------------------------------------------------------------------------------------------------
# decode
# ---------------------------------------------------------------------------------------------------------------
def dec_param_shapes(cfg: VQConfig, n_embed: int = 0, disentangled_dim: int = 0):
    """(name, shape) of the decode half in the reference's registration order: ``decoder.*`` (conv_in, mid, up.0 .. up.L-1 --
    Decoder.__init__ PREPENDS the levels it creates from the deepest one, model.py:505-527 -- norm_out, conv_out),
    ``quantize.embedding.weight`` when n_embed > 0, ``post_quant_conv.*`` (embed_dim + disentangled_dim inputs,
    autoencoder.py:50-53)."""
    out = []

    def conv(name, co, ci, k):
        out.append((name + ".weight", (co, ci, k, k)))
        out.append((name + ".bias", (co,)))

    def norm(name, c):
        out.append((name + ".weight", (c,)))
        out.append((name + ".bias", (c,)))

    def res(name, ci, co):
        norm(name + ".norm1", ci)
        conv(name + ".conv1", co, ci, 3)
        norm(name + ".norm2", co)
        conv(name + ".conv2", co, co, 3)
        if ci != co:
            conv(name + ".nin_shortcut", co, ci, 1)

    L = len(cfg.ch_mult)
    block_in = cfg.ch * cfg.ch_mult[L - 1]
    conv("decoder.conv_in", block_in, cfg.z_channels, 3)
    res("decoder.mid.block_1", block_in, block_in)
    norm("decoder.mid.attn_1.norm", block_in)
    for n in ("q", "k", "v", "proj_out"):
        conv(f"decoder.mid.attn_1.{n}", block_in, block_in, 1)
    res("decoder.mid.block_2", block_in, block_in)
    per_level = {}
    for lvl in reversed(range(L)):
        names_before = len(out)
        block_out = cfg.ch * cfg.ch_mult[lvl]
        for i in range(cfg.num_res_blocks + 1):
            res(f"decoder.up.{lvl}.block.{i}", block_in, block_out)
            block_in = block_out
        if lvl != 0:
            conv(f"decoder.up.{lvl}.upsample.conv", block_in, block_in, 3)
        per_level[lvl] = out[names_before:]
        del out[names_before:]
    for lvl in range(L):
        out.extend(per_level[lvl])
    norm("decoder.norm_out", block_in)
    conv("decoder.conv_out", 3, block_in, 3)
    if n_embed > 0:
        out.append(("quantize.embedding.weight", (n_embed, cfg.embed_dim)))
    conv("post_quant_conv", cfg.z_channels, cfg.embed_dim + disentangled_dim, 1)
    return out


def make_dec_params(cfg: VQConfig, seed: int = 2, n_embed: int = 0, disentangled_dim: int = 0) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    p = {}
    for name, shape in dec_param_shapes(cfg, n_embed, disentangled_dim):
        if name == "quantize.embedding.weight":
            p[name] = torch.randn(shape, generator=g)                   # codes on the scale of the latents
        elif ".norm" in name and name.endswith(".weight"):
            p[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif ".norm" in name:
            p[name] = 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".weight"):
            fan_in = shape[1] * shape[2] * shape[3]
            p[name] = (torch.rand(shape, generator=g) * 2 - 1) * (3.0 / fan_in) ** 0.5
        else:
            p[name] = 0.05 * torch.randn(shape, generator=g)
    return p


def vq_nearest(z: torch.Tensor, codebook: torch.Tensor) -> torch.Tensor:
    """The forward value of taming's VectorQuantizer2 (submodule src/taming-transformers @3ba01b24, ABSENT from the reference
    tree -- restated from its published algorithm, taming/modules/vqvae/quantize.py: d = |z|^2 + |e|^2 - 2 z.e over the
    flattened (b h w) positions, argmin, embedding lookup; the straight-through / loss terms do not change the value).
    Parity of this one function is unpinned; everything after it is pinned to the reference Decoder."""
    b, c, h, w = z.shape
    zf = z.permute(0, 2, 3, 1).reshape(-1, c)
    d = (zf ** 2).sum(1, keepdim=True) + (codebook ** 2).sum(1)[None] - 2 * zf @ codebook.t()
    idx = d.argmin(1)
    return codebook[idx].reshape(b, h, w, c).permute(0, 3, 1, 2).contiguous()


def upsample(p, name, x):                    # Upsample.forward with_conv, model.py:53-57: nearest x2 then conv3x3
    return _conv(p, name + ".conv", F.interpolate(x, scale_factor=2.0, mode="nearest"))


def decoder_forward(p: Dict[str, torch.Tensor], cfg: VQConfig, z: torch.Tensor) -> torch.Tensor:
    """Decoder.forward, model.py:532-568 (give_pre_end False, tanh_out False)."""
    h = _conv(p, "decoder.conv_in", z)
    h = resnet_block(p, "decoder.mid.block_1", h)
    h = attn_block(p, "decoder.mid.attn_1", h)
    h = resnet_block(p, "decoder.mid.block_2", h)
    for lvl in reversed(range(len(cfg.ch_mult))):
        for i in range(cfg.num_res_blocks + 1):
            h = resnet_block(p, f"decoder.up.{lvl}.block.{i}", h)
        if lvl != 0:
            h = upsample(p, f"decoder.up.{lvl}.upsample", h)
    return _conv(p, "decoder.conv_out", _swish(_norm(p, "decoder.norm_out", h)))


def vq_decode(p: Dict[str, torch.Tensor], cfg: VQConfig, h: torch.Tensor, force_not_quantize: bool = False,
              disentangled_repr=None, disentangled_dim: int = 0) -> torch.Tensor:
    """VQModelInterface.decode, autoencoder.py:328-369: quantise unless told not to, concatenate the disentangled
    representation broadcast over the latent grid (zeros when none is given) if the model was built with
    use_disentangled_concat, post_quant_conv, Decoder."""
    quant = h if force_not_quantize else vq_nearest(h, p["quantize.embedding.weight"])
    if disentangled_dim > 0:
        B, _, H, W = quant.shape
        s = (disentangled_repr[:, :, None, None].expand(-1, -1, H, W) if disentangled_repr is not None
             else torch.zeros(B, disentangled_dim, H, W, dtype=quant.dtype))
        quant = torch.cat([quant, s], dim=1)
    quant = _conv(p, "post_quant_conv", quant, padding=0)
    return decoder_forward(p, cfg, quant)
