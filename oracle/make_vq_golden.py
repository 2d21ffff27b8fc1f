"""Generate tests/golden/vq_encoder_{tiny,shapes3d}.npz by RUNNING THE UNMODIFIED REFERENCE first-stage encoder
(ldm.modules.diffusionmodules.model.Encoder imported from /root/reference, followed by the 1x1 ``quant_conv`` exactly as
VQModel builds it, ldm/models/autoencoder.py:30-51,313-316) on seeded parameters and images.  Only the inputs, the latent
and a few intermediate norms are stored; the parameters are regenerated from the seed (oracle.vq_oracle.make_params).
Run in the build container only:  python -m oracle.make_vq_golden"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle import vq_oracle as VO  # noqa: E402
from oracle.make_golden import GOLDEN  # noqa: E402


def reference_encoder(cfg: VO.VQConfig, params):
    ref_shim.reference_modules()                     # installs the stubs and puts the reference tree on sys.path
    from ldm.modules.diffusionmodules.model import Encoder
    enc = Encoder(ch=cfg.ch, out_ch=3, ch_mult=cfg.ch_mult, num_res_blocks=cfg.num_res_blocks, attn_resolutions=[],
                  dropout=0.0, in_channels=cfg.in_channels, resolution=cfg.resolution, z_channels=cfg.z_channels,
                  double_z=cfg.double_z)
    quant_conv = torch.nn.Conv2d(cfg.z_channels, cfg.embed_dim, 1)          # autoencoder.py:49
    missing, unexpected = enc.load_state_dict({k[len("encoder."):]: v for k, v in params.items() if k.startswith("encoder.")},
                                              strict=True)
    assert not missing and not unexpected
    quant_conv.load_state_dict({"weight": params["quant_conv.weight"], "bias": params["quant_conv.bias"]})
    names = [("encoder." + k) for k, _ in enc.named_parameters()] + ["quant_conv.weight", "quant_conv.bias"]
    assert names == [n for n, _ in VO.param_shapes(cfg)], "parameter order differs from the reference registration order"
    return enc.eval(), quant_conv.eval()


def reference_decoder(cfg: VO.VQConfig, params, disentangled_dim):
    ref_shim.reference_modules()
    from ldm.modules.diffusionmodules.model import Decoder
    dec = Decoder(ch=cfg.ch, out_ch=3, ch_mult=cfg.ch_mult, num_res_blocks=cfg.num_res_blocks, attn_resolutions=[],
                  dropout=0.0, in_channels=cfg.in_channels, resolution=cfg.resolution, z_channels=cfg.z_channels)
    pqc = torch.nn.Conv2d(cfg.embed_dim + disentangled_dim, cfg.z_channels, 1)        # autoencoder.py:50-53
    missing, unexpected = dec.load_state_dict({k[len("decoder."):]: v for k, v in params.items() if k.startswith("decoder.")},
                                              strict=True)
    assert not missing and not unexpected
    pqc.load_state_dict({"weight": params["post_quant_conv.weight"], "bias": params["post_quant_conv.bias"]})
    names = ["decoder." + k for k, _ in dec.named_parameters()]
    want = [n for n, _ in VO.dec_param_shapes(cfg, 0, disentangled_dim) if n.startswith("decoder.")]
    assert names == want, "decoder parameter order differs from the reference registration order"
    return dec.eval(), pqc.eval()


def decode_goldens():
    """Decode direction: the reference Decoder + post_quant_conv on an already-quantised latent (force_not_quantize=True
    semantics, autoencoder.py:341-366) with the disentangled representation concatenated as the Shapes3D yaml configures it
    (use_disentangled_concat, disentangled_dim 20).  The quantiser itself (taming, absent) is not part of these fixtures."""
    for name, cfg, B, dd in (("tiny", VO.TINY_VQ, 3, 4), ("shapes3d", VO.SHAPES3D_VQ, 2, 20)):
        params = VO.make_dec_params(cfg, seed=2, n_embed=64, disentangled_dim=dd)
        dec, pqc = reference_decoder(cfg, params, dd)
        g = torch.Generator().manual_seed(13)
        r = cfg.resolution // (2 ** (len(cfg.ch_mult) - 1))
        h = torch.randn(B, cfg.embed_dim, r, r, generator=g)
        s = torch.randn(B, dd, generator=g)
        with torch.no_grad():
            q = torch.cat([h, s[:, :, None, None].expand(-1, -1, r, r)], dim=1)
            img = dec(pqc(q))
            mine = VO.vq_decode(params, cfg, h, force_not_quantize=True, disentangled_repr=s, disentangled_dim=dd)
            img0 = dec(pqc(torch.cat([h, torch.zeros(B, dd, r, r)], dim=1)))
            mine0 = VO.vq_decode(params, cfg, h, force_not_quantize=True, disentangled_repr=None, disentangled_dim=dd)
        print(name, "img", tuple(img.shape), "absmax", img.abs().max().item(), "oracle vs reference",
              ((mine - img).abs().max() / img.abs().max()).item(), ((mine0 - img0).abs().max() / img0.abs().max()).item())
        np.savez_compressed(os.path.join(GOLDEN, f"vq_decoder_{name}.npz"), h=h.numpy(), s=s.numpy(), img=img.numpy(),
                            img_zero_s=img0.numpy(), disentangled_dim=dd,
                            n_params=sum(v.numel() for k, v in params.items() if not k.startswith("quantize.")))


def main():
    decode_goldens()
    for name, cfg, B in (("tiny", VO.TINY_VQ, 3), ("shapes3d", VO.SHAPES3D_VQ, 2)):
        params = VO.make_params(cfg, seed=1)
        enc, qc = reference_encoder(cfg, params)
        g = torch.Generator().manual_seed(11)
        x = torch.rand(B, cfg.in_channels, cfg.resolution, cfg.resolution, generator=g) * 2 - 1
        with torch.no_grad():
            h = enc(x)
            z = qc(h)
            zo = VO.vq_encode(params, cfg, x)
        print(name, "z", tuple(z.shape), "absmax", z.abs().max().item(), "oracle vs reference",
              ((zo - z).abs().max() / z.abs().max()).item())
        np.savez_compressed(os.path.join(GOLDEN, f"vq_encoder_{name}.npz"), x=x.numpy(), z=z.numpy(),
                            pre_quant=h.numpy(), n_params=sum(v.numel() for v in params.values()))


if __name__ == "__main__":
    main()
