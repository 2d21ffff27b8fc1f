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


def main():
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
