"""Pins oracle/vq_oracle.py (the CPU restatement of the first-stage VQ-f4 encode path) to the
fixtures oracle/make_vq_golden.py produced by running the UNMODIFIED reference Encoder + quant_conv
(ldm/modules/diffusionmodules/model.py:368-459, ldm/models/autoencoder.py:313-316) on the same seeded parameters."""
import os

import numpy as np
import pytest
import torch

from oracle import vq_oracle as VO

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name,cfg", [("tiny", VO.TINY_VQ), ("shapes3d", VO.SHAPES3D_VQ)])
def test_vq_oracle_vs_reference_golden(name, cfg):
    g = np.load(os.path.join(GOLDEN, f"vq_encoder_{name}.npz"))
    params = VO.make_params(cfg, seed=1)
    assert sum(v.numel() for v in params.values()) == int(g["n_params"])
    with torch.no_grad():
        z = VO.vq_encode(params, cfg, torch.tensor(g["x"]))
    err = (z - torch.tensor(g["z"])).abs().max().item() / np.abs(g["z"]).max()
    assert z.shape == g["z"].shape and err <= 2e-5, err


def test_vq_param_inventory_matches_reference_order():
    """encoder.* in the reference's registration order, then quant_conv (asserted against the live module by
    oracle/make_vq_golden.py when the fixtures are generated); Shapes3D first stage: ch 32, ch_mult (1, 2, 4)."""
    names = [n for n, _ in VO.param_shapes(VO.SHAPES3D_VQ)]
    assert names[0] == "encoder.conv_in.weight" and names[-1] == "quant_conv.bias"
    assert "encoder.down.1.block.0.nin_shortcut.weight" in names and "encoder.down.0.block.0.nin_shortcut.weight" not in names
    assert "encoder.down.1.downsample.conv.weight" in names and "encoder.down.2.downsample.conv.weight" not in names
    assert "encoder.mid.attn_1.proj_out.bias" in names


@pytest.mark.parametrize("name,cfg", [("tiny", VO.TINY_VQ), ("shapes3d", VO.SHAPES3D_VQ)])
def test_vq_decode_oracle_vs_reference_golden(name, cfg):
    g = np.load(os.path.join(GOLDEN, f"vq_decoder_{name}.npz"))
    dd = int(g["disentangled_dim"])
    params = VO.make_dec_params(cfg, seed=2, n_embed=64, disentangled_dim=dd)
    assert sum(v.numel() for k, v in params.items() if not k.startswith("quantize.")) == int(g["n_params"])
    with torch.no_grad():
        img = VO.vq_decode(params, cfg, torch.tensor(g["h"]), force_not_quantize=True, disentangled_repr=torch.tensor(g["s"]),
                           disentangled_dim=dd)
        img0 = VO.vq_decode(params, cfg, torch.tensor(g["h"]), force_not_quantize=True, disentangled_dim=dd)
    assert (img - torch.tensor(g["img"])).abs().max().item() / np.abs(g["img"]).max() <= 2e-5
    assert (img0 - torch.tensor(g["img_zero_s"])).abs().max().item() / np.abs(g["img_zero_s"]).max() <= 2e-5


def test_vq_nearest_is_argmin_of_squared_distance():
    g = torch.Generator().manual_seed(3)
    z, e = torch.randn(2, 3, 4, 4, generator=g), torch.randn(50, 3, generator=g)
    q = VO.vq_nearest(z, e)
    zf = z.permute(0, 2, 3, 1).reshape(-1, 3)
    idx = ((zf[:, None] - e[None]) ** 2).sum(-1).argmin(1)
    assert torch.equal(q.permute(0, 2, 3, 1).reshape(-1, 3), e[idx])
