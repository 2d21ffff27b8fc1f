"""Generate tests/golden/concept_mlp_l5.npz by RUNNING THE UNMODIFIED REFERENCE Encoder4 (imported from /root/reference):
its ``warp`` on seeded codes, plus the gradients of a fixed scalar loss w.r.t. codes and every MLP parameter.
Run in the build container only:  python -m oracle.make_concept_golden"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import GOLDEN  # noqa: E402


def main():
    _, openaimodel_enc, _ = ref_shim.reference_modules()
    torch.manual_seed(7)
    L, Dc, B = 5, 16, 11
    enc = openaimodel_enc.Encoder4(d=128, context_dim=Dc, latent_unit=L).eval()
    g = torch.Generator().manual_seed(3)
    codes = (torch.randn(B, L, generator=g) * 1.5).requires_grad_(True)
    w = torch.randn(B, L * Dc, generator=g)
    y3 = enc.warp(codes.reshape(B, L, 1))                # openaimodel_enc.py:1037-1041: [B, L, Dc] for 3-D codes
    assert y3.shape == (B, L, Dc) and torch.allclose(y3.reshape(B, -1), enc.warp(codes), atol=1e-6)   # 2-D codes: [B, L * Dc]
    y = y3.reshape(B, -1)
    (y * w).sum().backward()
    out = {"codes": codes.detach().numpy(), "loss_weight": w.numpy(), "tokens": y.detach().numpy(),
           "dcodes": codes.grad.numpy(), "latent_unit": L, "context_dim": Dc}
    for k, v in enc.net.state_dict().items():
        out["param::" + k] = v.numpy()
    for k, p in enc.net.named_parameters():
        out["grad::" + k] = p.grad.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "concept_mlp_l5.npz"), **out)
    print("tokens absmax", y.abs().max().item(), "keys", len(out))


if __name__ == "__main__":
    main()
