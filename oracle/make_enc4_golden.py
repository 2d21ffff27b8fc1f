"""Generate tests/golden/enc4_trunk.npz by RUNNING THE UNMODIFIED REFERENCE ``Encoder4`` (ldm.modules.diffusionmodules.
openaimodel_enc, imported from /root/reference) on seeded parameters, running statistics and images, in training and eval
mode, forward and backward.  Stored: codes, the small gradients in full, for every gradient its L2 norm and a seeded random
projection, and the updated running statistics.  Parameters and inputs are regenerated from the seed
(oracle.enc4_oracle).  Run in the build container only:  python -m oracle.make_enc4_golden"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import enc4_oracle as EO  # noqa: E402
from oracle import ref_shim  # noqa: E402
from oracle.make_golden import GOLDEN  # noqa: E402

B, SEED = 6, 11


def projection(name: str, g: torch.Tensor) -> float:
    r = torch.randn(g.shape, generator=torch.Generator().manual_seed(abs(hash_name(name)) % (2 ** 31)), dtype=torch.float64)
    return float((g.double() * r).sum())


def hash_name(name: str) -> int:
    h = 0
    for ch in name:
        h = (h * 131 + ord(ch)) % 1000003
    return h


def reference_run(training: bool):
    ref_shim.reference_modules()
    from ldm.modules.diffusionmodules.openaimodel_enc import Encoder4
    enc = Encoder4(d=128, context_dim=16, latent_unit=20)
    p, st = EO.make_params(seed=SEED), EO.make_stats(seed=SEED)
    sd = enc.state_dict()
    for k, v in list(p.items()) + list(st.items()):
        assert sd[k].shape == v.shape, k
        sd[k] = v.clone()
    enc.load_state_dict(sd, strict=True)
    names = ["encoder." + n for n, _ in enc.encoder.named_parameters()]
    assert names == [n for n, _ in EO.param_shapes()], "parameter order differs from the reference registration order"
    enc.train(training)
    x, dcodes = EO.seeded_case(B, SEED)
    codes = enc.encoding(x)
    codes.backward(dcodes)
    grads = {"encoder." + n: q.grad for n, q in enc.encoder.named_parameters()}
    stats = {k: v for k, v in enc.state_dict().items() if "running" in k}
    return codes.detach(), grads, stats


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    out = {"B": np.int64(B), "seed": np.int64(SEED)}
    for mode in ("train", "eval"):
        codes, grads, stats = reference_run(mode == "train")
        out[f"{mode}.codes"] = codes.numpy()
        for k, g in grads.items():
            out[f"{mode}.gnorm.{k}"] = np.float64(g.double().norm())
            out[f"{mode}.gproj.{k}"] = np.float64(projection(k, g))
            if g.numel() <= 128:
                out[f"{mode}.grad.{k}"] = g.numpy()
        if mode == "train":
            for k, v in stats.items():
                out[f"train.stat.{k}"] = v.numpy()
    path = os.path.join(GOLDEN, "enc4_trunk.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
