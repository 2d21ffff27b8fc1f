"""The trunk oracle (oracle/enc4_oracle.py) against tests/golden/enc4_trunk.npz, which the UNMODIFIED reference Encoder4
produced (oracle/make_enc4_golden.py): codes, gradients (small ones in full, all by norm and random projection) and the
running statistics after one training-mode forward.  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import enc4_oracle as EO
from oracle.make_enc4_golden import projection

GOLD = os.path.join(os.path.dirname(__file__), "golden", "enc4_trunk.npz")


def test_param_inventory():
    shapes = EO.param_shapes()
    assert len(shapes) == 32 and shapes[0] == ("encoder.0.weight", (128, 3, 4, 4)) and shapes[-1] == ("encoder.16.bias", (20,))
    assert sum(int(np.prod(s)) for _, s in shapes) == 1_164_052


@pytest.mark.parametrize("mode", ["train", "eval"])
def test_oracle_matches_reference_fixture(mode):
    g = np.load(GOLD)
    B, seed = int(g["B"]), int(g["seed"])
    p, st = EO.make_params(seed=seed), EO.make_stats(seed=seed)
    x, dcodes = EO.seeded_case(B, seed)
    codes, grads = EO.forward_backward(p, st, x, dcodes, mode == "train")
    ref = torch.from_numpy(g[f"{mode}.codes"])
    assert (codes - ref).abs().max() <= 1e-5 * ref.abs().max()
    for k, gr in grads.items():
        n_ref, p_ref = float(g[f"{mode}.gnorm.{k}"]), float(g[f"{mode}.gproj.{k}"])
        assert abs(float(gr.double().norm()) - n_ref) <= 1e-4 * n_ref + 1e-9, k
        assert abs(projection(k, gr) - p_ref) <= 1e-3 * n_ref * gr.numel() ** 0.5 + 1e-9, k
        if f"{mode}.grad.{k}" in g:
            full = torch.from_numpy(g[f"{mode}.grad.{k}"])
            assert (gr - full).abs().max() <= 1e-4 * full.abs().max() + 1e-7, k
    if mode == "train":
        for k, v in st.items():
            assert np.allclose(v.numpy(), g[f"train.stat.{k}"], rtol=1e-5, atol=1e-6), k
