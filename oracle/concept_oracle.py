"""CPU oracle of the concept-token MLPs -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (only tests/ import it).

Restates the tail of ``Encoder4.forward`` / ``Encoder4.warp`` (ldm/modules/diffusionmodules/openaimodel_enc.py:1017-1041):
for every latent unit i, ``net[i] = Linear(1,64) - ELU - Linear(64,128) - ELU - Linear(128,context_dim)`` applied to
``x[:, i][:, None]`` and the results concatenated along dim 1.  Pinned to the reference by
tests/golden/concept_mlp_l5.npz (oracle/make_concept_golden.py ran the reference's own Encoder4)."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F


def concept_mlp(state: Dict[str, torch.Tensor], codes: torch.Tensor, latent_unit: int) -> torch.Tensor:
    """state: the reference's ``Encoder4.net.state_dict()`` ('{i}.0.weight' [64,1], '{i}.0.bias', '{i}.2.*', '{i}.4.*');
    codes [B, latent_unit] (or [B, latent_unit, 1]) -> tokens [B, latent_unit * context_dim]."""
    if codes.dim() == 3:
        codes = codes.reshape(codes.shape[0], -1)
    out = []
    for i in range(latent_unit):                                   # openaimodel_enc.py:1029-1031 / 1039-1040
        h = codes[:, i][:, None]
        h = F.elu(F.linear(h, state[f"{i}.0.weight"], state[f"{i}.0.bias"]))
        h = F.elu(F.linear(h, state[f"{i}.2.weight"], state[f"{i}.2.bias"]))
        out.append(F.linear(h, state[f"{i}.4.weight"], state[f"{i}.4.bias"]))
    return torch.cat(out, dim=1)


def make_state(latent_unit: int, context_dim: int, seed: int) -> Dict[str, torch.Tensor]:
    """Seeded parameters in the reference's layout (nn.Linear default init, reference construction order)."""
    g = torch.random.get_rng_state()
    torch.manual_seed(seed)
    state = {}
    for i in range(latent_unit):
        for idx, lin in ((0, torch.nn.Linear(1, 64)), (2, torch.nn.Linear(64, 128)), (4, torch.nn.Linear(128, context_dim))):
            state[f"{i}.{idx}.weight"] = lin.weight.detach().clone()
            state[f"{i}.{idx}.bias"] = lin.bias.detach().clone()
    torch.random.set_rng_state(g)
    return state
