"""Concept-token MLPs (Encoder4.net / Encoder4.warp, openaimodel_enc.py:1017-1041).
CPU: the oracle against the reference-generated golden vectors (tokens and every gradient), the state_dict key mapping
of the stacked-parameter module.  GPU: the one-launch CUDA forward / backward against the golden vectors and against the
oracle at the shipped size (20 units, ragged batch)."""
import os

import numpy as np
import pytest
import torch

from oracle import concept_oracle as CO

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "concept_mlp_l5.npz")


def _golden():
    g = np.load(GOLDEN)
    state = {k[len("param::"):]: torch.tensor(g[k]) for k in g.files if k.startswith("param::")}
    grads = {k[len("grad::"):]: torch.tensor(g[k]) for k in g.files if k.startswith("grad::")}
    return g, state, grads


def test_oracle_matches_reference_golden():
    g, state, grads = _golden()
    L = int(g["latent_unit"])
    p = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    codes = torch.tensor(g["codes"]).requires_grad_(True)
    y = CO.concept_mlp(p, codes, L)
    assert torch.allclose(y, torch.tensor(g["tokens"]), rtol=0, atol=1e-6)
    (y * torch.tensor(g["loss_weight"])).sum().backward()
    assert torch.allclose(codes.grad, torch.tensor(g["dcodes"]), rtol=1e-5, atol=1e-6)
    for k, want in grads.items():
        assert torch.allclose(p[k].grad, want, rtol=1e-5, atol=1e-6), k


def test_state_dict_uses_the_reference_keys():
    from encdiff_b200.concept import ConceptMLP
    g, state, _ = _golden()
    L, Dc = int(g["latent_unit"]), int(g["context_dim"])
    m = ConceptMLP(L, Dc)
    assert set(m.state_dict().keys()) == set(state.keys())                 # '{i}.{0,2,4}.{weight,bias}'
    m.load_state_dict(state, strict=True)
    back = m.state_dict()
    for k, v in state.items():
        assert back[k].shape == v.shape and torch.equal(back[k], v), k
    assert m.w2.shape == (L, 128, 64) and torch.equal(m.w1[3], state["3.0.weight"].reshape(64))
    with pytest.raises(RuntimeError):                                       # no CPU fallback
        m(torch.zeros(2, L))


def test_seeded_init_follows_the_reference_construction_order():
    from encdiff_b200.concept import ConceptMLP
    torch.manual_seed(5)
    m = ConceptMLP(3, 16)
    want = CO.make_state(3, 16, seed=5)
    for k, v in m.state_dict().items():
        assert torch.equal(v, want[k]), k


@pytest.mark.gpu
def test_cuda_vs_reference_golden(cuda_device):
    from encdiff_b200.concept import ConceptMLP
    g, state, grads = _golden()
    L, Dc = int(g["latent_unit"]), int(g["context_dim"])
    m = ConceptMLP(L, Dc)
    m.load_state_dict(state, strict=True)
    m = m.to(cuda_device)
    codes = torch.tensor(g["codes"]).to(cuda_device).requires_grad_(True)
    y = m(codes)
    assert torch.allclose(y.cpu(), torch.tensor(g["tokens"]), rtol=0, atol=2e-6)
    y3 = m.warp(codes.detach().reshape(-1, L, 1))
    assert y3.shape == (codes.shape[0], L, Dc) and torch.equal(y3.reshape(y.shape), y.detach())
    (y * torch.tensor(g["loss_weight"]).to(cuda_device)).sum().backward()
    assert torch.allclose(codes.grad.cpu(), torch.tensor(g["dcodes"]), rtol=1e-4, atol=2e-6)
    got = {k: v for k, v in zip(["w1", "b1", "w2", "b2", "w3", "b3"], [m.w1, m.b1, m.w2, m.b2, m.w3, m.b3])}
    for mine, ref in ConceptMLP._MAP:
        for i in range(L):
            want = grads[f"{i}.{ref}"].reshape(got[mine].grad[i].shape)
            assert torch.allclose(got[mine].grad[i].cpu(), want, rtol=1e-4, atol=2e-6), (mine, i)


@pytest.mark.gpu
@pytest.mark.parametrize("B", [1, 37, 256])
def test_cuda_vs_oracle_shipped_size(cuda_device, B):
    from encdiff_b200.concept import ConceptMLP
    L, Dc = 20, 16
    state = CO.make_state(L, Dc, seed=2)
    m = ConceptMLP(L, Dc)
    m.load_state_dict(state, strict=True)
    m = m.to(cuda_device)
    gen = torch.Generator().manual_seed(B)
    codes = torch.randn(B, L, generator=gen) * 2.0
    w = torch.randn(B, L * Dc, generator=gen)
    p = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    co = codes.clone().requires_grad_(True)
    yo = CO.concept_mlp(p, co, L)
    (yo * w).sum().backward()
    cd = codes.to(cuda_device).requires_grad_(True)
    y = m(cd)
    (y * w.to(cuda_device)).sum().backward()
    assert torch.allclose(y.detach().cpu(), yo.detach(), rtol=0, atol=5e-6)
    assert torch.allclose(cd.grad.cpu(), co.grad, rtol=1e-4, atol=1e-5)
    sd_grad = {"w1": m.w1.grad, "b1": m.b1.grad, "w2": m.w2.grad, "b2": m.b2.grad, "w3": m.w3.grad, "b3": m.b3.grad}
    worst = 0.0
    for mine, ref in ConceptMLP._MAP:
        want = torch.stack([p[f"{i}.{ref}"].grad.reshape(sd_grad[mine][i].shape) for i in range(L)])
        err = ((sd_grad[mine].cpu() - want).abs().max() / want.abs().max().clamp_min(1e-12)).item()
        worst = max(worst, err)
    print(f"B={B}: worst parameter-gradient rel err {worst:.2e}")
    assert worst < 1e-4            # fp32 on both sides; only the summation order over the samples differs
