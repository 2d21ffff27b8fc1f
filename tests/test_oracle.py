"""Pins the CPU oracle (oracle/) to the reference: against the committed golden vectors that
oracle/make_golden.py produced by running the unmodified reference, and — when /root/reference is present (build
container) — against the live reference modules.  fp32 on both sides: tolerance 2e-5 relative (op order only)."""
import os

import numpy as np
import pytest
import torch

from oracle import diffusion_oracle as DO
from oracle import ref_shim
from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def rel_err(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def test_param_tree_matches_reference_anchor():
    g = load("init_anchor.npz")
    specs = UO.param_specs(UO.SHAPES3D)
    assert len(specs) == int(g["n_tensors"]) == 734
    assert sum(int(np.prod(s)) for _, s in specs) == int(g["n_params"]) == 37469635
    assert [n for n, _ in specs] == list(g["names"])
    assert [str(tuple(s)) for _, s in specs] == list(g["shapes"])
    # zero-initialised output conv => eps == 0 and first-step L1 == E|N(0,1)| ~ 0.798
    assert float(g["eps_absmax"]) == 0.0
    assert abs(float(g["loss"]) - 0.798) < 0.01


def test_schedule_and_embedding():
    g = load("schedule.npz")
    betas = DO.linear_betas()
    assert np.array_equal(betas, g["betas"])
    acp = DO.alphas_cumprod(betas)
    for S in (50, 200):
        co = DO.ddim_coefficients(acp, S, 0.0)
        assert np.array_equal(co["timesteps"], g[f"timesteps_{S}"])
        assert np.array_equal(co["a_t"], g[f"alphas_{S}"])
        assert np.allclose(co["a_prev"], g[f"alphas_prev_{S}"].astype(np.float32), rtol=0, atol=0)
    assert list(DO.ddim_timesteps(50)[:3]) == [1, 21, 41] and DO.ddim_timesteps(50)[-1] == 981
    emb = UO.timestep_embedding(torch.tensor(g["temb_t"]), 64)
    assert rel_err(emb, g["temb"]) < 1e-6


@pytest.mark.parametrize("fixture,cfgname", [("unet_tiny_b3.npz", "tiny"), ("unet_shapes3d_b2.npz", "shapes3d")])
def test_unet_forward_backward_vs_golden(fixture, cfgname):
    g = load(fixture)
    cfg = CONFIGS[cfgname]
    p = {k: v.requires_grad_(True) for k, v in UO.make_params(cfg, seed=1).items()}
    x = torch.tensor(g["x"]).requires_grad_(True)
    ctx = torch.tensor(g["ctx"]).requires_grad_(True)
    t = torch.tensor(g["t"])
    taps = {}
    eps = UO.unet_forward(p, cfg, x, t, [ctx], taps=taps)
    assert rel_err(eps.detach(), g["eps"]) < 2e-5
    for k, v in taps.items():
        assert rel_err(v.detach()[0, :8, :2, :2], g["tap_corner::" + k]) < 5e-5, k
        assert abs(v.detach().double().abs().mean().item() - float(g["tap_absmean::" + k])) < 1e-5 * (1 + float(g["tap_absmean::" + k]))
    loss = DO.l1_loss(eps, torch.tensor(g["noise"]))
    assert abs(loss.item() - float(g["loss"])) < 1e-6
    loss.backward()
    assert rel_err(x.grad, g["dx"]) < 1e-4
    assert rel_err(ctx.grad, g["dctx"]) < 1e-4
    names = list(g["param_names"])
    assert names == [n for n, _ in UO.param_specs(cfg)]
    norms = np.array([p[n].grad.norm().item() for n in names])
    assert np.allclose(norms, g["grad_norms"], rtol=2e-4, atol=1e-9)
    for k in g.files:
        if k.startswith("grad::"):
            assert rel_err(p[k[6:]].grad, g[k]) < 2e-4, k


@pytest.mark.parametrize("fixture,cfgname", [("ddim_tiny_b2_s10_eta000.npz", "tiny"),
                                             ("ddim_tiny_b2_s10_eta100.npz", "tiny")])
def test_ddim_vs_golden(fixture, cfgname):
    g = load(fixture)
    cfg = CONFIGS[cfgname]
    p = UO.make_params(cfg, seed=1)
    ctx = torch.tensor(g["ctx"])
    acp = DO.alphas_cumprod(DO.linear_betas())
    S, eta = int(g["S"]), float(g["eta"])
    co = DO.ddim_coefficients(acp, S, eta)
    assert np.array_equal(co["timesteps"], g["ddim_timesteps"])
    assert np.array_equal(co["a_t"], g["ddim_alphas"])
    assert np.array_equal(co["sqrt_one_minus_at"], g["ddim_sqrt_one_minus_alphas"])
    assert np.array_equal(co["sigma_t"], g["ddim_sigmas"].astype(np.float32))
    noises = [torch.tensor(n) for n in g["noises"]] if eta > 0 else None
    with torch.no_grad():
        out, inter = DO.ddim_sample(lambda x, ts: UO.unet_forward(p, cfg, x, ts, [ctx]), torch.tensor(g["x_T"]), acp,
                                    S, eta, noises, log_every_t=10)
    assert rel_err(out, g["samples"]) < 1e-4
    assert len(inter["x_inter"]) == g["x_inter"].shape[0]
    assert rel_err(torch.stack(inter["pred_x0"]), g["pred_x0"]) < 1e-4


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree only exists in the build container")
def test_live_reference_submodules():
    """SpatialTransformer / CrossAttention / ResBlock of the reference vs the oracle functions, same weights."""
    attention, openaimodel, util = ref_shim.reference_modules()
    torch.manual_seed(0)
    st = attention.SpatialTransformer(128, 8, 16, depth=1, context_dim=16).eval()
    for prm in st.parameters():
        torch.nn.init.normal_(prm, std=0.05)
    p = {"st." + k: v.detach() for k, v in st.state_dict().items()}
    x, ctx = torch.randn(2, 128, 8, 8), torch.randn(2, 20, 16)
    lay = UO.Layer("st", "st", 128, 128, heads=8, d_head=16)
    with torch.no_grad():
        assert rel_err(UO.spatial_transformer(p, lay, x, ctx), st(x, ctx)) < 2e-5
        ca = st.transformer_blocks[0].attn2
        tok = torch.randn(2, 64, 128)
        assert rel_err(UO.attention(p, "st.transformer_blocks.0.attn2", tok, ctx, 8), ca(tok, ctx)) < 2e-5
    rb = openaimodel.ResBlock(128, 256, 0, out_channels=64, use_scale_shift_norm=True, down=False).eval()
    for prm in rb.parameters():
        torch.nn.init.normal_(prm, std=0.05)
    pr = {"rb." + k: v.detach() for k, v in rb.state_dict().items()}
    emb = torch.randn(2, 256)
    with torch.no_grad():
        assert rel_err(UO.resblock(pr, UO.Layer("res", "rb", 128, 64), x, emb), rb(x, emb)) < 2e-5
