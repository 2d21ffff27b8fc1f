"""Concept-token swap generation (BASELINE.json configs[3]; reference: the sample_swap branch of LatentDiffusion.log_images,
ldm/models/diffusion/ddpm_enc.py:1522-1535).  CPU: the folded conditioning batch against the literal restatement of the
reference loop (and, in the build container, through the reference's own Encoder4.warp); the reference-order noise draws.
GPU: one folded DDIM run == the reference's latent_unit sequential runs on the same seed."""
import os
import sys

import pytest
import torch

from oracle import diffusion_oracle as DO
from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS

REF = "/root/reference"


def _toy_warp(L, d_ctx, seed=0):
    """A per-unit map like cond_stage_model.warp (openaimodel_enc.py:1037-1041): unit i has its own weights."""
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(L, 1, d_ctx, generator=g)
    b = torch.randn(L, d_ctx, generator=g)

    def warp(x):                                   # x [M, L, 1] -> [M, L * d_ctx]
        return torch.cat([torch.tanh(x[:, i] @ w[i] + b[i]) for i in range(L)], dim=1)
    return warp


@pytest.mark.parametrize("three_d", [False, True])
def test_swap_codes_match_reference_loop(three_d):
    from encdiff_b200.swap import swap_codes
    L, N = 20, 5
    orc = torch.randn(N, L, generator=torch.Generator().manual_seed(3))
    warp = _toy_warp(L, 16)
    want = DO.swap_conditionings(orc.reshape(N, L, 1) if three_d else orc, L, warp)     # list of [N, 320], one per cdx
    codes = swap_codes(orc.reshape(N, L, 1) if three_d else orc, L)
    assert codes.shape == (L, N, L, 1)
    got = warp(codes.reshape(L * N, L, 1)).reshape(L, N, -1)
    for cdx in range(L):
        assert torch.equal(got[cdx], want[cdx])
        assert torch.equal(codes[cdx, :, cdx], orc[0, cdx].expand(N).reshape(N, 1))      # unit cdx is image 0's everywhere
    assert torch.equal(codes[3, :, 4], orc[:, 4].reshape(N, 1))                          # the other units are untouched


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout only exists in the build container")
def test_swap_fold_through_reference_encoder4_warp():
    """Folding the 20 runs into the batch is valid for the reference's own warp (no batch statistics inside)."""
    from encdiff_b200.swap import swap_codes
    from oracle import ref_shim
    _, openaimodel_enc, _ = ref_shim.reference_modules()
    Encoder4 = openaimodel_enc.Encoder4
    torch.manual_seed(0)
    enc = Encoder4(d=128, context_dim=16, latent_unit=20).eval()
    N = 4
    orc = torch.randn(N, 20)
    with torch.no_grad():
        want = DO.swap_conditionings(orc, 20, enc.warp)
        got = enc.warp(swap_codes(orc, 20).reshape(20 * N, 20, 1)).reshape(20, N, -1)
    for cdx in range(20):
        assert torch.allclose(got[cdx], want[cdx], rtol=0, atol=1e-6)


def test_reference_order_noise_follows_the_reference_draws():
    from encdiff_b200.swap import reference_order_noise
    shape, runs, S = (2, 3, 4, 4), 3, 5
    for eta in (0.0, 1.0):
        torch.manual_seed(11)
        xT, noise = reference_order_noise(runs, shape, S, eta, "cpu")
        torch.manual_seed(11)
        for r in range(runs):                       # ddim.py:122-125 then one noise_like per step (ddim.py:203)
            assert torch.equal(xT[r], torch.randn(shape))
            for s in range(S):
                n = torch.randn(shape)
                if eta:
                    assert torch.equal(noise[s, r], n)
        assert (noise is None) == (eta == 0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("eta", [0.0, 1.0])
def test_sample_swap_folded_equals_sequential_runs(cuda_device, eta):
    from encdiff_b200 import DDIMSampler, UNetModel
    from encdiff_b200.diffusion import DenoiserHost
    from encdiff_b200.swap import sample_swap
    cfg = CONFIGS["tiny"]
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    m.load_state_dict(UO.make_params(cfg, seed=1), strict=True)
    host = DenoiserHost(m.eval()).to(cuda_device)
    L, N, S = cfg.latent_unit, 3, 8
    shape = (cfg.in_channels, cfg.image_size, cfg.image_size)
    orc = torch.randn(N, L, generator=torch.Generator().manual_seed(5)).to(cuda_device)
    tw = _toy_warp(L, cfg.context_dim)
    warp = lambda x: tw(x.cpu()).to(cuda_device)
    # the reference's loop: one sampler call per swapped unit, all drawing from the same generator
    torch.manual_seed(77)
    seq = []
    for cond in DO.swap_conditionings(orc, L, warp):
        s, _ = DDIMSampler(host).sample(S, N, shape, cond, verbose=False, eta=eta)
        seq.append(s)
    want = torch.cat(seq, 0)
    torch.manual_seed(77)
    got = sample_swap(DDIMSampler(host), orc, warp, ddim_steps=S, eta=eta, shape=shape)
    assert got.shape == want.shape == (L * N,) + shape
    err = (got - want).abs().max().item()
    print(f"eta={eta}: folded vs sequential max abs diff {err:.2e}")
    assert torch.allclose(got, want, rtol=0, atol=2e-3)
