"""Generate tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE (imported from /root/reference) on seeded inputs.
Run in the build container only:  python -m oracle.make_golden
The fixtures pin the oracle (tests/test_oracle.py) and, through it and directly, the CUDA path (tests -m gpu).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import diffusion_oracle as DO  # noqa: E402
from oracle import ref_shim  # noqa: E402
from oracle import unet_oracle as UO  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

# small architecture with the same block types (down/up ResBlocks, skip-conv ResBlocks, ST at two resolutions)
TINY = UO.UNetConfig(image_size=8, model_channels=64, channel_mult=(1, 2), num_res_blocks=1,
                     attention_resolutions=(1, 2), latent_unit=20)
# BASELINE.json configs[4] (CelebA-64, KL-f4 first stage => 16x16x3 latent, larger UNet): the reference tree ships no yaml
# for it, so the configuration is synthesised as SURVEY.md suggests: twice the Shapes3D width
CELEBA64 = UO.UNetConfig(image_size=16, model_channels=128, channel_mult=(1, 2, 4, 4), num_res_blocks=2,
                         attention_resolutions=(1, 2, 4), latent_unit=20)
CONFIGS = {"shapes3d": UO.SHAPES3D, "tiny": TINY, "celeba64": CELEBA64}


def seeded_inputs(cfg, B, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cfg.in_channels, cfg.image_size, cfg.image_size, generator=g)
    t = torch.randint(0, 1000, (B,), generator=g)
    ctx = torch.randn(B, cfg.latent_unit * cfg.context_dim, generator=g)
    noise = torch.randn(B, cfg.in_channels, cfg.image_size, cfg.image_size, generator=g)
    return x, t, ctx, noise


def unet_case(name, B, seed):
    cfg = CONFIGS[name]
    params = UO.make_params(cfg, seed=1)
    ref = ref_shim.reference_unet(cfg, params)
    for p in ref.parameters():
        p.requires_grad_(True)
    x, t, ctx, noise = seeded_inputs(cfg, B, seed)
    x = x.requires_grad_(True)
    ctx = ctx.requires_grad_(True)
    taps = {}
    hooks = []
    for i, m in enumerate(ref.input_blocks):
        hooks.append(m.register_forward_hook(lambda mod, inp, out, k=f"input_blocks.{i}": taps.__setitem__(k, out.detach())))
    hooks.append(ref.middle_block.register_forward_hook(lambda mod, inp, out: taps.__setitem__("middle_block", out.detach())))
    for i, m in enumerate(ref.output_blocks):
        hooks.append(m.register_forward_hook(lambda mod, inp, out, k=f"output_blocks.{i}": taps.__setitem__(k, out.detach())))
    eps = ref(x, t, context=[ctx])
    loss = (noise - eps).abs().mean(dim=[1, 2, 3]).mean()   # ddpm_enc.py:297-301,1198-1213
    loss.backward()
    for h in hooks:
        h.remove()
    out = {
        "x": x.detach().numpy(), "t": t.numpy(), "ctx": ctx.detach().numpy(), "noise": noise.numpy(),
        "eps": eps.detach().numpy(), "loss": np.float32(loss.item()),
        "dx": x.grad.numpy(), "dctx": ctx.grad.numpy(),
    }
    names = [n for n, _ in ref.named_parameters()]
    out["param_names"] = np.array(names)
    out["grad_norms"] = np.array([p.grad.norm().item() for _, p in ref.named_parameters()], dtype=np.float64)
    out["grad_sums"] = np.array([p.grad.double().sum().item() for _, p in ref.named_parameters()], dtype=np.float64)
    keep = ["time_embed.0.bias", "input_blocks.0.0.weight", "input_blocks.1.0.in_layers.0.weight",
            "input_blocks.1.0.emb_layers.1.bias", "input_blocks.1.1.transformer_blocks.0.attn2.to_k.weight",
            "input_blocks.1.1.transformer_blocks.0.norm2.bias", "out.2.weight", "out.0.weight"]
    pd = dict(ref.named_parameters())
    for k in keep:
        out["grad::" + k] = pd[k].grad.numpy()
    for k, v in taps.items():
        out["tap_mean::" + k] = np.float64(v.double().mean().item())
        out["tap_absmean::" + k] = np.float64(v.double().abs().mean().item())
        out["tap_corner::" + k] = v[0, :8, :2, :2].numpy()
    np.savez_compressed(os.path.join(GOLDEN, f"unet_{name}_b{B}.npz"), **out)
    print(f"unet_{name}_b{B}: loss {loss.item():.6f} eps absmax {eps.abs().max().item():.4f}")


def init_case():
    """Reference initialisation anchors (SURVEY.md §8c pins i, ii, v)."""
    torch.manual_seed(23)
    ref = ref_shim.reference_unet(UO.SHAPES3D)
    n_t = sum(1 for _ in ref.parameters())
    n_p = sum(p.numel() for p in ref.parameters())
    x, t, ctx, noise = seeded_inputs(UO.SHAPES3D, 8, 7)
    with torch.no_grad():
        eps = ref(x, t, context=[ctx])
    loss = (noise - eps).abs().mean().item()
    names = np.array([n for n, _ in ref.named_parameters()])
    shapes = np.array([str(tuple(p.shape)) for _, p in ref.named_parameters()])
    np.savez_compressed(os.path.join(GOLDEN, "init_anchor.npz"), n_tensors=n_t, n_params=n_p,
                        eps_absmax=np.float32(eps.abs().max().item()), loss=np.float32(loss), names=names,
                        shapes=shapes)
    print("init:", n_t, n_p, eps.abs().max().item(), loss)


def ddim_case(name, B, S, eta, seed):
    cfg = CONFIGS[name]
    params = UO.make_params(cfg, seed=1)
    ref = ref_shim.reference_unet(cfg, params)
    betas = DO.linear_betas()
    model = ref_shim.FakeLatentDiffusion(ref, betas)
    sampler = ref_shim.reference_ddim_sampler(model)
    x, t, ctx, noise = seeded_inputs(cfg, B, seed)
    x_T = x
    shape = (cfg.in_channels, cfg.image_size, cfg.image_size)
    torch.manual_seed(1234)
    samples, inter = sampler.sample(S, B, shape, conditioning=ctx, eta=eta, x_T=x_T, verbose=False, log_every_t=10)
    # the per-step noise the sampler drew (noise_like is the only RNG consumer once x_T is explicit, ddim.py:203)
    torch.manual_seed(1234)
    noises = np.stack([torch.randn(B, *shape).numpy() for _ in range(S)])
    out = {
        "x_T": x_T.numpy(), "ctx": ctx.numpy(), "samples": samples.numpy(),
        "x_inter": np.stack([v.numpy() for v in inter["x_inter"]]),
        "pred_x0": np.stack([v.numpy() for v in inter["pred_x0"]]),
        "ddim_timesteps": np.asarray(sampler.ddim_timesteps),
        "ddim_alphas": sampler.ddim_alphas.numpy(), "ddim_alphas_prev": np.asarray(sampler.ddim_alphas_prev),
        "ddim_sigmas": np.asarray(sampler.ddim_sigmas),
        "ddim_sqrt_one_minus_alphas": sampler.ddim_sqrt_one_minus_alphas.numpy(),
        "S": S, "eta": eta,
    }
    if eta > 0:
        out["noises"] = noises
    np.savez_compressed(os.path.join(GOLDEN, f"ddim_{name}_b{B}_s{S}_eta{int(eta * 100):03d}.npz"), **out)
    print(f"ddim_{name}_b{B}_s{S}_eta{eta}: samples absmax {samples.abs().max().item():.4f}")


def schedule_case():
    betas = DO.linear_betas()
    model = ref_shim.FakeLatentDiffusion(None, betas)
    out = {}
    for S in (50, 200):
        s = ref_shim.reference_ddim_sampler(model)
        s.make_schedule(S, ddim_eta=0.0, verbose=False)
        out[f"timesteps_{S}"] = np.asarray(s.ddim_timesteps)
        out[f"alphas_{S}"] = s.ddim_alphas.numpy()
        out[f"alphas_prev_{S}"] = np.asarray(s.ddim_alphas_prev)
    ref_shim._install_stubs()
    from ldm.modules.diffusionmodules.util import make_beta_schedule, timestep_embedding
    out["betas"] = make_beta_schedule("linear", 1000, linear_start=0.0015, linear_end=0.0155)
    out["temb_t"] = np.array([0, 1, 17, 500, 999], dtype=np.int64)
    out["temb"] = timestep_embedding(torch.tensor(out["temb_t"]), 64).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "schedule.npz"), **out)
    print("schedule ok")


def main():
    assert ref_shim.available(), "reference tree not found"
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    schedule_case()
    init_case()
    unet_case("tiny", 3, seed=11)
    unet_case("shapes3d", 2, seed=23)
    ddim_case("tiny", 2, 10, 0.0, seed=5)
    ddim_case("tiny", 2, 10, 1.0, seed=6)
    ddim_case("shapes3d", 2, 50, 0.0, seed=23)


if __name__ == "__main__":
    main()
