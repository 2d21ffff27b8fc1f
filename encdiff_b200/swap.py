"""Concept-token swap generation (BASELINE.json configs[3]): the ``sample_swap`` branch of
``LatentDiffusion.log_images`` (ldm/models/diffusion/ddpm_enc.py:1522-1535, driven by generate_swap_from_ckpt.py:103-105
with DDIM-200, eta 0).

The reference runs ``latent_unit`` (20) independent DDIM samplings of N images one after the other: in run ``cdx`` every
image keeps its own concept code except unit ``cdx``, which is replaced by image 0's.  The runs share nothing, so here
they are folded into ONE batch of ``latent_unit * N`` samples (run-major, i.e. exactly the order of the reference's final
``torch.cat(x_samples_list, dim=0)``) and sampled by one replayed step graph; across GPUs the folded batch is sharded
with no communication inside the loop.

The concept encoder (``cond_stage_model.warp``: 20 per-unit MLPs, openaimodel_enc.py:1037-1041) and the first-stage
decoder stay the reference's frozen PyTorch front / back ends: ``warp`` is passed in as a callable and the result is
returned in latent space.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

from . import parallel


def swap_codes(orc: torch.Tensor, latent_unit: int) -> torch.Tensor:
    """All swapped concept codes of ddpm_enc.py:1525-1529 at once.

    ``orc``: the encoder's codes, [N, latent_unit * d] (chunked along dim 1 as the reference does) or [N, latent_unit, d].
    Returns [latent_unit, N, latent_unit, d]: out[cdx] is the reference's ``swapped_c`` of run ``cdx`` (a copy of ``orc``
    whose unit ``cdx`` is image 0's for every image)."""
    if orc.dim() == 2:
        if orc.shape[1] % latent_unit != 0:
            raise ValueError(f"swap_codes: {orc.shape[1]} code values do not split into {latent_unit} units")
        orc = torch.stack(orc.chunk(latent_unit, dim=1), dim=1)
    if orc.dim() != 3 or orc.shape[1] != latent_unit:
        raise ValueError(f"swap_codes: expected [N, {latent_unit}, d] codes, got {tuple(orc.shape)}")
    N = orc.shape[0]
    out = orc.unsqueeze(0).repeat(latent_unit, 1, 1, 1)             # [L, N, L, d]
    idx = torch.arange(latent_unit, device=orc.device)
    out[idx, :, idx] = orc[0].unsqueeze(1).expand(latent_unit, N, orc.shape[2])
    return out


def reference_order_noise(runs: int, shape: Tuple[int, ...], steps: int, eta: float, device,
                          x_T: Optional[torch.Tensor] = None):
    """x_T and per-step noise of ``runs`` consecutive reference samplings, drawn from the current torch generator in the
    reference's own order (per run: one randn for x_T unless given, ddim.py:122-125, then one randn per step even at
    eta = 0, ddim.py:203), so a seeded folded run reproduces a seeded reference loop.  Returns (x_T [runs, *shape],
    step_noise [steps, runs, *shape] or None when eta == 0)."""
    xs, noise = [], []
    for r in range(runs):
        xs.append(torch.randn(shape, device=device) if x_T is None else x_T[r])
        per_step = [torch.randn(shape, device=device) for _ in range(steps)]
        if eta != 0.:
            noise.append(torch.stack(per_step, 0))
    xT = torch.stack(xs, 0)
    return xT, (torch.stack(noise, 1) if eta != 0. else None)


@torch.no_grad()
def sample_swap(sampler, orc: torch.Tensor, warp: Optional[Callable[[torch.Tensor], torch.Tensor]] = None,
                ddim_steps: int = 200, eta: float = 0., shape: Tuple[int, int, int] = (3, 16, 16),
                x_T: Optional[torch.Tensor] = None, group=None, gather: bool = True, shard: bool = True,
                reference_rng: bool = True, **sample_kwargs):
    """The whole swap loop of ``log_images(sample_swap=True)`` as one folded, batch-sharded DDIM run.

    sampler:  an ``encdiff_b200.DDIMSampler`` (its model supplies ``apply_model`` and the UNet's ``latent_unit``)
    orc:      [N, latent_unit * d] or [N, latent_unit, d] concept codes of the N source images (``cond_stage_model.encoding``)
    warp:     the frozen ``cond_stage_model.warp`` (codes [M, latent_unit, d] -> tokens [M, latent_unit * context_dim]);
              None when ``orc`` already holds context_dim-wide tokens
    x_T:      optional [latent_unit, N, C, H, W] start noise (run-major); drawn in the reference's order otherwise
    reference_rng: draw x_T / step noise call by call in the reference loop's order (a seeded run then reproduces a
              seeded reference loop; 20 x (1 + S) small randn launches); False = two bulk draws
    shard:    under torch.distributed, split the folded batch over the ranks of ``group`` (every rank passes the same
              ``orc``); False = this rank samples all of its own ``orc`` (each rank has its own source images)
    Returns the latents [latent_unit * N, C, H, W] in the reference's concatenation order (run ``cdx`` = rows
    [cdx * N, (cdx + 1) * N)); with ``gather=False`` under torch.distributed, only this rank's shard and its row range."""
    unet = sampler.model.model.diffusion_model
    L = int(unet.latent_unit)
    codes = swap_codes(orc, L)                                       # [L, N, L, d]
    N = codes.shape[1]
    folded = codes.reshape(L * N, L, codes.shape[3])
    cond = (warp(folded) if warp is not None else folded).reshape(L * N, -1)
    dev = cond.device
    if reference_rng:
        xT, step_noise = reference_order_noise(L, (N,) + tuple(shape), ddim_steps, eta, dev, x_T)
    else:
        xT = torch.randn((L, N) + tuple(shape), device=dev) if x_T is None else x_T
        step_noise = torch.randn((ddim_steps, L, N) + tuple(shape), device=dev) if eta != 0. else None
    xT = xT.reshape((L * N,) + tuple(shape))
    if step_noise is not None:
        step_noise = step_noise.reshape((ddim_steps, L * N) + tuple(shape))
    rank, world = 0, 1
    if shard and torch.distributed.is_available() and torch.distributed.is_initialized():
        rank, world = torch.distributed.get_rank(group), torch.distributed.get_world_size(group)
    b, e = parallel.shard_range(L * N, rank, world)
    if world > 1 and gather and (L * N) % world != 0:
        raise ValueError("sample_swap: gather=True needs latent_unit * N divisible by the world size")
    keep = sampler.consume_rng_at_eta0
    sampler.consume_rng_at_eta0 = False                              # the draws above already followed the reference
    try:
        kw = dict(sample_kwargs)
        if step_noise is not None:
            kw["step_noise"] = step_noise[:, b:e].contiguous()
        samples, _ = sampler.sample(ddim_steps, e - b, tuple(shape), cond[b:e].contiguous(), verbose=False, eta=eta,
                                    x_T=xT[b:e].contiguous(), **kw)
    finally:
        sampler.consume_rng_at_eta0 = keep
    if world > 1 and gather:
        return parallel.gather_samples(samples, group)
    return samples if world == 1 else (samples, (b, e))
