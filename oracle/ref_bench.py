"""Timing harness for the UNMODIFIED reference on the same hot path bench.py measures (SURVEY.md
section 8d): ``LatentDiffusion.p_losses`` (ddpm_enc.py:1183-1213) -> ``apply_model`` -> ``DiffusionWrapper`` -> the
reference ``UNetModel``, ``loss.backward()``, ``torch.optim.AdamW`` over ``model.model.parameters()``
(ddpm_enc.py:1598-1604) and ``LitEma`` (ddpm_enc.py:399-401), on latents / concept tokens of the benched shape; and
``DDIMSampler.sample`` (ddim.py:56) for the sampling leg.  Runs on the host cores (the CPU baseline / reference arm) or
in eager mode on a GPU (the honest GPU comparator).  Imported ONLY by bench.py's baseline legs and tests; the reference
package comes from /root/reference or oracle/_ref through oracle/ref_shim.py.  Baseline infrastructure, not product."""
from __future__ import annotations

import contextlib
import os
import statistics
import time

import torch

from oracle import ref_shim
from oracle import unet_oracle as UO


def available() -> bool:
    return ref_shim.available()


def build(device="cpu", seed=23):
    """The real LatentDiffusion from the Shapes3D yaml values, reference UNet, zero-initialised tensors re-drawn
     so that the step does real work."""
    cfg = UO.SHAPES3D
    with contextlib.redirect_stdout(open(os.devnull, "w")):
        model = ref_shim.reference_latent_diffusion(cfg, device=device, seed=seed,
                                                    unet_params=UO.make_params(cfg, seed=1))
    return model.train()


def _batch(B, seed, device):
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(B, 3, 16, 16, generator=g)
    t = torch.randint(0, 1000, (B,), generator=g)
    cond = torch.randn(B, 320, generator=g)
    noise = torch.randn(B, 3, 16, 16, generator=g)
    return [v.to(device) for v in (x0, cond, t, noise)]


def train_step_time(B, steps, warmup, device="cpu", autocast_dtype=None, tf32=False, model=None):
    """Median seconds per reference training step (p_losses fwd + bwd + AdamW + LitEma) at batch B."""
    dev = torch.device(device)
    if dev.type == "cpu":
        torch.set_num_threads(os.cpu_count() or 1)
    else:
        torch.backends.cuda.matmul.allow_tf32 = tf32           # main_val.py:524-525 sets 'high' (= TF32) on GPUs
        torch.backends.cudnn.allow_tf32 = tf32
    model = build(dev) if model is None else model
    opt = torch.optim.AdamW(list(model.model.parameters()), lr=1e-4)
    x0, cond, t, noise = _batch(B, 23, dev)
    times = []
    for i in range(warmup + steps):
        if dev.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        ctx = torch.autocast("cuda", dtype=autocast_dtype) if autocast_dtype is not None else contextlib.nullcontext()
        with ctx:
            loss, _ = model.p_losses(x0, cond, t, noise=noise)
            if autocast_dtype is not None:
                # the reference's CheckpointFunction.backward (diffusionmodules/util.py:147-176) re-runs the block forward
                # WITHOUT restoring the autocast state, so under autocast its recompute mixes bf16 saved inputs with fp32
                # weights and raises; keeping backward() inside the autocast region is the only way the unmodified
                # reference trains in bf16 at all
                loss.backward()
        if autocast_dtype is None:
            loss.backward()
        opt.step()
        if model.use_ema:
            model.model_ema(model.model)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return statistics.median(times), (torch.get_num_threads() if dev.type == "cpu" else 0), float(loss.detach())


def ddim_time(B, S, reps, device, tf32=False, autocast_dtype=None, model=None):
    """Seconds per DDIMSampler.sample(S, B, ...) call of the reference sampler around the reference UNet (eta 0)."""
    dev = torch.device(device)
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    model = build(dev) if model is None else model
    model.eval()
    g = torch.Generator().manual_seed(5)
    cond = torch.randn(B, 320, generator=g).to(dev)
    xT = torch.randn(B, 3, 16, 16, generator=g).to(dev)
    times = []
    for i in range(1 + reps):
        if dev.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx = torch.autocast("cuda", dtype=autocast_dtype) if autocast_dtype is not None else contextlib.nullcontext()
        with ctx, contextlib.redirect_stdout(open(os.devnull, "w")):
            model.sample_log(cond=cond, batch_size=B, ddim=True, ddim_steps=S, eta=0., x_T=xT)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        if i >= 1:
            times.append(time.perf_counter() - t0)
    model.train()
    return statistics.median(times)


def full_step_time(B, steps, warmup, device="cpu", tf32=False, model=None):
    """configs[0] as defined (SURVEY.md 8d "Config 1"): the COMPLETE reference training step from images --
    ``LatentDiffusion.shared_step`` (ddpm_enc.py:1018-1021) = get_input (VQ-f4 encode of the images, frozen) + forward
    (Encoder4 on the images, random t, p_losses through the UNet) -- then ``loss.backward()``, AdamW over the UNet and the
    trainable concept encoder (configure_optimizers, ddpm_enc.py:1598-1612) and LitEma.  Median seconds per step."""
    dev = torch.device(device)
    if dev.type == "cpu":
        torch.set_num_threads(os.cpu_count() or 1)
    else:
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
    model = build(dev) if model is None else model
    params = list(model.model.parameters()) + list(model.cond_stage_model.parameters())
    opt = torch.optim.AdamW(params, lr=1e-4)
    g = torch.Generator().manual_seed(23)
    batch = {"image": (torch.rand(B, 64, 64, 3, generator=g) * 2 - 1).to(dev)}
    times = []
    for i in range(warmup + steps):
        if dev.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        out = model.shared_step(batch)
        loss = out[0] if isinstance(out, tuple) else out
        loss = loss[0] if isinstance(loss, tuple) else loss
        loss.backward()
        opt.step()
        if model.use_ema:
            model.model_ema(model.model)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return statistics.median(times), (torch.get_num_threads() if dev.type == "cpu" else 0), float(loss.detach())
