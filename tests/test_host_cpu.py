"""CPU tests of the host logic: DDIM schedule tables of the drop-in sampler against the reference-generated golden
vectors, the beta schedule, LitEma / LR-scaling rules, and the data-parallel plumbing on a world_size-2 gloo group
(batch sharding + ONE flat-bucket gradient all-reduce reproduces the full-batch gradient of the oracle)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import diffusion_oracle as DO
from oracle import unet_oracle as UO
from oracle.make_golden import TINY, seeded_inputs

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def tiny_unet():
    from encdiff_b200 import UNetModel
    cfg = TINY
    return UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                     model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                     num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                     use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                     context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)


def test_beta_schedule_matches_reference():
    from encdiff_b200.train import make_beta_schedule
    g = np.load(os.path.join(GOLDEN, "schedule.npz"))
    assert np.array_equal(make_beta_schedule(), g["betas"])
    with pytest.raises(NotImplementedError):
        make_beta_schedule("cosine")


@pytest.mark.parametrize("S", [50, 200])
def test_sampler_schedule_tables_match_reference(S):
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.diffusion import DenoiserHost
    g = np.load(os.path.join(GOLDEN, "schedule.npz"))
    host = DenoiserHost(tiny_unet())
    s = DDIMSampler(host)
    s.make_schedule(S, ddim_eta=0.0, verbose=False)
    assert np.array_equal(s.ddim_timesteps, g[f"timesteps_{S}"])
    assert np.array_equal(s.ddim_alphas.numpy(), g[f"alphas_{S}"])
    assert np.array_equal(np.asarray(s.ddim_alphas_prev), g[f"alphas_prev_{S}"])
    assert s.ddim_timesteps[0] == 1 and s.ddim_timesteps[-1] == 1000 - 1000 // S + 1


def test_sampler_sigmas_match_reference_eta1():
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.diffusion import DenoiserHost
    g = np.load(os.path.join(GOLDEN, "ddim_tiny_b2_s10_eta100.npz"))
    s = DDIMSampler(DenoiserHost(tiny_unet()))
    s.make_schedule(int(g["S"]), ddim_eta=float(g["eta"]), verbose=False)
    assert np.array_equal(s.ddim_timesteps, g["ddim_timesteps"])
    assert np.array_equal(np.asarray(s.ddim_sigmas), g["ddim_sigmas"])
    assert np.array_equal(s.ddim_sqrt_one_minus_alphas.numpy(), g["ddim_sqrt_one_minus_alphas"])


def test_q_sample_and_apply_model_plumbing():
    from encdiff_b200.diffusion import DenoiserHost
    host = DenoiserHost(tiny_unet())
    x, t, ctx, noise = seeded_inputs(TINY, 3, seed=2)
    acp = DO.alphas_cumprod(DO.linear_betas())
    want = DO.q_sample(x, t, noise, acp)
    assert torch.allclose(host.q_sample(x, t, noise), want, rtol=0, atol=1e-6)
    with pytest.raises(RuntimeError, match="CUDA"):        # reaches the UNet, which refuses CPU tensors
        host.apply_model(x, t, ctx)


def test_ema_warmup_and_lr_scaling_rules():
    from encdiff_b200 import parallel
    # LitEma (ema.py:31-33): decay = min(0.9999, (1 + n) / (10 + n)), n = number of updates so far, counted from 1
    for n, want in [(1, 2 / 11), (10, 11 / 20), (10 ** 7, 0.9999)]:
        assert abs(min(0.9999, (1 + n) / (10 + n)) - want) < 1e-12
    assert parallel.scaled_lr(1e-6, 8, 256) == pytest.approx(8 * 256 * 1e-6)
    assert [parallel.shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [parallel.shard_range(8, r, 2) for r in range(2)] == [(0, 4), (4, 8)]
    with pytest.raises(ValueError):
        parallel.shard_range(4, 2, 2)


# ---- world_size 2, gloo ---------------------------------------------------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _flat_oracle_grad(params, cfg, x, t, ctx, noise):
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    acp = DO.alphas_cumprod(DO.linear_betas())
    loss = DO.l1_loss(UO.unet_forward(p, cfg, DO.q_sample(x, t, noise, acp), t, [ctx]), noise)
    loss.backward()
    return torch.cat([p[n].grad.reshape(-1) for n, _ in UO.param_specs(cfg)]), loss.detach()


def _dp_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    from encdiff_b200 import parallel
    r, _, w = parallel.init_from_env("gloo")
    assert (r, w) == (rank, world)
    cfg = TINY
    params = UO.make_params(cfg, seed=1 + rank)               # replicas start different ...
    flat = torch.cat([params[n].reshape(-1) for n, _ in UO.param_specs(cfg)])
    parallel.broadcast_(flat)                                 # ... and take rank 0's parameters
    off = 0
    for n, s in UO.param_specs(cfg):
        k = int(np.prod(s))
        params[n] = flat[off:off + k].view(*s).clone()
        off += k
    x, t, ctx, noise = seeded_inputs(cfg, 4, seed=9)
    b, e = parallel.shard_range(4, rank, world)
    g, loss = _flat_oracle_grad(params, cfg, x[b:e], t[b:e], ctx[b:e], noise[b:e])
    parallel.allreduce_mean_(g)
    samples = parallel.gather_samples(x[b:e])
    ms = parallel.max_over_ranks(float(rank + 1))
    if rank == 0:
        torch.save({"grad": g, "samples": samples, "ms": ms}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_gloo_world2(tmp_path):
    out = str(tmp_path / "dp.pt")
    mp.spawn(_dp_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    cfg = TINY
    x, t, ctx, noise = seeded_inputs(cfg, 4, seed=9)
    want, _ = _flat_oracle_grad(UO.make_params(cfg, seed=1), cfg, x, t, ctx, noise)
    err = (got["grad"] - want).abs().max() / want.abs().max()
    assert err < 1e-5, err                                    # mean of equal-shard mean-gradients == batch-mean gradient
    assert torch.equal(got["samples"], x)                     # rank-ordered gather
    assert got["ms"] == 2.0                                   # max over ranks
