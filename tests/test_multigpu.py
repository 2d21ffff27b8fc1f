"""The N > 1 paths on real GPUs (skipped on a single-GPU box; run with `gpurun --gpus 2`): batch-sharded data-parallel
training step == the single-GPU step on the whole batch (one NCCL all-reduce of the flat gradient bucket), and the
folded swap generation sharded over two ranks == the unsharded run (no communication inside the loop, one final gather).
The host-side logic of both is also covered on CPU by tests/test_host_cpu.py::test_data_parallel_gloo_world2."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS, seeded_inputs

pytestmark = pytest.mark.gpu


def _need_two():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _model(cfg, dev):
    from encdiff_b200 import UNetModel
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    m.load_state_dict(UO.make_params(cfg, seed=1), strict=True)
    return m.to(dev)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from encdiff_b200 import DDIMSampler, parallel
    from encdiff_b200.diffusion import DenoiserHost
    from encdiff_b200.swap import sample_swap
    from encdiff_b200.train import DenoiserTrainer
    r, local, w = parallel.init_from_env("nccl")
    dev = torch.device("cuda", local)
    cfg = CONFIGS["tiny"]
    # ---- training: each rank steps on its half of the batch ------------------------------------------------------
    B = 8
    x, t, ctx, noise = seeded_inputs(cfg, B, seed=31)
    b, e = parallel.shard_range(B, r, w)
    m = _model(cfg, dev).train()
    tr = DenoiserTrainer(m, lr=1e-4, use_cuda_graph=False)
    loss, _ = tr.step(x[b:e].to(dev), ctx[b:e].to(dev), t[b:e].to(dev), noise[b:e].to(dev))
    grad = (m._gflat / w).cpu()                       # the all-reduced SUM; the optimiser kernel applies 1 / world
    loss_mean = parallel.allreduce_mean_(loss.clone()).cpu()
    # ---- overlapped all-reduce (bucket tail under the deferred weight gradients of the rest) vs one all-reduce after the
    # backward: four graph-mode steps each (eager warm-up, capture, two replays); world 2 => sums are bit-identical -------
    finals = []
    for overlap in (True, False):
        mo = _model(cfg, dev).train()
        tro = DenoiserTrainer(mo, lr=1e-3, overlap_allreduce=overlap)
        assert tro.overlap_allreduce == overlap
        for k in range(4):
            xs, ts, cs, ns = seeded_inputs(cfg, B, seed=200 + k)
            tro.step(xs[b:e].to(dev), cs[b:e].to(dev), ts[b:e].to(dev), ns[b:e].to(dev))
        assert (tro._split > 0) == overlap
        finals.append(tro.opt.flat.clone())
    start = torch.cat([q.detach().flatten() for q in _model(cfg, dev).parameters()])
    # (gradients differ in the last bits from run to run -- float atomics, TMA reduce-add order -- so not torch.equal)
    overlap_equal = bool((finals[0] - finals[1]).norm() <= 2e-2 * (finals[1] - start).norm())
    both_f = [torch.empty_like(finals[0]) for _ in range(w)]
    torch.distributed.all_gather(both_f, finals[0])
    overlap_replicas = bool(torch.equal(both_f[0], both_f[1]))
    # ---- swap generation: same codes on every rank, folded batch sharded ---------------------------------------------
    host = DenoiserHost(_model(cfg, dev).eval()).to(dev)
    L, N, S = cfg.latent_unit, 2, 6
    g = torch.Generator().manual_seed(3)
    tokens = torch.randn(N, L, cfg.context_dim, generator=g).to(dev)
    xT = torch.randn(L, N, cfg.in_channels, cfg.image_size, cfg.image_size, generator=g).to(dev)
    shape = (cfg.in_channels, cfg.image_size, cfg.image_size)
    sharded = sample_swap(DDIMSampler(host), tokens, None, ddim_steps=S, eta=0., shape=shape, x_T=xT,
                          reference_rng=False)
    # ---- torch DistributedDataParallel around the drop-in module (the reference's Lightning 'ddp' strategy, main_val.py:
    # 656-660, find_unused_parameters as Lightning sets it): every parameter's AccumulateGrad hook must fire so the bucket
    # reducer sees all tensors; two iterations, the second one is where a missing hook raises -------------------------
    from torch.nn.parallel import DistributedDataParallel as DDP
    md = _model(cfg, dev).train()
    ddp = DDP(md, device_ids=[local], find_unused_parameters=True)
    gw = torch.Generator().manual_seed(77)
    wgt = torch.randn(x.shape, generator=gw)
    for _ in range(2):
        ddp.zero_grad(set_to_none=True)
        out_d = ddp(x[b:e].to(dev), t[b:e].to(dev), context=[ctx[b:e].to(dev)])
        (out_d * wgt[b:e].to(dev)).sum().backward()
    g_ddp = torch.cat([q.grad.flatten() for q in md.parameters()])
    both_g = [torch.empty_like(g_ddp) for _ in range(w)]
    torch.distributed.all_gather(both_g, g_ddp)
    ddp_equal = bool(torch.equal(both_g[0], both_g[1]))
    g_ddp = g_ddp.cpu()
    del ddp
    # ---- the full step from images (EncDiffTrainer): different initial weights and different images on each rank; the
    # constructor's broadcast and the two gradient all-reduces must leave the replicas bit-identical ----------------------
    from encdiff_b200 import Encoder4, VQEncoder
    from encdiff_b200.train import EncDiffTrainer
    cfg_s = CONFIGS["shapes3d"]
    torch.manual_seed(7)
    vq = VQEncoder(embed_dim=3, ddconfig=dict(double_z=False, z_channels=3, resolution=64, in_channels=3, out_ch=3, ch=32,
                                              ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[], dropout=0.0)).to(dev)
    torch.manual_seed(100 + r)
    enc = Encoder4(d=128, context_dim=cfg_s.context_dim, latent_unit=cfg_s.latent_unit).to(dev).train()
    unet = _model(cfg_s, dev).train()
    with torch.no_grad():
        for q in unet.parameters():
            q.add_(0.01 * r)
    full = EncDiffTrainer(vq, enc, unet, lr=1e-3, use_cuda_graph=False)
    start = full.cond_opt.flat.clone()
    gi = torch.Generator().manual_seed(500 + r)
    for _ in range(2):
        full.step((torch.rand(4, 3, 64, 64, generator=gi) * 2 - 1).to(dev))
    mine = torch.cat([full.cond_opt.flat, full.denoiser.opt.flat])
    both = [torch.empty_like(mine) for _ in range(w)]
    torch.distributed.all_gather(both, mine)
    replicas_equal = bool(torch.equal(both[0], both[1]))
    moved = float((full.cond_opt.flat - start).abs().max())
    if r == 0:
        torch.save({"grad": grad, "loss": loss_mean, "swap": sharded.cpu(), "replicas_equal": replicas_equal,
                    "moved": moved, "ddp_grad": g_ddp, "ddp_equal": ddp_equal, "overlap_equal": overlap_equal,
                    "overlap_replicas": overlap_replicas}, out)
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_dp_step_and_sharded_swap_nccl_world2(cuda_device, tmp_path):
    _need_two()
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.diffusion import DenoiserHost
    from encdiff_b200.swap import sample_swap
    from encdiff_b200.train import DenoiserTrainer
    out = str(tmp_path / "mg.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    cfg = CONFIGS["tiny"]
    dev = cuda_device
    x, t, ctx, noise = seeded_inputs(cfg, 8, seed=31)
    m = _model(cfg, dev).train()
    tr = DenoiserTrainer(m, lr=1e-4, use_cuda_graph=False)
    loss, _ = tr.step(x.to(dev), ctx.to(dev), t.to(dev), noise.to(dev))
    want = m._gflat.cpu()
    err = ((got["grad"] - want).abs().max() / want.abs().max()).item()
    print(f"dp2 vs single-GPU gradient: max rel err {err:.2e}; loss {got['loss'].item():.6f} vs {loss.item():.6f}")
    assert err < 2e-3                                  # per-rank gradient range scales differ: 16-bit rounding only
    assert abs(got["loss"].item() - loss.item()) < 1e-5
    host = DenoiserHost(_model(cfg, dev).eval()).to(dev)
    L, N, S = cfg.latent_unit, 2, 6
    g = torch.Generator().manual_seed(3)
    tokens = torch.randn(N, L, cfg.context_dim, generator=g).to(dev)
    xT = torch.randn(L, N, cfg.in_channels, cfg.image_size, cfg.image_size, generator=g).to(dev)
    whole = sample_swap(DDIMSampler(host), tokens, None, ddim_steps=S, eta=0.,
                        shape=(cfg.in_channels, cfg.image_size, cfg.image_size), x_T=xT, reference_rng=False)
    assert torch.equal(got["swap"], whole.cpu())       # a sample does not depend on its batch neighbours: bit-identical
    print(f"overlapped all-reduce: parameter update over 4 steps within 2 % (L2) of the un-overlapped run's: {got['overlap_equal']}; replicas identical: "
          f"{got['overlap_replicas']}")
    assert got["overlap_equal"] and got["overlap_replicas"]
    # DDP averages the ranks' gradients: 2 x that == the gradient of the whole-batch sum loss on one GPU
    md = _model(cfg, dev).train()
    gw = torch.Generator().manual_seed(77)
    wgt = torch.randn(x.shape, generator=gw).to(dev)
    (md(x.to(dev), t.to(dev), context=[ctx.to(dev)]) * wgt).sum().backward()
    want_d = torch.cat([q.grad.flatten() for q in md.parameters()]).cpu()
    err_d = ((2 * got["ddp_grad"] - want_d).abs().max() / want_d.abs().max()).item()
    print(f"torch DDP around UNetModel: ranks hold identical gradients: {got['ddp_equal']}; vs single-GPU whole batch {err_d:.2e}")
    assert got["ddp_equal"] and err_d < 5e-3
    print(f"EncDiffTrainer dp2: replicas identical after 2 steps: {got['replicas_equal']}; concept encoder moved {got['moved']:.2e}")
    assert got["replicas_equal"] and got["moved"] > 0
