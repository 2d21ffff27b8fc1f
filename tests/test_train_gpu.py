"""Training hot path on the GPU (LatentDiffusion.p_losses forward + backward, AdamW + LitEma) against the CPU oracle:
loss value, the fused graph-captured step, the optimiser kernel, the EMA rule and a short loss curve (north star: loss
curve within 2 % of the reference; here on the tiny configuration so that the fp32 oracle finishes in seconds)."""
import numpy as np
import pytest
import torch

from oracle import diffusion_oracle as DO
from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS, seeded_inputs

pytestmark = pytest.mark.gpu


def make_unet(cfg, dev, seed=1):
    from encdiff_b200 import UNetModel
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    params = UO.make_params(cfg, seed=seed)
    m.load_state_dict(params, strict=True)
    return m.to(dev).train(), params


def test_p_losses_matches_oracle(cuda_device):
    from encdiff_b200.train import DenoiserTrainer
    cfg = CONFIGS["tiny"]
    m, params = make_unet(cfg, cuda_device)
    tr = DenoiserTrainer(m, lr=1e-4, use_cuda_graph=False)
    x, t, ctx, noise = seeded_inputs(cfg, 6, seed=21)
    acp = DO.alphas_cumprod(DO.linear_betas())
    want = DO.l1_loss(UO.unet_forward(params, cfg, DO.q_sample(x, t, noise, acp), t, [ctx]), noise)
    cd = ctx.to(cuda_device).requires_grad_(True)
    loss, ld = tr.p_losses(x.to(cuda_device), cd, t.to(cuda_device), noise.to(cuda_device))
    assert abs(loss.item() - want.item()) <= 2e-3 * want.item()
    assert set(ld) == {"train/loss_simple", "train/loss"}
    loss.backward()
    assert cd.grad is not None and all(p.grad is not None for p in m.parameters())


def test_adamw_kernel_matches_torch(cuda_device):
    """encdiff_adamw_step_dev (+EMA) against torch.optim.AdamW and the LitEma rule, 5 steps on one flat tensor."""
    from encdiff_b200.train import FlatAdamWEMA
    gen = torch.Generator().manual_seed(4)
    w0 = torch.randn(10007, generator=gen)
    ref = torch.nn.Parameter(w0.clone())
    opt_ref = torch.optim.AdamW([ref], lr=3e-3, weight_decay=0.01)
    prm = torch.nn.Parameter(w0.clone().to(cuda_device))
    opt = FlatAdamWEMA([prm], lr=3e-3, weight_decay=0.01, use_ema=True)
    shadow = w0.clone()
    for step in range(1, 6):
        g = torch.randn(10007, generator=gen)
        ref.grad = g.clone()
        opt_ref.step()
        opt.write_hyper()
        opt.launch(g.to(cuda_device) * 2.0, grad_scale=0.5)        # the 1/world factor is applied inside the kernel
        decay = min(0.9999, (1 + step) / (10 + step))              # LitEma.forward, ema.py:31-42
        shadow = shadow - (1.0 - decay) * (shadow - ref.detach())
    torch.cuda.synchronize()
    assert torch.allclose(prm.detach().cpu(), ref.detach(), rtol=1e-5, atol=1e-6)
    assert torch.allclose(opt.ema.cpu(), shadow, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("graph,steps", [(False, 24), (True, 120)])
def test_fused_step_loss_curve_vs_oracle(cuda_device, graph, steps):
    """North-star bar: training loss curve within 2 % of the reference's (here every step of a 120-step run of the tiny
    configuration, same batches / noise / timesteps on both sides; the fp32 oracle needs ~0.2 s per step on the host)."""
    from encdiff_b200.train import DenoiserTrainer
    cfg = CONFIGS["tiny"]
    B, lr = 8, 2e-4
    m, params = make_unet(cfg, cuda_device)
    tr = DenoiserTrainer(m, lr=lr, weight_decay=0.01, use_ema=True, use_cuda_graph=graph)
    # oracle: same batches, torch autograd + torch.optim.AdamW on CPU
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    opt = torch.optim.AdamW([p[n] for n, _ in UO.param_specs(cfg)], lr=lr, weight_decay=0.01)
    acp = DO.alphas_cumprod(DO.linear_betas())
    got, want = [], []
    for i in range(steps):
        x, t, ctx, noise = seeded_inputs(cfg, B, seed=1000 + i)
        opt.zero_grad(set_to_none=True)
        lo = DO.l1_loss(UO.unet_forward(p, cfg, DO.q_sample(x, t, noise, acp), t, [ctx]), noise)
        lo.backward()
        opt.step()
        want.append(lo.item())
        loss, dcond = tr.step(x.to(cuda_device), ctx.to(cuda_device), t.to(cuda_device), noise.to(cuda_device))
        got.append(loss.item())
        assert dcond.shape == ctx.shape
    got, want = np.array(got), np.array(want)
    rel = np.abs(got - want) / want
    print(f"graph={graph}: loss {want[0]:.4f} -> {want[-1]:.4f} (oracle), {got[0]:.4f} -> {got[-1]:.4f} (cuda); "
          f"max rel diff {rel.max():.2e}")
    # two floating-point trajectories of the same optimisation drift apart step by step (batch of 8: every later loss
    # depends on all earlier roundings), so the per-step bar covers the first 60 steps and the bar on the CURVE (10-step
    # moving mean, what the north star's "loss curve within 2 %" is about) covers the whole run
    k = np.ones(10) / 10
    rel_curve = np.abs(np.convolve(got, k, mode="valid") - np.convolve(want, k, mode="valid")) / np.convolve(want, k, mode="valid")
    print(f"first 60 steps max rel diff {rel[:60].max():.2e}; 10-step mean max rel diff {rel_curve.max() if steps >= 10 else 0:.2e}")
    assert rel[:60].max() <= 2e-2
    assert steps < 10 or rel_curve.max() <= 2e-2
    assert rel.max() <= 5e-2
    assert want[-1] < want[0]                                   # it actually trains
    # parameters after the run: same direction of travel as the oracle's
    d_cuda = torch.cat([(q.detach().cpu() - params[n]).flatten() for n, q in m.named_parameters()]).double()
    d_orac = torch.cat([(p[n].detach() - params[n]).flatten() for n, _ in m.named_parameters()]).double()
    cos = torch.dot(d_cuda, d_orac) / (d_cuda.norm() * d_orac.norm())
    print("parameter-update cosine", cos.item())
    assert cos.item() > 0.98
    ema = tr.ema_state()
    assert set(ema) == {n for n, _ in m.named_parameters()}


def test_loss_curve_1k_vs_reference(cuda_device):
    """North-star bar as stated: a 1 000-step training loss curve within 2 % of the reference's.  The reference side is the
    committed fixture tests/golden/loss_curve_tiny_1k.npz: the UNMODIFIED reference UNetModel trained on the CPU with
    torch.optim.AdamW on the same seeded batches (oracle/make_loss_curve.py, run in the build container).  The CUDA side
    is the graph-captured fused step.  Two fp trajectories of a 1 000-step optimisation drift apart step by step, so the
    bar is applied to the curve (20-step moving mean) over the whole run and to every single step of the first 100."""
    import os
    from encdiff_b200.train import DenoiserTrainer
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_curve_tiny_1k.npz"))
    want = fx["losses"]
    steps, B, lr, wd = int(fx["steps"]), int(fx["batch"]), float(fx["lr"]), float(fx["weight_decay"])
    assert steps == 1000 and len(want) == steps
    cfg = CONFIGS["tiny"]
    m, params = make_unet(cfg, cuda_device)
    tr = DenoiserTrainer(m, lr=lr, weight_decay=wd, use_ema=True, use_cuda_graph=True)
    got = []
    for i in range(steps):
        x, t, ctx, noise = seeded_inputs(cfg, B, seed=1000 + i)
        loss, _ = tr.step(x.to(cuda_device), ctx.to(cuda_device), t.to(cuda_device), noise.to(cuda_device))
        got.append(loss.detach().clone())                 # step() returns its static loss buffer: copy, read back at the end
    got = torch.stack([g.reshape(()) for g in got]).cpu().double().numpy()
    rel = np.abs(got - want) / want
    k = np.ones(20) / 20
    sm_got, sm_want = np.convolve(got, k, mode="valid"), np.convolve(want, k, mode="valid")
    rel_sm = np.abs(sm_got - sm_want) / sm_want
    print(f"1k steps: reference {want[0]:.4f} -> {want[-20:].mean():.4f}, cuda {got[0]:.4f} -> {got[-20:].mean():.4f}; "
          f"per-step rel diff max {rel.max():.2e} (first 200: {rel[:200].max():.2e}), median {np.median(rel):.2e}; "
          f"20-step mean rel diff max {rel_sm.max():.2e}")
    assert np.isfinite(got).all()
    # fp32 atomics / TMA reduce-adds make the gradient summation order vary from run to run, so single late steps of the
    # chaotic batch-8 trajectory move by a few per cent between two runs of the SAME binary (observed per-step maxima
    # 4.2e-2 and 5.1e-2, first 200 steps 1.2e-2 and 1.6e-2); the 2 % bar is on the curve, the per-step bars are sanity
    assert rel[:100].max() <= 2e-2
    assert rel[:200].max() <= 3e-2
    assert rel.max() <= 1e-1
    assert rel_sm.max() <= 2e-2
    assert got[-50:].mean() < 0.75 * got[0]               # it trains


def test_loss_curve_shapes3d_vs_reference(cuda_device):
    """The north star's loss-curve bar on the Shapes3D configuration itself (16x16x3 latents of 64x64x3 images, the UNet
    of shapes3d-vq-4-16-encdiff.yaml:33-48): tests/golden/loss_curve_shapes3d_400.npz is the UNMODIFIED reference UNetModel
    trained for 400 AdamW steps on the CPU in fp32 (oracle/make_loss_curve.py shapes3d 400, same seeded batches of 8).
    CUDA side: the graph-captured fused step.  Bar: the CURVE within 2 % over the whole run, read as the 50-step moving
    mean.  Single steps of this batch-8 trajectory are chaotic and the summation order of the TMA reduce-adds in the
    weight-gradient kernel differs from run to run: five runs of ONE build (tools/gpu_r2_n.sh,
    profiles/experiments/r02_loss_curve_spread.txt) gave a worst single step of 4.3-5.9 % (always step 71), a worst 20-step
    mean of 1.1-1.8 % and a worst 50-step mean of 0.75-1.26 %, so single steps and the 20-step mean get sanity bars
    (10 % / 3 %) and the 2 % bar sits on the 50-step mean."""
    import os
    from encdiff_b200.train import DenoiserTrainer
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_curve_shapes3d_400.npz"))
    want = fx["losses"]
    steps, B, lr, wd = int(fx["steps"]), int(fx["batch"]), float(fx["lr"]), float(fx["weight_decay"])
    cfg = CONFIGS["shapes3d"]
    m, params = make_unet(cfg, cuda_device)
    tr = DenoiserTrainer(m, lr=lr, weight_decay=wd, use_ema=True, use_cuda_graph=True)
    got = []
    for i in range(steps):
        x, t, ctx, noise = seeded_inputs(cfg, B, seed=1000 + i)
        loss, _ = tr.step(x.to(cuda_device), ctx.to(cuda_device), t.to(cuda_device), noise.to(cuda_device))
        got.append(loss.detach().clone())
    got = torch.stack([g.reshape(()) for g in got]).cpu().double().numpy()
    rel = np.abs(got - want) / want
    def smooth_rel(n):
        k = np.ones(n) / n
        a, b = np.convolve(got, k, mode="valid"), np.convolve(want, k, mode="valid")
        return np.abs(a - b) / b
    rel_20, rel_50 = smooth_rel(20), smooth_rel(50)
    print(f"shapes3d {steps} steps: reference {want[0]:.4f} -> {want[-20:].mean():.4f}, cuda {got[0]:.4f} -> "
          f"{got[-20:].mean():.4f}; per-step rel diff max {rel.max():.2e} at step {int(rel.argmax())} (first 30: "
          f"{rel[:30].max():.2e}); 20-step mean rel diff max {rel_20.max():.2e}; 50-step mean rel diff max {rel_50.max():.2e}")
    assert np.isfinite(got).all()
    assert rel_50.max() <= 2e-2          # the curve (north star: within 2 %)
    assert rel_20.max() <= 3e-2
    assert rel.max() <= 1e-1
    assert got[-50:].mean() < 0.6 * got[0]
