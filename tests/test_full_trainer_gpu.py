"""encdiff_b200.train.EncDiffTrainer -- the whole training step from images (frozen VQ-f4 encode, trainable concept encoder,
UNet p_losses, ONE AdamW over UNet + concept encoder, EMA on the UNet) -- against the reference's own ``LatentDiffusion``
(ddpm_enc.py: get_input / forward / p_losses) trained with ``torch.optim.AdamW`` over the parameter list of
``configure_optimizers`` (:1598-1604) in fp32 on the same GPU, same images, timesteps and noise."""
import pytest
import torch

from oracle import ref_shim
from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS

pytestmark = pytest.mark.gpu


needs_ref = pytest.mark.skipif(not ref_shim.available(), reason="no reference tree (/root/reference or oracle/_ref)")


def build_pair(cuda_device, lr):
    """(reference LatentDiffusion + torch AdamW over configure_optimizers' parameter list, EncDiffTrainer over this package's
    three modules holding the same weights)."""
    from encdiff_b200 import Encoder4, UNetModel, first_stage
    from encdiff_b200.train import EncDiffTrainer
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    cfg = CONFIGS["shapes3d"]
    params = UO.make_params(cfg, seed=1)
    ref = ref_shim.reference_latent_diffusion(cfg, device=cuda_device, unet_params=params).train()
    assert ref.cond_stage_trainable and type(ref.cond_stage_model).__name__ == "Encoder4"
    vq, _ = first_stage.from_reference(ref.first_stage_model)
    enc = Encoder4(d=128, context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    enc.load_state_dict(ref.cond_stage_model.state_dict(), strict=True)
    enc = enc.to(cuda_device).train()
    unet = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                     model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                     num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                     use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                     context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    unet.load_state_dict(params, strict=True)
    unet = unet.to(cuda_device)
    trainer = EncDiffTrainer(vq, enc, unet, lr=lr, scale_factor=float(ref.scale_factor))
    opt = torch.optim.AdamW(list(ref.model.parameters()) + list(ref.cond_stage_model.parameters()), lr=lr)
    return ref, opt, trainer, enc


def run_steps(ref, opt, trainer, cuda_device, B, steps, seed):
    g = torch.Generator().manual_seed(seed)
    l_ref, l_mine = [], []
    for _ in range(steps):
        img = (torch.rand(B, 64, 64, 3, generator=g) * 2 - 1).to(cuda_device)
        t = torch.randint(0, ref.num_timesteps, (B,), generator=g).to(cuda_device)
        noise = torch.randn(B, 3, 16, 16, generator=g).to(cuda_device)
        # reference: get_input (frozen encode, scale_factor) + forward's conditioning + p_losses, then AdamW
        x, c = ref.get_input({"image": img}, "image")
        cond = ref.get_learned_conditioning(c)
        loss, _ = ref.p_losses(x, cond, t, noise=noise)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        l_ref.append(loss.item())
        l_mine.append(trainer.step(img.permute(0, 3, 1, 2).contiguous(), t, noise).item())
    return l_ref, l_mine


@needs_ref
def test_full_step_from_images_vs_reference_latent_diffusion(cuda_device):
    ref, opt, trainer, enc = build_pair(cuda_device, 1e-4)
    l_ref, l_mine = run_steps(ref, opt, trainer, cuda_device, 16, 6, 41)
    rel = [abs(a - b) / abs(a) for a, b in zip(l_ref, l_mine)]
    print("full step from images, loss per step: reference " + " ".join(f"{v:.5f}" for v in l_ref) + " | EncDiffTrainer " +
          " ".join(f"{v:.5f}" for v in l_mine) + " | rel " + " ".join(f"{v:.1e}" for v in rel))
    assert rel[0] <= 1e-3                      # same weights: forward parity of the whole chain
    assert max(rel) <= 1e-2                    # six optimiser steps on both sides (AdamW's first steps are sign-like)
    # BatchNorm running statistics after six training steps
    sd = enc.state_dict()
    for k, v in ref.cond_stage_model.state_dict().items():
        if "running_mean" in k or "running_var" in k:
            err = float((sd[k] - v).abs().max() / v.abs().max())
            assert err <= 2e-2, (k, err)


@needs_ref
def test_full_pipeline_loss_curve_200_steps(cuda_device):
    """SURVEY.md 8(d) config 2's loss-curve bar (within 2 %) for the WHOLE pipeline: 200 optimiser steps from images with the
    concept encoder training along (BatchNorm in training mode), same images / timesteps / noise on both sides; the
    reference in fp32 eager mode on the same GPU.  Individual steps decorrelate (L1's sign gradient, ReLU masks, float
    atomics), so the bars are on window means, as in test_loss_curve_shapes3d_vs_reference."""
    ref, opt, trainer, _ = build_pair(cuda_device, 1e-4)
    l_ref, l_mine = run_steps(ref, opt, trainer, cuda_device, 16, 200, 43)
    a, b = torch.tensor(l_ref), torch.tensor(l_mine)
    win = [(abs(a[i:i + 25].mean() - b[i:i + 25].mean()) / a[i:i + 25].mean()).item() for i in range(0, 200, 25)]
    last = (abs(a[-50:].mean() - b[-50:].mean()) / a[-50:].mean()).item()
    print(f"full pipeline, 200 steps: reference loss {a[0]:.4f} -> {a[-50:].mean():.4f}, EncDiffTrainer {b[0]:.4f} -> "
          f"{b[-50:].mean():.4f}; 25-step window means differ by " + " ".join(f"{w:.1e}" for w in win) +
          f"; last 50 steps {last:.1e}")
    assert a[-50:].mean() < 0.8 * a[0]            # it does train
    assert last <= 2e-2 and max(win) <= 3e-2
