"""Per-block error of the CUDA UNet against the CPU oracle (diagnostic; run on the GPU box).
usage: python tools/diag_blocks.py [tiny|shapes3d] [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import unet_oracle as UO  # noqa: E402
from oracle.make_golden import CONFIGS, seeded_inputs  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "tiny"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    cfg = CONFIGS[name]
    from encdiff_b200 import UNetModel
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    params = UO.make_params(cfg, seed=1)
    m.load_state_dict(params, strict=True)
    m = m.cuda()
    x, t, ctx, _ = seeded_inputs(cfg, B, seed=101)
    taps = {}
    with torch.no_grad():
        eps_o = UO.unet_forward(params, cfg, x, t, [ctx], taps=taps)
        eps = m(x.cuda(), t.cuda(), context=[ctx.cuda()]).cpu()

    def report(k, a, b):
        d = (a - b).abs()
        print(f"{k:20s} max|b| {b.abs().max():8.3f} rms|b| {b.pow(2).mean().sqrt():7.3f}  maxerr/max {d.max() / b.abs().max():.3e}"
              f"  rmserr/rms {d.pow(2).mean().sqrt() / b.pow(2).mean().sqrt():.3e}")

    for k, v in taps.items():
        if k.startswith("input_blocks."):
            sec, idx = 0, int(k.split(".")[1])
        elif k == "middle_block":
            sec, idx = 1, 0
        else:
            sec, idx = 2, int(k.split(".")[1])
        report(k, m.debug_block(sec, idx).cpu(), v)
    report("eps", eps, eps_o)


if __name__ == "__main__":
    main()
