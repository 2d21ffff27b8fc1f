"""Generate tests/golden/loss_curve_<config>_<steps>.npz: the training loss curve of the UNMODIFIED reference UNetModel
(imported from /root/reference through oracle/ref_shim.py) with torch autograd + torch.optim.AdamW on the CPU, batch 8,
the seeded batches of oracle.make_golden.seeded_inputs(seed = 1000 + step).  It pins the north star's "1k-step training
loss curve within 2 %" bar (tests/test_train_gpu.py) on the tiny configuration (1 000 steps) and on the Shapes3D
configuration itself (shapes3d-vq-4-16-encdiff.yaml:33-48; 400 steps -- the fp32 reference needs ~1 s per step here).
Run in the build container only:  python -m oracle.make_loss_curve [tiny|shapes3d] [steps]
Test infrastructure: nothing under encdiff_b200/ imports this.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import diffusion_oracle as DO  # noqa: E402
from oracle import ref_shim  # noqa: E402
from oracle import unet_oracle as UO  # noqa: E402
from oracle.make_golden import CONFIGS, GOLDEN, seeded_inputs  # noqa: E402

STEPS, BATCH, LR, WD = 1000, 8, 2e-4, 0.01


def main(name="tiny", steps=STEPS):
    # denormal intermediates (softmax tails) put the CPU on its slow path after a few hundred steps (100x slower);
    # flushing them to zero changes no loss digit that the 2 % bar could see
    torch.set_flush_denormal(True)
    cfg = CONFIGS[name]
    params = UO.make_params(cfg, seed=1)
    ref = ref_shim.reference_unet(cfg, params).train()
    for p in ref.parameters():
        p.requires_grad_(True)
    # ddpm_enc.py:1598-1604 (AdamW over model.parameters()), loss = p_losses' L1 (ddpm_enc.py:297-301,1198-1213)
    opt = torch.optim.AdamW(list(ref.parameters()), lr=LR, weight_decay=WD)
    acp = DO.alphas_cumprod(DO.linear_betas())
    losses = []
    t0 = time.time()
    for i in range(steps):
        x, t, ctx, noise = seeded_inputs(cfg, BATCH, seed=1000 + i)
        opt.zero_grad(set_to_none=True)
        eps = ref(DO.q_sample(x, t, noise, acp), t, context=[ctx])
        loss = (noise - eps).abs().mean(dim=[1, 2, 3]).mean()
        loss.backward()
        opt.step()
        losses.append(loss.item())
        if i % 50 == 0:
            print(i, loss.item(), f"{time.time() - t0:.0f}s", flush=True)
    names = [n for n, _ in ref.named_parameters()]
    delta_norm = np.array([(p.detach() - params[n]).norm().item() for n, p in ref.named_parameters()])
    np.savez_compressed(os.path.join(GOLDEN, f"loss_curve_{name}_{'1k' if steps == 1000 else steps}.npz"), losses=np.array(losses, dtype=np.float64),
                        steps=steps, batch=BATCH, lr=LR, weight_decay=WD, param_names=np.array(names),
                        param_delta_norm=delta_norm)
    print("final", losses[-1])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "tiny", int(sys.argv[2]) if len(sys.argv) > 2 else STEPS)
