"""Generate tests/golden/swap_shapes3d_n2_s10.npz by running the UNMODIFIED reference's own swap loop:
``LatentDiffusion.log_images(batch, N, sample_swap=True, ddim_steps=S, ddim_eta=0, ...)`` (ddpm_enc.py:1473-1535, as
generate_swap_from_ckpt.py:99-109 calls it) on the real ``LatentDiffusion`` built by ``instantiate_from_config`` from the
Shapes3D yaml values (oracle/ref_shim.py), CPU, seeded.  Recorded: the encoder's codes ``orc``, the weights of
``Encoder4.net`` (what ``warp`` applies), the start noise of each of the 20 runs in the order the reference drew it, and
the 20 x N latents handed to ``decode_first_stage`` -- i.e. the reference-side pin of SURVEY.md section 8 row a17.
Run in the build container only:  python -m oracle.make_swap_golden        Test infrastructure.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from oracle import unet_oracle as UO  # noqa: E402
from oracle.make_golden import GOLDEN  # noqa: E402

N, S = 2, 10


def main():
    cfg = UO.SHAPES3D
    model = ref_shim.reference_latent_diffusion(cfg, unet_params=UO.make_params(cfg, seed=1)).eval()
    import ldm.models.diffusion.ddpm_enc as ddpm_enc
    x_Ts, latents = [], []

    class RecordingSampler(model._cpu_ddim_sampler):
        def ddim_sampling(self, cond, shape, x_T=None, **kw):
            if x_T is None:
                x_T = torch.randn(shape)          # the draw ddim.py:122-125 would have made at this point
            x_Ts.append(x_T.clone())
            return super().ddim_sampling(cond, shape, x_T=x_T, **kw)

    ddpm_enc.DDIMSampler = RecordingSampler        # sample_log resolves the name at call time (ddpm_enc.py:1446)
    dec = model.decode_first_stage

    def recording_decode(z, *a, **k):
        latents.append(z.detach().clone())
        return dec(z, *a, **k)

    g = torch.Generator().manual_seed(23)
    batch = {"image": torch.rand(N, 64, 64, 3, generator=g) * 2 - 1}
    torch.manual_seed(4242)
    with torch.no_grad():
        z, c, x, xrec, xc, orc = model.get_input(batch, "image", return_first_stage_outputs=True, force_c_encode=True,
                                                 return_original_cond=True, bs=N, return_false=True)
        model.decode_first_stage = recording_decode
        log = model.log_images(batch, N=N, sample=False, sample_swap=True, ddim_steps=S, ddim_eta=0.0, inpaint=False,
                               plot_denoise_rows=False, plot_progressive_rows=False, plot_diffusion_rows=False)
    # log_images decodes the reconstruction first (get_input), then one decode per swapped unit
    L = cfg.latent_unit
    swaps = latents[-L:]
    assert len(x_Ts) == L and log["samples_swapping"].shape[0] == L * N
    out = {"orc": orc.numpy(), "x_T": torch.stack(x_Ts).numpy(), "samples": torch.cat(swaps, 0).numpy(),
           "N": N, "S": S, "eta": 0.0}
    for k, v in model.cond_stage_model.net.state_dict().items():
        out["net::" + k] = v.numpy()
    np.savez_compressed(os.path.join(GOLDEN, f"swap_shapes3d_n{N}_s{S}.npz"), **out)
    print("swap golden: orc", orc.shape, "samples", out["samples"].shape, "absmax", np.abs(out["samples"]).max())


if __name__ == "__main__":
    main()
