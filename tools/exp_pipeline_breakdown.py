"""Where the full-pipeline step (bench.py full_pipeline) goes: CUDA-event time per phase, B = 256; the concept encoder's
trunk through encdiff_enc4_* (the product path) and, for comparison, as the PyTorch nn.Sequential (torch_trunk=True) with
TF32 off / on, channels_last and bf16 autocast."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    B = 256
    dev = torch.device("cuda:0")
    from encdiff_b200 import Encoder4, VQEncoder
    from encdiff_b200.train import DenoiserTrainer
    vq = VQEncoder(embed_dim=3, ddconfig=dict(double_z=False, z_channels=3, resolution=64, in_channels=3, out_ch=3, ch=32,
                                              ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[], dropout=0.0)).to(dev)
    unet = bench.build_model(dev)
    trainer = DenoiserTrainer(unet, lr=1e-4)
    host = (torch.rand(B, 64, 64, 3) * 2 - 1).pin_memory()
    for tf32, cl, ac in (("cuda", False, None), (False, False, None), (True, False, None), (True, True, None),
                         (True, True, torch.bfloat16)):
        cuda_trunk = tf32 == "cuda"
        tf32 = bool(tf32)
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        enc4 = Encoder4(d=128, context_dim=16, latent_unit=20, torch_trunk=not cuda_trunk).to(dev).train()
        if cl:
            enc4.encoder = enc4.encoder.to(memory_format=torch.channels_last)
        opt_c = torch.optim.AdamW(enc4.parameters(), lr=1e-4, fused=True)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
        acc = [0.0] * 7
        n = 0
        for it in range(14):
            ev[0].record()
            img = host.to(dev, non_blocking=True).permute(0, 3, 1, 2)
            img = img.contiguous(memory_format=torch.channels_last) if cl else img.contiguous()
            ev[1].record()
            with torch.no_grad():
                z = vq.encode(img)
            ev[2].record()
            if ac is not None:
                with torch.autocast("cuda", dtype=ac):
                    codes = enc4.encoder(img)
                cond = enc4.net(codes.float())
            else:
                cond = enc4(img)
            ev[3].record()
            t = torch.randint(0, 1000, (B,), device=dev)
            noise = torch.randn_like(z)
            ev[4].record()
            loss, dcond = trainer.step(z, cond.detach(), t, noise)
            ev[5].record()
            opt_c.zero_grad(set_to_none=True)
            cond.backward(dcond)
            ev[6].record()
            opt_c.step()
            ev[7].record()
            torch.cuda.synchronize()
            if it >= 4:
                n += 1
                for k in range(7):
                    acc[k] += ev[k].elapsed_time(ev[k + 1])
        names = ["h2d+permute", "vq.encode", "enc4 fwd", "rng", "unet step", "enc4 bwd", "enc4 adamw"]
        print(f"{'CUDA trunk' if cuda_trunk else 'torch trunk'} tf32={tf32} channels_last={cl} autocast={ac}: total {sum(acc)/n:.2f} ms  " +
              "  ".join(f"{nm} {a/n:.2f}" for nm, a in zip(names, acc)), flush=True)


if __name__ == "__main__":
    main()
