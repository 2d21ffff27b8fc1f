"""One eager training step (or forward, or the full step from images) of the bench workload bracketed by
cudaProfilerStart/Stop, for `ncu --profile-from-start off`.  usage: python tools/one_step.py [train|fwd|full] [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "train"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    dev = torch.device("cuda:0")
    model = bench.build_model(dev)
    batch = bench.synthetic_batch(B, 23, device=dev)
    if mode == "full":               # EncDiffTrainer.step: VQ encode -> concept encoder -> UNet step -> encoder backward + AdamW
        from encdiff_b200 import Encoder4, VQEncoder
        from encdiff_b200.train import EncDiffTrainer
        vq = VQEncoder(embed_dim=3, ddconfig=dict(double_z=False, z_channels=3, resolution=64, in_channels=3, out_ch=3, ch=32,
                                                  ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[],
                                                  dropout=0.0)).to(dev)
        enc4 = Encoder4(d=128, context_dim=16, latent_unit=20).to(dev).train()
        full = EncDiffTrainer(vq, enc4, model, use_cuda_graph=False)
        img = torch.rand(B, 3, 64, 64, device=dev) * 2 - 1
        run = lambda: full.step(img)
    elif mode == "train":
        from encdiff_b200.train import DenoiserTrainer
        tr = DenoiserTrainer(model, use_cuda_graph=False)
        run = lambda: tr.step(*batch)
    else:
        x0, cond, t, _ = batch
        model.eval()
        run = lambda: model(x0, t, context=[cond])
    with torch.no_grad() if mode == "fwd" else torch.enable_grad():
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        run()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    print("done")


if __name__ == "__main__":
    main()
