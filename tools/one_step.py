"""One eager training step (or forward) of the bench workload bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off`.  usage: python tools/one_step.py [train|fwd] [B]"""
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
    if mode == "train":
        from encdiff_b200.train import DenoiserTrainer
        tr = DenoiserTrainer(model, use_cuda_graph=False)
        run = lambda: tr.step(*batch)
    else:
        x0, cond, t, _ = batch
        model.eval()
        run = lambda: model(x0, t, context=[cond])
    with torch.no_grad() if mode != "train" else torch.enable_grad():
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
