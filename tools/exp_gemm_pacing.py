"""Times conv3x3 / linear shapes of the training step through the operator-level C ABI (20 back-to-back launches in a
replayed graph, L2-hot): the micro-benchmark behind profiles/experiments/r01_gemm_pacing.txt.  The ENCDIFF_GEMM_DEBUG
bits that experiment used (drop the loads / the MMAs / the epilogue work) lived in sgemm2_kernel only for that study
(commit "sgemm2 pacing experiment", removed afterwards because every instruction in the single-thread producer / MMA
loops costs step time).  usage: python tools/exp_gemm_pacing.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from encdiff_b200 import _lib, ops  # noqa: E402

SHAPES = [  # (B, H, W, Cin, Cout, taps)
    (256, 16, 16, 64, 64, 9), (256, 16, 16, 128, 128, 9), (256, 8, 8, 256, 256, 9), (256, 4, 4, 256, 256, 9),
    (256, 2, 2, 256, 256, 9), (256, 16, 16, 64, 64, 1), (256, 16, 16, 256, 64, 1), (256, 8, 8, 128, 1024, 1)]


def main():
    dev = torch.device("cuda:0")
    dt = _lib.act_dtype()
    dbg = os.environ.get("ENCDIFF_TAG", "-")
    for (B, H, W, ci, co, taps) in SHAPES:
        x = torch.randn(B, H, W, ci, device=dev).to(dt)
        if taps == 9:
            w = ops.pack_conv3x3(torch.randn(co, ci, 3, 3, device=dev) * 0.05)
        else:
            w = ops.pack_linear(torch.randn(co, ci, device=dev) * 0.05)
        out = torch.empty(B * H * W, co, device=dev, dtype=dt)
        run = lambda: ops.sgemm([x], [taps], w, co, H=H, W=W, B=B, out=out)
        for _ in range(5):
            run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                run()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 100 * 1e3
        M, K = B * H * W, ci * taps
        print(f"{dbg} M={M} N={co} K={K} taps={taps}: {us:7.2f} us  {2.0 * M * co * K / us / 1e6:7.1f} TF/s",
              flush=True)


if __name__ == "__main__":
    main()
