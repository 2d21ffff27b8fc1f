"""Per-launch time of back-to-back dependent kernels inside one CUDA graph: a trivial elementwise kernel vs small GEMMs."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from encdiff_b200 import _lib, ops  # noqa: E402


def time_graph(fn, n=200):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5 / n * 1e3


def main():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    dt = _lib.act_dtype()
    x = torch.randn(4096, device=dev)
    y = torch.empty(4096, dtype=dt, device=dev)
    us = time_graph(lambda: _lib.check(lib.encdiff_timestep_embedding(x.data_ptr(), y.data_ptr(), 64, 64, _lib.current_stream())))
    print(f"temb (trivial kernel)               {us:6.2f} us/launch")
    for (B, H, C, N, taps) in [(256, 4, 256, 256, 1), (256, 16, 64, 64, 1), (256, 8, 128, 128, 1), (256, 2, 256, 256, 9),
                               (256, 4, 256, 256, 9), (256, 16, 64, 64, 9), (256, 8, 128, 128, 9), (256, 16, 128, 128, 9)]:
        a = torch.randn(B, H, H, C, device=dev).to(dt)
        o = torch.empty(B * H * H, N, device=dev, dtype=dt)
        w = torch.randn(N, C * taps, device=dev) * 0.05
        if taps == 9:
            wp = ops.pack_conv3x3(w.view(N, C, 3, 3), rows_pad=256)
        else:
            wp = ops.pack_linear(w, rows_pad=256)
        res = torch.randn(B * H * H, N, device=dev).to(dt)
        us = time_graph(lambda: ops.sgemm([a], [taps], wp, N, H=H, W=H, B=B, res=res, out=o))
        fl = 2.0 * B * H * H * N * C * taps
        print(f"sgemm M={B * H * H:6d} N={N:4d} K={C * taps:5d} taps={taps}   {us:6.2f} us/launch  {fl / us / 1e6:7.1f} TF/s")


if __name__ == "__main__":
    main()
