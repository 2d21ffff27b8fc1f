"""Experiment: one B=256 fwd+bwd chain vs N independent chains of B=256/N captured on N streams in one graph
(are the kernels latency-bound enough that independent chains overlap?).  usage: python tools/exp_chains.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from encdiff_b200.train import DenoiserTrainer  # noqa: E402


def prep(tr, B, dev, seed):
    st = tr._make_static(B, (3, 16, 16), 320)
    x0, cond, t, noise = bench.synthetic_batch(B, seed, device=dev)
    st["x0"].copy_(x0); st["cond"].copy_(cond); st["t"].copy_(t); st["noise"].copy_(noise)
    u = tr.unet
    u._bind(need_grads=True)
    u._prepare_grads()
    u.refresh_weights()
    u.cache_packed_weights = True
    return st


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dev = torch.device("cuda:0")
    for n in (1, 2, 4):
        trs = [DenoiserTrainer(bench.build_model(dev), use_cuda_graph=False) for _ in range(n)]
        sts = [prep(tr, B // n, dev, 5 + i) for i, tr in enumerate(trs)]
        for tr, st in zip(trs, sts):
            tr._fwd_bwd(st)
        torch.cuda.synchronize()
        streams = [torch.cuda.Stream() for _ in range(n - 1)]
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            cur = torch.cuda.current_stream()
            for s in streams:
                s.wait_stream(cur)
            trs[0]._fwd_bwd(sts[0])
            for s, tr, st in zip(streams, trs[1:], sts[1:]):
                with torch.cuda.stream(s):
                    tr._fwd_bwd(st)
            for s in streams:
                cur.wait_stream(s)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        print(f"chains {n} x B {B // n}: {e0.elapsed_time(e1) / 10:.3f} ms per fwd+bwd of {B} images", flush=True)
        del g, trs, sts
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
