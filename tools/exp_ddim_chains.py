"""Experiment: DDIM-50 of B images as n independent sampler chains of B/n on n streams (step graphs replayed round-robin)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from encdiff_b200 import DDIMSampler  # noqa: E402
from encdiff_b200.diffusion import DenoiserHost  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    dev = torch.device("cuda:0")
    for n in (1, 2, 4):
        b = B // n
        samplers, streams, args = [], [], []
        for i in range(n):
            m = bench.build_model(dev).eval()
            s = DDIMSampler(DenoiserHost(m).to(dev))
            g = torch.Generator().manual_seed(5 + i)
            cond = torch.randn(b, 320, generator=g).to(dev)
            xT = torch.randn(b, 3, 16, 16, generator=g).to(dev)
            s.sample(50, b, (3, 16, 16), cond, eta=0., x_T=xT, verbose=False)     # captures the step graph
            samplers.append(s)
            streams.append(torch.cuda.Stream())
            args.append((cond, xT))
        graphs = [next(iter(s._graph_cache.values()))["graph"] for s in samplers]
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        e0.record()
        cur = torch.cuda.current_stream()
        for st in streams:
            st.wait_stream(cur)
        for _ in range(reps * 50):
            for st, g in zip(streams, graphs):
                with torch.cuda.stream(st):
                    g.replay()
        for st in streams:
            cur.wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"chains {n} x B {b}: {ms:.1f} ms per DDIM-50 of {B} -> {B / ms * 1e3:.0f} img/s", flush=True)
        del samplers, graphs
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
