"""Experiment: does running the step as TWO independent half-batch chains on two streams hide the per-launch latency that
prices the captured step (~800 dependent launches at ~5 us)?  Two separate models / trainers (so no
shared state), each with B = 128, replayed concurrently on two streams, against ONE model at B = 256.  Upper bound on what a
shared-parameter two-chain step could give."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from encdiff_b200.train import DenoiserTrainer  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    steps = 20

    def make(B, seed):
        m = bench.build_model(dev).train()
        tr = DenoiserTrainer(m, lr=1e-4)
        batch = bench.synthetic_batch(B, seed, device=dev)
        return tr, batch

    def run(tr, batch):
        x0, cond, t, noise = batch
        tr.step(x0, cond, t, noise)

    whole = make(256, 1)
    for _ in range(5):
        run(*whole)
    ms = bench.timed(lambda i: run(*whole), steps, False) / steps
    print(f"one chain  B=256: {ms:.3f} ms/step  ({256 / ms:.1f} k img/s)")
    for parts, Bp in ((2, 128), (4, 64)):
        chains = [make(Bp, 10 + k) for k in range(parts)]
        streams = [torch.cuda.Stream() for _ in range(parts)]
        for c, s in zip(chains, streams):
            with torch.cuda.stream(s):
                for _ in range(5):
                    run(*c)
        torch.cuda.synchronize()
        ms_seq = bench.timed(lambda i: [run(*c) for c in chains], steps, False) / steps
        cur = torch.cuda.current_stream()

        def concurrent(i):
            ev = torch.cuda.Event()
            ev.record(cur)
            for c, s in zip(chains, streams):
                s.wait_event(ev)
                with torch.cuda.stream(s):
                    run(*c)
                done = torch.cuda.Event()
                done.record(s)
                cur.wait_event(done)

        ms_con = bench.timed(concurrent, steps, False) / steps
        print(f"{parts} chains B={Bp}: sequential {ms_seq:.3f} ms, concurrent on {parts} streams {ms_con:.3f} ms per {parts * Bp} images "
              f"({parts * Bp / ms_con:.1f} k img/s)")
        del chains


if __name__ == "__main__":
    main()
