"""Per-launch-family CUDA-event times of the concept encoder's trunk (encdiff_enc4_forward / backward) at B = 256, and the
whole forward / backward timed with events around the calls (eager launches, no profiling marks)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from encdiff_b200 import Encoder4, _lib  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    enc = Encoder4(d=128, context_dim=16, latent_unit=20).to(dev).train()
    x = torch.rand(B, 3, 64, 64, device=dev) * 2 - 1
    dcodes = torch.randn(B, 20, device=dev) / B
    lib = enc._lib

    def fb():
        c = enc.encoding(x)
        c.backward(dcodes)

    for _ in range(3):
        fb()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tf = tb = 0.0
    for _ in range(10):
        ev[0].record()
        c = enc.encoding(x)
        ev[1].record()
        c.backward(dcodes)
        ev[2].record()
        torch.cuda.synchronize()
        tf += ev[0].elapsed_time(ev[1]) / 10
        tb += ev[1].elapsed_time(ev[2]) / 10
    print(f"B={B}: forward {tf:.3f} ms, backward {tb:.3f} ms (eager launches incl. weight pack / grad bucket copy)")
    _lib.check(lib.encdiff_unet_profile(enc._h, 1), "profile")
    agg = {}
    for _ in range(5):
        fb()
        buf = C.create_string_buffer(1 << 16)
        n = lib.encdiff_unet_profile_read(enc._h, buf, len(buf))
        assert n >= 0
        for line in buf.value.decode().splitlines():
            name, k, ms, fl, by = line.split(",")
            a = agg.setdefault(name, [0, 0.0, 0.0])
            a[0] += int(k) / 5
            a[1] += float(ms) / 5
            a[2] += float(fl) / 5
    tot = sum(a[1] for a in agg.values())
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        tfs = f"{a[2] / a[1] / 1e9:8.1f} TF/s" if a[2] else ""
        print(f"  {name:28s} {a[0]:5.0f} launches {a[1] * 1e3:9.1f} us  {tfs}")
    print(f"  total {tot:.3f} ms (serialised, event-timed)")


if __name__ == "__main__":
    main()
