"""Time the one-launch concept-token MLPs (encdiff_b200.ConceptMLP) against the same arithmetic as a PyTorch loop of
nn.Sequential modules on the same GPU (what Encoder4.forward / warp does), forward and forward + backward, B = 256."""
import json
import os
import sys

import torch
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from encdiff_b200.concept import ConceptMLP  # noqa: E402


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3      # us


def main():
    dev = torch.device("cuda:0")
    L, Dc, B = 20, 16, 256
    torch.manual_seed(0)
    net = nn.ModuleList([nn.Sequential(nn.Linear(1, 64), nn.ELU(True), nn.Linear(64, 128), nn.ELU(True), nn.Linear(128, Dc))
                         for _ in range(L)]).to(dev)
    fast = ConceptMLP.from_reference(net)
    x = torch.randn(B, L, device=dev, requires_grad=True)
    w = torch.randn(B, L * Dc, device=dev)

    def loop_fwd():
        return torch.cat([net[i](x[:, i][:, None]) for i in range(L)], dim=1)

    def loop_fb():
        (loop_fwd() * w).sum().backward()

    def fast_fb():
        (fast(x) * w).sum().backward()

    with torch.no_grad():
        a, b = loop_fwd(), fast(x)
    out = {"max_abs_diff": (a - b).abs().max().item(),
           "torch_loop_fwd_us": timed(lambda: loop_fwd()), "cuda_fwd_us": timed(lambda: fast(x)),
           "torch_loop_fwd_bwd_us": timed(loop_fb), "cuda_fwd_bwd_us": timed(fast_fb), "B": B, "latent_unit": L}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
