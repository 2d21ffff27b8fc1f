import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, os.getcwd())
import torch, bench
from encdiff_b200.train import DenoiserTrainer
dev = torch.device("cuda:0")
batches = [bench.synthetic_batch(256, 23 + i, device=dev) for i in range(4)]
base = None
for name, pat in [("none", ""), ("sp", "self_attn_bwd_sp"), ("mma", "self_attn_bwd_mma"), ("fwd", "self_attn_fwd_mma"), ("gnb", "gn_bwd_kernel")]:
    os.environ["ENCDIFF_DEBUG_SKIP_KERNELS"] = pat
    m = bench.build_model(dev); tr = DenoiserTrainer(m, lr=1e-4, use_cuda_graph=True)
    for i in range(4): tr.step(*batches[i % 4])
    ms = bench.timed(lambda i: tr.step(*batches[i % 4]), 20, False) / 20
    base = base or ms
    print(f"{name:6s} {ms:.3f} marginal {base-ms:.3f}", flush=True)
    del tr, m; torch.cuda.empty_cache()
