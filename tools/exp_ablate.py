"""Marginal in-graph cost of each kernel family of the B = 256 training step: the step graph is re-captured with one family
skipped (ENCDIFF_DEBUG_SKIP_KERNELS, common.cuh launch_k) and the step-time difference is that family's cost INSIDE the
replayed graph (launch latency + execution - whatever overlapped under programmatic dependent launch).  The per-family
event timings of bench.py are eager and serialised (sum 13 ms for a 9.8 ms step); this is the number that says where the
graph step really goes.  Results of an ablated run are garbage by construction; only the time is used.
usage: python tools/exp_ablate.py [B] [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

FAMILIES = [
    ("(none)", ""),
    ("sgemm2 all", "sgemm2_kernel"),
    ("sgemm2<64,3,2,1> shallow", "sgemm2_kernelILi64ELi3ELi2ELi1"),
    ("sgemm2<64,3,1,2> deep", "sgemm2_kernelILi64ELi3ELi1ELi2"),
    ("sgemm2<128,2,2,1> shallow", "sgemm2_kernelILi128ELi2ELi2ELi1"),
    ("sgemm2<128,2,1,2> deep", "sgemm2_kernelILi128ELi2ELi1ELi2"),
    ("sgemm2<256,3,1,1>", "sgemm2_kernelILi256ELi3ELi1ELi1"),
    ("wgrad2_group", "wgrad2_group_kernel"),
    ("gn_fwd", "gn_fwd_kernel"),
    ("gn_bwd", "gn_bwd_kernel"),
    ("ln_fwd", "ln_fwd_kernel"),
    ("ln_bwd", "ln_bwd_kernel"),
    ("geglu_fwd", "geglu_fwd_kernel"),
    ("geglu_bwd", "geglu_bwd_kernel"),
    ("self_attn_fwd", "self_attn_fwd_mma_kernel"),
    ("self_attn_bwd", "self_attn_bwd_mma_kernel"),
    ("self_attn_bwd<8>", "self_attn_bwd_mma_kernelILi8"),
    ("cross_attn_fwd", "cross_attn_fwd_mma_kernel"),
    ("cross_attn_bwd", "cross_attn_bwd_mma_kernel"),
    ("cross_attn_bwd_proj", "cross_attn_bwd_proj_group_kernel"),
    ("adamw", "adamw_dev"),
    ("pack", "pack_kernel"),
    ("wgrad_finalize", "wgrad_finalize_kernel"),
    ("all norms+geglu", "gn_fwd_kernel,gn_bwd_kernel,ln_fwd_kernel,ln_bwd_kernel,geglu_fwd_kernel,geglu_bwd_kernel"),
    ("all attention", "self_attn_fwd_mma,self_attn_bwd_mma,cross_attn_fwd_mma,cross_attn_bwd_mma"),
]


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    dev = torch.device("cuda:0")
    from encdiff_b200.train import DenoiserTrainer
    batches = [bench.synthetic_batch(B, 23 + i, device=dev) for i in range(4)]
    base = None
    for name, pat in FAMILIES:
        os.environ["ENCDIFF_DEBUG_SKIP_KERNELS"] = pat
        model = bench.build_model(dev)
        tr = DenoiserTrainer(model, lr=1e-4, use_cuda_graph=True)
        for i in range(4):
            tr.step(*batches[i % 4])
        ms = bench.timed(lambda i: tr.step(*batches[i % 4]), steps, False) / steps
        if base is None:
            base = ms
        print(f"{name:28s} {ms:8.3f} ms/step   marginal {base - ms:7.3f} ms", flush=True)
        del tr, model
        torch.cuda.empty_cache()
    os.environ["ENCDIFF_DEBUG_SKIP_KERNELS"] = ""


if __name__ == "__main__":
    main()
