"""Benchmark of the EncDiff denoising hot path on B200 (contract: see the task statement / DESIGN.md §Measurement).

  python bench.py --gpus 1 --steps 20 --warmup 5                       # this repo's CUDA path
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
         bench.py --gpus N --steps K --warmup W                         # data-parallel, one rank per GPU
  python bench.py --impl reference --steps 3 --warmup 1                # the reference algorithm on the host cores

Workload = BASELINE.json configs[1]: Shapes3D EncDiff training step (p_losses forward + backward + gradient all-reduce
+ AdamW/EMA), batch 256 per GPU, synthetic 16x16x3 latents and 20x16 concept tokens, random-init weights with the
zero-initialised layers re-drawn.  One "step" = one optimisation step on one batch per GPU.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GFLOP_FWD = 1.987          # per image, hook-counted on the reference module
GFLOP_TRAIN = 5.960        # fwd + dgrad + wgrad
METRIC = "train_img_per_s"
UNIT = "img/s"
SHAPES3D = dict(image_size=16, in_channels=3, out_channels=3, model_channels=64, attention_resolutions=[1, 2, 4],
                num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8, use_scale_shift_norm=True,
                resblock_updown=True, use_spatial_transformer=True, context_dim=16, latent_unit=20)


# BASELINE.json configs[4]: no yaml in the reference tree; synthesised as SURVEY.md section 7 suggests (twice the width)
CELEBA64 = dict(SHAPES3D, model_channels=128)
WORKLOADS = {"shapes3d": SHAPES3D, "celeba64": CELEBA64}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


def synthetic_batch(B, seed, device="cpu", pin=False):
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(B, 3, 16, 16, generator=g)
    t = torch.randint(0, 1000, (B,), generator=g)
    cond = torch.randn(B, 320, generator=g)
    noise = torch.randn(B, 3, 16, 16, generator=g)
    out = [x0, cond, t, noise]
    if pin:
        out = [v.pin_memory() for v in out]
    elif device != "cpu":
        out = [v.to(device) for v in out]
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                       "200", "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        rows = [r.split(", ") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm = [float(r[1]) for r in rows if len(r) >= 9]
        reasons = set()
        for r in rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": float(rows[0][2]) if rows and len(rows[0]) >= 9 else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port of the reference's torch modules) on the host cores
# ---------------------------------------------------------------------------------------------------------------------
def cpu_train_step_time(B, steps, warmup):
    """(seconds per step, host threads, kind) of the reference's training step on the host cores at batch B.
    kind "reference": the UNMODIFIED reference modules (LatentDiffusion.p_losses -> DiffusionWrapper -> UNetModel, autograd,
    torch.optim.AdamW, LitEma) imported from /root/reference or oracle/_ref (oracle/ref_bench.py).
    kind "port": the oracle's functional port of the same step, only when no reference tree travelled to this box."""
    from oracle import ref_bench
    if ref_bench.available():
        sec, cores, _ = ref_bench.train_step_time(B, steps, warmup, "cpu")
        return sec, cores, "reference"
    from oracle import diffusion_oracle as DO
    from oracle import unet_oracle as UO
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = UO.SHAPES3D
    params = {k: v.clone().requires_grad_(True) for k, v in UO.make_params(cfg, seed=1).items()}
    opt = torch.optim.AdamW(list(params.values()), lr=1e-4)
    acp = DO.alphas_cumprod(DO.linear_betas())
    x0, cond, t, noise = synthetic_batch(B, 23)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        c = cond.clone().requires_grad_(True)
        eps = UO.unet_forward(params, cfg, DO.q_sample(x0, t, noise, acp), t, [c])
        DO.l1_loss(eps, noise).backward()
        opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return statistics.median(times), torch.get_num_threads(), "port"


def run_reference(args):
    """The reference arm: the reference's own CPU implementation of the step at the SAME per-GPU batch as the CUDA arm
    (BASELINE.json configs[1]: 256), all host threads.  One process, whatever --gpus says (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = args.batch
    sec, cores, kind = cpu_train_step_time(B, args.steps, args.warmup)
    v = B / sec
    sample = (f"{args.steps} timed steps (median) of the full per-GPU batch {B}: LatentDiffusion.p_losses fwd + bwd + "
              f"AdamW + LitEma of the {'unmodified reference modules' if kind == 'reference' else 'oracle port'}, fp32")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"{args.workload}_train_step_b{B}_per_gpu", "global_batch": B, "parallelism": "cpu",
                       "step": "q_sample+UNet fwd+L1+UNet bwd+AdamW/EMA"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------------
def measured_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/r02_sgemm2_dram.json, written
    by tools/ncu_dram_summary.py from `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum` of one step of the final
    round-2 build; the round-1 file as a fallback); None if absent."""
    p = os.path.join(ROOT, "profiles", "r02_sgemm2_dram.json")
    if not os.path.exists(p):
        p = os.path.join(ROOT, "profiles", "r01_sgemm2_dram.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return d.get("dram_bytes_per_launch"), d.get("algorithmic_bytes_per_launch")


def build_model(dev, workload="shapes3d"):
    """Reference initialisation under seed 23 (main_val.py:173), then every all-zero parameter tensor (zero_module convs,
    norm biases) re-drawn N(0, 0.02^2) from Generator(1) so that outputs and gradients are non-trivial."""
    from encdiff_b200 import UNetModel
    torch.manual_seed(23)
    m = UNetModel(**WORKLOADS[workload])
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for p in m.parameters():
            if p.abs().max().item() == 0.0:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return m.to(dev)


def timed(fn, steps, dist_on):
    """CUDA-event time of `steps` calls of fn on the current stream, barrier + synchronize on both sides, max over ranks."""
    if dist_on:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if dist_on:
        torch.distributed.barrier()
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return ms.item()


def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist_on = world > 1
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if dist_on:
        torch.distributed.init_process_group("nccl", device_id=dev)
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.train import DenoiserTrainer
    from encdiff_b200 import _lib as edb_lib
    act = "bf16" if edb_lib.act_dtype() == torch.bfloat16 else "fp16"      # 16-bit operands, fp32 accumulate / masters
    B, K, W = args.batch, args.steps, args.warmup
    model = build_model(dev, args.workload)
    trainer = DenoiserTrainer(model, lr=1e-4, use_cuda_graph=not args.no_graph,
                              allreduce_chunks=int(os.environ.get("ENCDIFF_AR_CHUNKS", "1")))
    pk = peaks()

    # ---- (1) device-resident inputs: `value` ---------------------------------------------------------------------
    nb = 4
    dev_batches = [synthetic_batch(B, 23 + 7 * rank + i, device=dev) for i in range(nb)]
    losses = []

    def step_dev(i):
        loss, _ = trainer.step(*dev_batches[i % nb])
        if i == 0 or i == K - 1:
            losses.append(loss.clone())

    for i in range(max(W, 3)):
        trainer.step(*dev_batches[i % nb])
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(step_dev, K, dist_on)
    # ---- (2) host buffers through the public step(): `e2e` (H2D of the batch + D2H of the loss inside the region) --
    host_batches = [synthetic_batch(B, 123 + 7 * rank + i, pin=True) for i in range(nb)]
    loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    seen = []

    def step_host(i):
        # H2D of the batch (inside trainer.step) -> step -> D2H of the loss, every step.  The host reads each loss one
        # step late (after its copy event), the way a training loop logs it, so launching step i+1 overlaps step i.
        loss, _ = trainer.step(*host_batches[i % nb])
        loss_host[i & 1].copy_(loss, non_blocking=True)
        loss_ev[i & 1].record()
        if i > 0:
            loss_ev[(i - 1) & 1].synchronize()
            seen.append(float(loss_host[(i - 1) & 1][0]))

    for i in range(3):
        step_host(i)
    torch.cuda.synchronize()
    ms_e2e = timed(step_host, K, dist_on)       # timed() synchronises at the end: the last loss has landed too
    clk = clocks.stop() if rank == 0 else None
    h2d = sum(v.numel() * v.element_size() for v in host_batches[0])

    # ---- (3) per-kernel-family device time of the same step (eager, one event per launch) --------------------------
    trainer.use_cuda_graph = False
    model.profile(True)
    model.profile_read()
    P = 3
    for i in range(P):
        trainer.step(*dev_batches[i % nb])
    torch.cuda.synchronize()
    prof = model.profile_read()
    model.profile(False)
    launches_per_step = sum(v[0] for v in prof.values()) // P + 3      # + l1_loss, adamw, pack
    fam = {}
    for name, (n, tms, fl, by) in prof.items():
        fam[name] = {"launches_per_step": n // P, "ms_per_step": tms / P, "gflop_per_step": fl / P / 1e9,
                     "gbyte_per_step": by / P / 1e9}
    tot_ms = sum(v["ms_per_step"] for v in fam.values())
    sg = fam.get("sgemm", {"ms_per_step": 0.0, "gflop_per_step": 0.0, "launches_per_step": 0})
    achieved = sg["gflop_per_step"] / max(sg["ms_per_step"], 1e-9)        # GFLOP/ms == TFLOP/s
    traffic = measured_traffic()[0] if args.workload == "shapes3d" else None     # the ncu capture is of the Shapes3D step
    alg_bytes = sg.get("gbyte_per_step", 0.0) * 1e9 / max(sg["launches_per_step"], 1)
    roof = {"bound": "tensor", "kernel": "sgemm2_kernel (persistent tcgen05 implicit GEMM: conv3x3/1x1/linear fwd + dgrad)",
            "achieved": achieved, "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sustained"],
            "traffic": traffic, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": pk["source"] + " (sustained cuBLAS bf16)",
            "launches_per_step": sg["launches_per_step"], "share_of_step": sg["ms_per_step"] / max(tot_ms, 1e-9),
            "families": {k: {"ms": round(v["ms_per_step"], 4), "n": v["launches_per_step"],
                             "share": round(v["ms_per_step"] / max(tot_ms, 1e-9), 4)}
                         for k, v in sorted(fam.items(), key=lambda kv: -kv[1]["ms_per_step"])}}

    # ---- (3b) the same family INSIDE the replayed graph: marginal cost by ablation ---------------------------------
    # The per-launch events above serialise an eager step (launch gaps and cold prologues included: the families sum to
    # ~13 ms for a ~9 ms graph step).  What the family costs in the captured step is measured by capturing the step once
    # more WITHOUT its launches (ENCDIFF_DEBUG_SKIP_KERNELS, common.cuh launch_k; results of that run are garbage, only
    # its time is used) on a second model: in-graph ms = step - step_without_family.  Reported next to the event figure.
    if world == 1 and not args.no_graph:
        os.environ["ENCDIFF_DEBUG_SKIP_KERNELS"] = "sgemm2_kernel"
        try:
            m2 = build_model(dev, args.workload)
            tr2 = DenoiserTrainer(m2, lr=1e-4, use_cuda_graph=True)
            for i in range(4):
                tr2.step(*dev_batches[i % nb])
            ms_wo = timed(lambda i: tr2.step(*dev_batches[i % nb]), 10, False) / 10
            del tr2, m2
            torch.cuda.empty_cache()
        finally:
            os.environ["ENCDIFF_DEBUG_SKIP_KERNELS"] = ""
        marg = ms / K - ms_wo
        # The headline `value` is the replayed graph, so the family's time INSIDE that graph is the consistent
        # denominator: `achieved` / `frac` are the in-graph figures, the eager per-launch event figures stay beside them.
        roof["eager_events"] = {"achieved": roof["achieved"], "frac": roof["frac"], "family_ms_per_step": sg["ms_per_step"],
                                "note": "one CUDA event after every launch of an eager (un-captured) step: launch gaps "
                                        "and cold prologues are inside each interval"}
        roof["achieved"] = sg["gflop_per_step"] / max(marg, 1e-9)
        roof["frac"] = roof["achieved"] / pk["tf_sustained"]
        roof["timing"] = ("in-graph: CUDA-event time of the captured step minus the time of the same step captured without "
                          "the family's launches (ENCDIFF_DEBUG_SKIP_KERNELS ablation, 10 steps each)")
        roof["family_ms_per_step"] = marg
        roof["ms_per_step_without_family"] = ms_wo
        roof["share_of_step"] = marg / (ms / K)

    # ---- (4) DDIM-50 sampling throughput (configs[2] shape: batch-sharded, no communication) -----------------------
    ddim = swap = None
    if not args.no_ddim:
        ddim = ddim50(model, dev, args.ddim_batch, dist_on, world)
        swap = swap200(model, dev, args.swap_sources, dist_on, world)

    if rank != 0:
        if dist_on:
            torch.distributed.destroy_process_group()
        return
    img_s = world * B * K / (ms / 1e3)
    wl = f"{args.workload}_train_step_b{B}_per_gpu"
    # algorithmic work of the step: SURVEY.md's hook count for the Shapes3D UNet, the executor's own per-launch count
    # (2 M N K of every GEMM it launched) for a synthesised configuration
    gflop_img = GFLOP_TRAIN if args.workload == "shapes3d" else sum(v["gflop_per_step"] for v in fam.values()) / B
    e2e_img_s = world * B * K / (ms_e2e / 1e3)
    line = {"metric": METRIC, "value": img_s, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": act,
            "data": "synthetic",
            "config": {"workload": wl,
                       "global_batch": world * B, "parallelism": f"dp{world}", "cuda_graph": not args.no_graph,
                       "l2": "working set per step (activation workspace %.2f GB) exceeds the 126 MB L2; 4 rotating "
                             "input batches" % (model._lib.encdiff_unet_workspace_bytes(model._h, B, 1) / 1e9),
                       "step": "q_sample+UNet fwd+L1+UNet bwd+grad all-reduce+AdamW/EMA+16-bit weight re-pack",
                       "precision": act + " tensor-core operands and activations, fp32 accumulation, fp32 master weights/optimizer"},
            "e2e": {"value": e2e_img_s, "unit": UNIT, "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches_per_step * K),
            "clocks": clk,
            "roofline": roof,
            "step_tensor_frac": img_s / world * gflop_img / 1e3 / pk["tf_sustained"],
            "gflop_per_img": gflop_img,
            "loss_first_last": [losses[0].item(), losses[-1].item()] if losses else None}
    if ddim is not None:
        line["ddim50"] = ddim
    if swap is not None:
        line["swap_ddim200"] = swap
    if world == 1 and not args.no_cpu:
        sec, cores, kind = cpu_train_step_time(B, 3, 1)
        cfg1 = None
        if kind == "reference":
            try:
                from oracle import ref_bench
                s1, c1, _ = ref_bench.full_step_time(8, 3, 1, "cpu")
                cfg1 = {"img_per_s": 8 / s1, "ms_per_step": s1 * 1e3, "batch": 8, "cores": c1,
                        "what": "configs[0]: the complete reference step from images (VQ-f4 encode + Encoder4 + UNet p_losses "
                                "fwd+bwd + AdamW + LitEma), batch 8, host cores"}
            except Exception as e:
                cfg1 = {"error": repr(e)[:200]}
        line["config1_cpu_full_step"] = cfg1
        line["cpu_baseline"] = {"value": B / sec, "unit": UNIT, "cores": cores, "kind": kind,
                                "sample": f"3 timed steps (median, 1 warm-up) of the full batch {B}, fp32, all host threads: "
                                          + ("the unmodified reference LatentDiffusion.p_losses fwd+bwd+AdamW+LitEma "
                                             "(oracle/_ref)" if kind == "reference" else "oracle port of p_losses fwd+bwd+AdamW")}
    if world == 1 and not args.no_comparators:
        line["comparators"] = gpu_comparators(model, dev, B, args.ddim_batch)
        if args.workload == "shapes3d":
            try:
                line["full_pipeline"] = full_pipeline(dev, B)
            except Exception as e:           # an extra leg must never take the headline down with it
                line["full_pipeline_error"] = repr(e)[:300]
    print(json.dumps(line), flush=True)
    if dist_on:
        torch.distributed.destroy_process_group()


def gpu_comparators(model, dev, B, ddim_B):
    """Same-GPU context for the headline, each a few steps after warm-up:
      ref_eager_gpu_tf32 / ref_eager_gpu_bf16: the UNMODIFIED reference modules (oracle/_ref) in PyTorch eager mode on this
        B200 -- TF32 matmul/conv as the reference ships (main_val.py:524-525) and under bf16 autocast -- training step at
        the benched batch and DDIM-50 at the benched sampling batch;
      dropin_autograd: what a reference user gets from the documented drop-in seam without the fused trainer:
        encdiff_b200.UNetModel.forward + loss.backward() through torch autograd + torch.optim.AdamW."""
    import gc
    out = {}
    try:
        from oracle import ref_bench
        if ref_bench.available():
            ref = ref_bench.build(dev)
            for name, kw in (("ref_eager_gpu_tf32", dict(tf32=True)),
                             ("ref_eager_gpu_bf16", dict(tf32=True, autocast_dtype=torch.bfloat16))):
                sec, _, _ = ref_bench.train_step_time(B, 5, 2, dev, model=ref, **kw)
                dsec = ref_bench.ddim_time(ddim_B, 50, 2, dev, model=ref, **kw)
                out[name] = {"train_img_per_s": B / sec, "train_ms_per_step": sec * 1e3, "batch": B,
                             "ddim50_img_per_s": ddim_B / dsec, "ddim50_batch": ddim_B}
                if name == "ref_eager_gpu_tf32":      # the COMPLETE reference step from images (configs[0] at the benched batch)
                    fsec, _, _ = ref_bench.full_step_time(B, 3, 2, dev, tf32=True, model=ref)
                    out[name]["full_step_from_images_img_per_s"] = B / fsec
                    out[name]["full_step_from_images_ms"] = fsec * 1e3
            del ref
            gc.collect()
            torch.cuda.empty_cache()
        else:
            out["ref_eager_gpu"] = "unavailable: no reference tree on this box (oracle/_ref missing)"
    except Exception as e:           # a comparator must never take the headline down with it
        out["ref_eager_gpu_error"] = repr(e)[:300]
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        x0, cond, t, noise = synthetic_batch(B, 23, device=dev)
        model.train()
        model.cache_packed_weights = False
        opt = torch.optim.AdamW(list(model.parameters()), lr=1e-4)
        for p in model.parameters():
            p.grad = None
        times = []
        for i in range(2 + 5):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            opt.zero_grad(set_to_none=True)
            c = cond.clone().requires_grad_(True)
            eps = model(x0, t, context=[c])
            (eps - noise).abs().mean().backward()
            opt.step()
            torch.cuda.synchronize()
            if i >= 2:
                times.append(time.perf_counter() - t0)
        sec = statistics.median(times)
        out["dropin_autograd"] = {"train_img_per_s": B / sec, "train_ms_per_step": sec * 1e3, "batch": B,
                                  "path": "encdiff_b200.UNetModel.forward + loss.backward() through torch autograd (the library replays captured graphs of its forward / backward launch sequences) + torch.optim.AdamW"}
    except Exception as e:
        out["dropin_autograd_error"] = repr(e)[:300]
    return out


def full_pipeline(dev, B, steps=10):
    """The COMPLETE EncDiff training step from images through this package (configs[0] / [1] as a training loop runs them):
    pinned host images -> H2D -> frozen VQ-f4 encode (encdiff_b200.VQEncoder) -> concept encoder (encdiff_b200.Encoder4:
    encdiff_enc4_* trunk + the one-launch ConceptMLP) -> fused UNet step (DenoiserTrainer: q_sample, forward, L1, backward,
    AdamW + EMA, re-pack) -> d loss / d tokens back through the concept encoder (encdiff_enc4_backward) -> its AdamW.
    Synthetic images, random-init weights; the loss is read back every step."""
    from encdiff_b200 import Encoder4, VQEncoder
    from encdiff_b200.train import EncDiffTrainer
    torch.manual_seed(23)
    vq = VQEncoder(embed_dim=3, ddconfig=dict(double_z=False, z_channels=3, resolution=64, in_channels=3, out_ch=3, ch=32,
                                              ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[], dropout=0.0)).to(dev)
    enc4 = Encoder4(d=128, context_dim=16, latent_unit=20).to(dev).train()
    trainer = EncDiffTrainer(vq, enc4, build_model(dev), lr=1e-4)
    g = torch.Generator().manual_seed(31)
    host = [(torch.rand(B, 64, 64, 3, generator=g) * 2 - 1).pin_memory() for _ in range(3)]
    loss_host = torch.zeros(1).pin_memory()

    def step(i):
        img = host[i % 3].to(dev, non_blocking=True).permute(0, 3, 1, 2).contiguous()    # 'b h w c -> b c h w', get_input
        loss = trainer.step(img)              # t = randint, noise = randn_like drawn on the device, as forward / p_losses do
        loss_host.copy_(loss, non_blocking=True)

    for i in range(4):
        step(i)
    ms = timed(step, steps, False) / steps
    return {"img_per_s": B / (ms / 1e3), "ms_per_step": ms, "batch": B,
            "h2d_bytes_per_step": host[0].numel() * 4,
            "path": "encdiff_b200.train.EncDiffTrainer.step: images (pinned host) -> VQEncoder.encode -> Encoder4 (CUDA trunk + "
                    "ConceptMLP) -> DenoiserTrainer.step -> concept-encoder backward -> its fused AdamW"}


def ddim50(model, dev, B, dist_on, world):
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.diffusion import DenoiserHost
    host = DenoiserHost(model).to(dev)
    sampler = DDIMSampler(host)
    g = torch.Generator().manual_seed(5)
    cond = torch.randn(B, 320, generator=g).to(dev)
    xT = torch.randn(B, 3, 16, 16, generator=g).to(dev)
    model.eval()
    sampler.sample(50, B, (3, 16, 16), cond, eta=0., x_T=xT, verbose=False)       # warm-up: captures the step graph
    reps = 3
    ms = timed(lambda i: sampler.sample(50, B, (3, 16, 16), cond, eta=0., x_T=xT, verbose=False), reps, dist_on)
    img_s = world * B * reps / (ms / 1e3)
    pk = peaks()
    return {"img_per_s": img_s, "batch_per_gpu": B, "ms_per_sample_call": ms / reps,
            "tensor_frac": img_s / world * 50 * GFLOP_FWD / 1e3 / pk["tf_sustained"]}


def swap200(model, dev, n_src, dist_on, world):
    """configs[3]: concept-token swap generation, DDIM-200, eta 0 (generate_swap_from_ckpt.py:103-105): n_src source images
    per GPU x 20 swapped units folded into one batch of 20 * n_src samples per GPU; no communication in the loop."""
    from encdiff_b200 import DDIMSampler
    from encdiff_b200.diffusion import DenoiserHost
    from encdiff_b200.swap import sample_swap
    host = DenoiserHost(model).to(dev)
    sampler = DDIMSampler(host)
    L, dctx = model.latent_unit, model.context_dim
    g = torch.Generator().manual_seed(9)
    tokens = torch.randn(n_src, L, dctx, generator=g).to(dev)     # stands for warp(encoding(images)): frozen front end
    model.eval()
    run = lambda i: sample_swap(sampler, tokens, None, ddim_steps=200, eta=0., shape=(3, 16, 16), shard=False,
                                reference_rng=False)
    run(0)                                                         # warm-up: captures the step graph
    reps = 2
    ms = timed(run, reps, dist_on)
    img_s = world * L * n_src * reps / (ms / 1e3)
    pk = peaks()
    return {"img_per_s": img_s, "sources_per_gpu": n_src, "folded_batch_per_gpu": L * n_src, "ddim_steps": 200,
            "ms_per_call": ms / reps, "tensor_frac": img_s / world * 200 * GFLOP_FWD / 1e3 / pk["tf_sustained"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch (BASELINE.json configs[1]: 256)")
    ap.add_argument("--ddim-batch", type=int, default=128, help="per-GPU DDIM-50 batch (configs[2]: 1024 over 8 GPUs)")
    ap.add_argument("--swap-sources", type=int, default=8, help="source images per GPU of the swap leg (x 20 units)")
    ap.add_argument("--workload", default="shapes3d", choices=sorted(WORKLOADS),
                    help="shapes3d = configs[1] (the headline); celeba64 = the synthesised configs[4] UNet")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-ddim", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-comparators", action="store_true",
                    help="skip the same-GPU comparators (reference modules in eager mode, drop-in autograd path)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
