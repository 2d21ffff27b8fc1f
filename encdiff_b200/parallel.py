"""Data-parallel host logic of the hot path: one process per GPU, batch sharding, ONE gradient
all-reduce per training step over a flat fp32 bucket, and no communication at all inside DDIM sampling (an optional
final all-gather of the latents).  The reference gets the same from Lightning's DDP strategy (main_val.py:656-660,
LR scaling :834-838).

Everything here is backend-agnostic torch.distributed (NCCL over NVLink on the B200 box, gloo in the CPU tests).
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """Join the process group torchrun describes (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*); single process when the
    variables are absent.  Returns (rank, local_rank, world_size)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, local, world


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) of the samples rank owns when n samples are split as evenly as possible (first n % world ranks get
    one extra).  Training uses equal shards (n % world == 0) so that the mean of shard-mean gradients is the batch mean."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} of {world}")
    base, extra = divmod(n, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def scaled_lr(base_lr: float, world: int, batch_per_gpu: int, accumulate: int = 1) -> float:
    """main_val.py:834-838: lr = accumulate_grad_batches * ngpu * batch_size * base_lr (``scale_lr: True``)."""
    return accumulate * world * batch_per_gpu * base_lr


def allreduce_mean_(flat_grad: torch.Tensor, group=None, prescaled: bool = False) -> torch.Tensor:
    """Sum the flat gradient bucket over the ranks, in place.  The 1/world factor is folded into the optimiser kernel
    (``encdiff_adamw_step_dev``'s grad_scale) on the CUDA path; ``prescaled=False`` applies it here instead."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return flat_grad
    dist.all_reduce(flat_grad, group=group)
    if not prescaled:
        flat_grad.mul_(1.0 / dist.get_world_size(group))
    return flat_grad


def broadcast_(flat_params: torch.Tensor, src: int = 0, group=None) -> torch.Tensor:
    """Rank ``src``'s parameters to every replica (what DDP does at construction)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(flat_params, src=src, group=group)
    return flat_params


def gather_samples(local: torch.Tensor, group=None) -> torch.Tensor:
    """Concatenate equally sized per-rank sample tensors in rank order (the only collective of the sampling path)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    out = [torch.empty_like(local) for _ in range(dist.get_world_size(group))]
    dist.all_gather(out, local.contiguous(), group=group)
    return torch.cat(out, 0)


def max_over_ranks(value: float, device=None, group=None) -> float:
    """Device-timed milliseconds -> max over ranks (bench.py's timing rule)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
