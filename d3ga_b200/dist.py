"""Camera-sharded data parallelism for multi-view training steps.

The reference has no distributed code at all (single process, one random camera per step: train.py:426-434); views
are independent given replicated parameters, so the only exchange a multi-view step needs is ONE all-reduce (sum) of
the hot-path parameter gradients -- Gaussian barys/scales/rotation/opacity/colour plus the cage vertices -- over a
single flat fp32 buffer.  One process per GPU, torch.distributed for the plumbing (NCCL on GPUs over NVLink/NVSwitch,
gloo in the CPU tests).  Rasterization itself never communicates.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None):
    """Join the process group torchrun describes (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT); no-op for 1 rank."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, local, world


def shard_cameras(n_cameras: int, rank: int, world: int):
    """Camera c belongs to rank c mod world (deterministic, balanced to within one camera)."""
    return list(range(rank, n_cameras, world))


class FlatGrads:
    """One flat fp32 buffer whose slices ARE the parameters' .grad tensors, so the all-reduce needs no packing."""

    def __init__(self, params):
        self.params = list(params)
        n = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        self.flat = torch.zeros(n, dtype=torch.float32, device=p0.device)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero_(self):
        self.flat.zero_()

    def nbytes(self):
        return self.flat.numel() * 4

    def all_reduce(self, async_op: bool = False):
        """Sum over ranks (the reference's global clip_grad_norm_, models/trainer.py:188, must run after this)."""
        if dist.is_initialized() and dist.get_world_size() > 1:
            return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)
        return None


def max_over_ranks(value: float, device) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return value


def sum_over_ranks(value: float, device) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())
    return value


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
