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

    def sum_lanes_and_ranks(self, lane_flats, chunks: int = 1):
        """self.flat = sum over `lane_flats` (the per-lane gradient buffers of training.PipelinedViews), then summed over
        ranks.  With chunks > 1 the two steps are pipelined: while the collective of chunk k runs on a communication
        stream, the lanes' chunk k + 1 is being added on the caller's stream.  MEASURED on 8 B200 (metric workload, same
        box): 1 chunk 14.10 ms per step, 4 chunks 14.31, 8 chunks 14.52 -- the 60 MB all-reduce over NVSwitch is not what
        separates 8 GPUs from one (13.48 ms); the extra collectives' launches cost more than their overlap hides.  Hence
        the default of one chunk = the plain add followed by one all-reduce; D3GA_REDUCE_CHUNKS overrides it.
        The caller's stream waits for the last collective (no host synchronisation).
        The global clip_grad_norm_ of models/trainer.py:188 must still come after this."""
        multi = dist.is_initialized() and dist.get_world_size() > 1
        n = self.flat.numel()
        chunks = int(os.environ.get("D3GA_REDUCE_CHUNKS", chunks))
        k = max(1, int(chunks)) if (multi and n >= 1 << 16) else 1
        edges = [((n * i // k) + 255) // 256 * 256 if 0 < i < k else (0 if i == 0 else n) for i in range(k + 1)]
        cuda = self.flat.is_cuda
        if cuda and multi and getattr(self, "_comm", None) is None:
            self._comm = torch.cuda.Stream(device=self.flat.device)
        works = []
        for i in range(k):
            a, b = edges[i], edges[i + 1]
            if b <= a:
                continue
            dst = self.flat[a:b]
            if len(lane_flats) == 1:
                dst.copy_(lane_flats[0][a:b])
            else:
                torch.add(lane_flats[0][a:b], lane_flats[1][a:b], out=dst)
                for f in lane_flats[2:]:
                    dst.add_(f[a:b])
            if not multi:
                continue
            if cuda:
                ev = torch.cuda.Event()
                ev.record()
                with torch.cuda.stream(self._comm):
                    self._comm.wait_event(ev)
                    works.append(dist.all_reduce(dst, op=dist.ReduceOp.SUM, async_op=True))
            else:
                works.append(dist.all_reduce(dst, op=dist.ReduceOp.SUM, async_op=True))
        for w in works:
            w.wait()          # CUDA: makes the current stream wait for the collective; gloo: blocks


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
