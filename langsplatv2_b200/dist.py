"""View-sharded data parallelism over one 8xB200 box.

Rendering shards by camera view with no data-path collective: view ``v`` belongs to rank
``v % world``.  Training (feature stage, geometry frozen: /root/reference/scene/gaussian_model.py:238-243)
renders a different view on every rank and sums the gradients of the sparse-coefficient logits and the
codebooks with ONE all-reduce over a flat fp32 bucket per step.  SUM (not mean) is the anchor: a world of B
ranks is then the reference's single-GPU ``--accum_iter B`` (train.py:261-263,348 -- gradients accumulate,
the loss is not rescaled).

One process per GPU (torchrun); backend NCCL on the box, gloo in the CPU tests.
"""
from __future__ import annotations

import os
from typing import Iterable, List, Sequence

import torch
import torch.distributed as dist


def world_info():
    """(rank, world, local_rank) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_views(num_views: int, rank: int, world: int) -> List[int]:
    """Round-robin ownership of views; every view has exactly one owner."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, num_views, world))


def step_views(step: int, num_views: int, rank: int, world: int) -> int:
    """View rendered by ``rank`` at training step ``step``: consecutive ranks take consecutive views, so a
    world of B ranks covers the same views per step as ``--accum_iter B`` does in B iterations."""
    return (step * world + rank) % num_views


class GradBucket:
    """Flat fp32 bucket over a fixed list of parameters: grads are packed, all-reduced once (SUM) and
    unpacked in place.  The bucket is allocated once and reused every step."""

    def __init__(self, params: Sequence[torch.Tensor]):
        self.params = list(params)
        if not self.params:
            raise ValueError("GradBucket needs at least one parameter")
        dev = self.params[0].device
        self.sizes = [p.numel() for p in self.params]
        self.flat = torch.zeros(sum(self.sizes), dtype=torch.float32, device=dev)
        self.views = []
        o = 0
        for p, n in zip(self.params, self.sizes):
            self.views.append(self.flat[o:o + n].view(p.shape))
            o += n

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def pack(self) -> None:
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad)

    def all_reduce(self) -> None:
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)

    def unpack(self) -> None:
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                p.grad = v.clone()
            else:
                p.grad.copy_(v)

    def sync(self) -> None:
        if self.attached:
            self.all_reduce()                       # the gradients already live in the bucket
            return
        self.pack()
        self.all_reduce()
        self.unpack()

    attached = False

    def attach(self) -> None:
        """Make every parameter's ``.grad`` a view of the flat bucket: backward accumulates straight into it and
        ``sync()`` is the all-reduce alone (no pack / unpack copies).  Pair with ``zero()`` instead of
        ``optimizer.zero_grad(set_to_none=True)``."""
        for p, v in zip(self.params, self.views):
            p.grad = v
        self.attached = True

    def zero(self) -> None:
        self.flat.zero_()

    # ---- all-reduce overlapped with the producer of the gradient (attached buckets) ----
    def begin_overlap(self) -> None:
        self._works = []
        if self.flat.is_cuda and getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=self.flat.device)

    def reduce_slice(self, a: int, b: int) -> None:
        """Flat elements [a, b) are final on the current stream: start their all-reduce(SUM) now, on a side stream, so
        it runs under whatever the current stream does next."""
        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1) or b <= a:
            return
        if not self.flat.is_cuda:
            dist.all_reduce(self.flat[a:b], op=dist.ReduceOp.SUM)
            return
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(self._side):
            self._side.wait_event(ev)
            self._works.append(dist.all_reduce(self.flat[a:b], op=dist.ReduceOp.SUM, async_op=True))

    def wait_slice(self, i: int) -> None:
        """Make the current stream wait for the i-th ``reduce_slice`` of this step (no-op on CPU: it was synchronous)."""
        if i < len(self._works) and self._works[i] is not None:
            self._works[i].wait()
            self._works[i] = None

    def finish_overlap(self, done_upto: int) -> None:
        """Reduce what ``reduce_slice`` has not covered ([done_upto, numel)) and make the current stream wait for
        every slice."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 and done_upto < self.flat.numel():
            dist.all_reduce(self.flat[done_upto:], op=dist.ReduceOp.SUM)
        for w in getattr(self, "_works", []):
            if w is not None:
                w.wait()
        self._works = []


def all_reduce_grads(params: Iterable[torch.Tensor]) -> None:
    """One-shot helper (allocates a bucket); training loops should hold a GradBucket instead."""
    GradBucket(list(params)).sync()
