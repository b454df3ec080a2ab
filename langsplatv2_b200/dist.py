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


class ShardedExchange:
    """Gradient exchange of a flat buffer in blocks with the optimizer step sharded over the ranks (ZeRO-1 shape):

        reduce-scatter(block) -> this rank steps ITS 1/world of the block -> all-gather(the block's parameters)

    moves the same bytes over NVLink as the all-reduce it replaces, but every rank runs 1/world of the optimizer
    traffic, and each parameter is computed by exactly one rank (bitwise identical replicas by construction).
    SUM semantics as ``GradBucket`` (a world of B ranks == the reference's ``--accum_iter B``, train.py:261-263,348).
    ``blocks``: [(a, b)] flat element ranges, each a multiple of 4 * world long, covering the buffers."""

    def __init__(self, flat_g: torch.Tensor, flat_p: torch.Tensor, blocks: Sequence):
        self.g, self.p = flat_g, flat_p
        self.blocks = [(int(a), int(b)) for a, b in blocks]
        self.rank, self.world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
        for a, b in self.blocks:
            if (b - a) % (4 * self.world) or a % 4:
                raise ValueError(f"block [{a}, {b}) is not a multiple of 4 * world elements")
        self.shard_grad = [torch.empty((b - a) // self.world, dtype=torch.float32, device=flat_g.device)
                           for a, b in self.blocks]
        self._side = torch.cuda.Stream(device=flat_g.device) if flat_g.is_cuda else None
        self._rs = [None] * len(self.blocks)
        self._ag = [None] * len(self.blocks)

    def shard_range(self, i: int):
        a, b = self.blocks[i]
        n = (b - a) // self.world
        return a + self.rank * n, a + (self.rank + 1) * n

    def reduce_scatter(self, i: int) -> None:
        """Block i's gradient is final on the current stream: start its reduce-scatter (side stream)."""
        a, b = self.blocks[i]
        if self.world == 1:
            self.shard_grad[i].copy_(self.g[a:b])
            return
        if self._side is None:
            dist.reduce_scatter_tensor(self.shard_grad[i], self.g[a:b], op=dist.ReduceOp.SUM)
            return
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(self._side):
            self._side.wait_event(ev)
            self._rs[i] = dist.reduce_scatter_tensor(self.shard_grad[i], self.g[a:b], op=dist.ReduceOp.SUM, async_op=True)

    def wait_grad(self, i: int) -> None:
        if self._rs[i] is not None:
            self._rs[i].wait()
            self._rs[i] = None

    def all_gather_params(self, i: int) -> None:
        """This rank's shard of block i's parameters is final on the current stream: gather the block (side stream)."""
        if self.world == 1:
            return
        a, b = self.blocks[i]
        lo, hi = self.shard_range(i)
        if self._side is None:
            dist.all_gather_into_tensor(self.p[a:b], self.p[lo:hi].clone())
            return
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(self._side):
            self._side.wait_event(ev)
            self._ag[i] = dist.all_gather_into_tensor(self.p[a:b], self.p[lo:hi], async_op=True)

    def all_gather_all(self) -> None:
        """Deferred form: every block's shard is final on the current stream (the previous iteration's Adam); start all
        the gathers on the side stream.  The caller runs work that does not read the parameters (the next iteration's
        preprocess + binning) and then calls ``finish()``."""
        for i in range(len(self.blocks)):
            self.all_gather_params(i)

    def finish(self) -> None:
        """The current stream waits for every outstanding gather (call before the next forward)."""
        for i, w in enumerate(self._ag):
            if w is not None:
                w.wait()
                self._ag[i] = None

    def full_reduced_grad(self) -> torch.Tensor:
        """The summed gradient of the last exchange, reassembled from every rank's shards (tests only)."""
        out = torch.empty_like(self.g)
        for i, (a, b) in enumerate(self.blocks):
            if self.world == 1:
                out[a:b].copy_(self.shard_grad[i])
            else:
                dist.all_gather_into_tensor(out[a:b], self.shard_grad[i])
        return out


def all_reduce_grads(params: Iterable[torch.Tensor]) -> None:
    """One-shot helper (allocates a bucket); training loops should hold a GradBucket instead."""
    GradBucket(list(params)).sync()
