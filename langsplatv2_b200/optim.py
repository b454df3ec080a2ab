"""Adam over flat fp32 ranges, one kernel per range (``lsv2_adam_step``, csrc/adam.cu).

Replaces ``torch.optim.Adam(l, lr=0.0, eps=1e-15)`` of the reference (scene/gaussian_model.py:255, stepped at
train.py:261-263).  Parameters, gradients and both moments live in four flat buffers with the same layout, so a
data-parallel trainer can step any sub-range (a row block, or this rank's shard after a reduce-scatter) and the
whole feature stage needs two launches per iteration instead of torch's multi-tensor kernels.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import _abi


class FlatAdam:
    """``p``, ``g`` are flat fp32 views of equal length; ``m``, ``v`` are allocated to match.  ``step_range(a, b)``
    applies one Adam step to elements [a, b); every range that is stepped together belongs to one step count
    (``begin_step()`` advances it once per iteration)."""

    def __init__(self, p: torch.Tensor, g: torch.Tensor, lr: float, betas=(0.9, 0.999), eps: float = 1e-15):
        if p.dim() != 1 or g.shape != p.shape or p.dtype != torch.float32 or g.dtype != torch.float32:
            raise ValueError("FlatAdam needs flat fp32 parameter / gradient buffers of equal length")
        _abi.require_cuda(p, "parameters")
        self.p, self.g = p, g
        self.m = torch.zeros_like(p)
        self.v = torch.zeros_like(p)
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.t = 0
        # CUDA-graph replays need the two step-dependent scalars on the device; they are written by two fill kernels
        # (value in the launch arguments, stream-ordered) OUTSIDE the graph.  Eager steps pass them by value.
        self.device_scalars = False
        self._scal_dev = torch.zeros(2, dtype=torch.float32, device=p.device)

    def scalars(self, t: int):
        b1, b2 = self.betas
        return self.lr / (1.0 - b1 ** t), math.sqrt(1.0 - b2 ** t)

    def begin_step(self) -> None:
        """Advance the step count (once per iteration, before the ranges are stepped; not inside a graph capture)."""
        self.t += 1
        if self.device_scalars:
            a, b = self.scalars(self.t)
            self._scal_dev[0:1].fill_(a)
            self._scal_dev[1:2].fill_(b)

    @_abi.guarded(lambda self, *a, **k: self.p.device)
    def step_range(self, a: int = 0, b: Optional[int] = None, zero_grad: bool = False,
                   grad: Optional[torch.Tensor] = None) -> None:
        """One Adam step on flat elements [a, b) (a multiple of 4).  ``grad``: take the gradient of that range from
        this tensor (length b - a, e.g. the output of a reduce-scatter) instead of ``g[a:b]``."""
        b = self.p.numel() if b is None else b
        if b <= a:
            return
        if a % 4:
            raise ValueError("range start must be a multiple of 4 elements (16-byte alignment)")
        gsrc = self.g[a:b] if grad is None else grad
        if gsrc.numel() != b - a:
            raise ValueError("gradient length does not match the range")
        L_ = _abi.lib()
        _abi.check(L_.lsv2_adam_step(self.p[a:].data_ptr(), gsrc.data_ptr(), self.m[a:].data_ptr(),
                                     self.v[a:].data_ptr(), b - a, self.lr, self.betas[0], self.betas[1], self.eps,
                                     max(self.t, 1), 1 if zero_grad else 0,
                                     self._scal_dev.data_ptr() if self.device_scalars else None,
                                     _abi.stream_ptr(self.p.device)), "lsv2_adam_step")

    def state_dict(self):
        return {"m": self.m, "v": self.v, "t": self.t, "lr": self.lr, "betas": self.betas, "eps": self.eps}

    def load_state_dict(self, sd) -> None:
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.t = int(sd["t"])
