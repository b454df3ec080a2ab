"""Sparse-coefficient producer: softmax -> top-k -> renormalise as ONE kernel
(replaces the eight elementwise/topk launches of /root/reference/utils/vq_utils.py:9-40)."""
from __future__ import annotations

import torch

from . import _abi


def _topk_dense(x: torch.Tensor, k: int) -> torch.Tensor:
    L_ = _abi.lib()
    N, Cc = x.shape
    dense = torch.empty_like(x)
    _abi.check(L_.lsv2_topk_soft_code(N, Cc, k, x.data_ptr(), dense.data_ptr(), None, None,
                                      _abi.stream_ptr(x.device)), "lsv2_topk_soft_code")
    return dense


class _TopkSoftCode(torch.autograd.Function):
    """One kernel forward, one kernel backward (the reference's autograd graph holds ~6 [N, C] temporaries)."""

    @staticmethod
    def forward(ctx, logits, k):
        x = logits.contiguous().float()
        dense = _topk_dense(x, k)
        ctx.save_for_backward(x, dense)
        return dense

    @staticmethod
    def backward(ctx, g):
        L_ = _abi.lib()
        x, dense = ctx.saved_tensors
        g = g.contiguous().float()
        dx = torch.empty_like(x)
        _abi.check(L_.lsv2_topk_soft_code_backward(x.shape[0], x.shape[1], x.data_ptr(), dense.data_ptr(),
                                                   g.data_ptr(), dx.data_ptr(), _abi.stream_ptr(x.device)),
                   "lsv2_topk_soft_code_backward")
        return dx, None


def softmax_to_topk_soft_code(logits: torch.Tensor, k: int) -> torch.Tensor:
    """utils/vq_utils.py:9-24: dense [N, C] with k non-zeros per row; differentiable w.r.t. logits
    (the gradient flows through the kept probabilities and the renormalising sum, as in the reference's
    ``torch.where`` / division graph)."""
    _abi.require_cuda(logits, "logits")
    if torch.is_grad_enabled() and logits.requires_grad:
        return _TopkSoftCode.apply(logits, k)
    return _topk_dense(logits.detach().contiguous().float(), k)


def get_weights_and_indices(logits: torch.Tensor, k: int):
    """utils/vq_utils.py:26-40: (weights [N,k], indices [N,k] as float, ascending channel order)."""
    L_ = _abi.lib()
    _abi.require_cuda(logits, "logits")
    x = logits.detach().contiguous().float()
    N, Cc = x.shape
    w = torch.empty((N, k), dtype=torch.float32, device=x.device)
    i = torch.empty((N, k), dtype=torch.float32, device=x.device)
    _abi.check(L_.lsv2_topk_soft_code(N, Cc, k, x.data_ptr(), None, w.data_ptr(), i.data_ptr(),
                                      _abi.stream_ptr(x.device)), "lsv2_topk_soft_code")
    return w, i
