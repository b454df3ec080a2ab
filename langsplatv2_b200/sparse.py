"""Sparse-coefficient producer: softmax -> top-k -> renormalise as ONE kernel
(replaces the eight elementwise/topk launches of /root/reference/utils/vq_utils.py:9-40)."""
from __future__ import annotations

import torch

from . import _abi


def softmax_to_topk_soft_code(logits: torch.Tensor, k: int) -> torch.Tensor:
    """utils/vq_utils.py:9-24 (forward only): dense [N, C] with k non-zeros per row."""
    L_ = _abi.lib()
    _abi.require_cuda(logits, "logits")
    x = logits.detach().contiguous().float()
    N, Cc = x.shape
    dense = torch.empty_like(x)
    _abi.check(L_.lsv2_topk_soft_code(N, Cc, k, x.data_ptr(), dense.data_ptr(), None, None,
                                      _abi.stream_ptr(x.device)), "lsv2_topk_soft_code")
    return dense


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
