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


SPARSE_CAP, PACKED_WORDS = 16, 24       # include/lsv2.h: entries per compacted row, words per packed row


def _topk_dense_rows(x: torch.Tensor, k: int):
    """Dense rows + the compacted / packed rows of the same coefficients, one kernel (``lsv2_topk_soft_code_sparse``)."""
    L_ = _abi.lib()
    N, Cc = x.shape
    if Cc != 64 or not 0 < k <= SPARSE_CAP:
        raise _abi.Lsv2Error("the fused row producer handles 64 codes per row and k <= 16")
    dev = x.device
    dense = torch.empty_like(x)
    rows = dict(sparse_count=torch.empty(N, dtype=torch.int32, device=dev),
                sparse_index=torch.empty((N, SPARSE_CAP), dtype=torch.int16, device=dev),
                sparse_weight=torch.empty((N, SPARSE_CAP), dtype=torch.float32, device=dev),
                packed=torch.empty((N, PACKED_WORDS), dtype=torch.int32, device=dev))
    _abi.check(L_.lsv2_topk_soft_code_sparse(N, k, x.data_ptr(), dense.data_ptr(), rows["sparse_count"].data_ptr(),
                                             rows["sparse_index"].data_ptr(), rows["sparse_weight"].data_ptr(),
                                             rows["packed"].data_ptr(), _abi.stream_ptr(dev)),
               "lsv2_topk_soft_code_sparse")
    return dense, rows


class GradSink:
    """Where ``_TopkSoftCode.backward`` writes d logits when a trainer wants the gradient chunk by chunk (N > 1 ranks:
    the all-reduce of chunk i runs under the kernel of chunk i + 1, ``dist.GradBucket.reduce_slice``).  ``target`` is
    the [N, C] view of the flat gradient bucket that already IS ``logits.grad``; ``on_chunk(a, b)`` is called with the
    flat element range that has just been written on the current stream."""

    def __init__(self, target: torch.Tensor, row_blocks, on_chunk):
        self.target, self.row_blocks, self.on_chunk = target, list(row_blocks), on_chunk
        self.used_upto = 0


_GRAD_SINK = None


def set_grad_sink(sink) -> None:
    global _GRAD_SINK
    _GRAD_SINK = sink


class _TopkSoftCode(torch.autograd.Function):
    """One kernel forward, one kernel backward (the reference's autograd graph holds ~6 [N, C] temporaries)."""

    rows = None                # side channel forward -> softmax_to_topk_soft_code_rows (never referenced from ctx)

    @staticmethod
    def forward(ctx, logits, k, want_rows=False):
        x = logits.contiguous().float()
        if want_rows:
            dense, _TopkSoftCode.rows = _topk_dense_rows(x, k)
        else:
            dense = _topk_dense(x, k)
        ctx.save_for_backward(x, dense)
        return dense

    @staticmethod
    def backward(ctx, g):
        L_ = _abi.lib()
        x, dense = ctx.saved_tensors
        g = g.contiguous().float()
        sink = _GRAD_SINK
        if sink is not None and sink.target.shape == x.shape and sink.target.is_contiguous():
            N, C = x.shape
            for a, b in sink.row_blocks:
                _abi.check(L_.lsv2_topk_soft_code_backward(b - a, C, x[a:].data_ptr(), dense[a:].data_ptr(),
                                                           g[a:].data_ptr(), sink.target[a:].data_ptr(),
                                                           _abi.stream_ptr(x.device)), "lsv2_topk_soft_code_backward")
                sink.on_chunk(a * C, b * C)
            sink.used_upto = N * C
            return None, None, None      # the gradient already sits in logits.grad (the bucket view)
        dx = torch.empty_like(x)
        _abi.check(L_.lsv2_topk_soft_code_backward(x.shape[0], x.shape[1], x.data_ptr(), dense.data_ptr(),
                                                   g.data_ptr(), dx.data_ptr(), _abi.stream_ptr(x.device)),
                   "lsv2_topk_soft_code_backward")
        return dx, None, None


def softmax_to_topk_soft_code(logits: torch.Tensor, k: int) -> torch.Tensor:
    """utils/vq_utils.py:9-24: dense [N, C] with k non-zeros per row; differentiable w.r.t. logits
    (the gradient flows through the kept probabilities and the renormalising sum, as in the reference's
    ``torch.where`` / division graph)."""
    _abi.require_cuda(logits, "logits")
    if torch.is_grad_enabled() and logits.requires_grad:
        return _TopkSoftCode.apply(logits, k)
    return _topk_dense(logits.detach().contiguous().float(), k)


def softmax_to_topk_soft_code_rows(logits: torch.Tensor, k: int):
    """``softmax_to_topk_soft_code`` + what the rasterizer would derive from its result: returns ``(dense, rows)``,
    ``rows`` being the ``sparse_rows=`` argument of ``rasterize_forward`` / ``GaussianRasterizer.sparse_rows`` for a
    frame whose ``language_feature_precomp`` is ``dense``.  64 codes per row (one level), k <= 16."""
    _abi.require_cuda(logits, "logits")
    if torch.is_grad_enabled() and logits.requires_grad:
        dense = _TopkSoftCode.apply(logits, k, True)
        rows, _TopkSoftCode.rows = _TopkSoftCode.rows, None
        return dense, rows
    return _topk_dense_rows(logits.detach().contiguous().float(), k)


def topk_backward_rows(logits: torch.Tensor, dense: torch.Tensor, rows: dict, d_rows: torch.Tensor,
                       out: torch.Tensor, row_range=None) -> torch.Tensor:
    """d logits from the coefficient gradient in row form (``rasterize_backward(rows_out=)``): ``out[r]`` for the rows
    ``row_range = (a, b)`` (default all) of ``logits`` / ``dense`` / ``rows``; ``d_rows`` and ``out`` hold exactly those
    rows ([b - a, stride] and [b - a, 64]).  One kernel (``lsv2_topk_soft_code_backward_rows``)."""
    L_ = _abi.lib()
    N, Cc = logits.shape
    a, b = (0, N) if row_range is None else (int(row_range[0]), int(row_range[1]))
    n = b - a
    if Cc != 64 or not (0 <= a <= b <= N):
        raise ValueError("64 codes per row, 0 <= a <= b <= N")
    if d_rows.shape[0] != n or d_rows.dim() != 2 or out.shape != (n, Cc) or not d_rows.is_contiguous() or \
            not out.is_contiguous() or not logits.is_contiguous() or not dense.is_contiguous():
        raise ValueError("d_rows must be [b - a, stride], out [b - a, 64], all contiguous")
    if n == 0:
        return out
    x, y = logits.detach(), dense.detach()
    _abi.check(L_.lsv2_topk_soft_code_backward_rows(
        n, x[a:].data_ptr(), y[a:].data_ptr(), rows["sparse_count"][a:].data_ptr(),
        rows["sparse_index"][a:].data_ptr(), d_rows.data_ptr(), int(d_rows.shape[1]), out.data_ptr(),
        _abi.stream_ptr(x.device)), "lsv2_topk_soft_code_backward_rows")
    return out


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
