"""Feature-stage loss as two kernels (train.py:157-164 + utils/loss_utils.py:24-25): see csrc/loss.cu."""
from __future__ import annotations

from typing import Optional

import torch

from . import _abi


class _CosLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, gt, mask, normalize):
        L_ = _abi.lib()
        _abi.require_cuda(feat, "language_feature")
        D = int(feat.shape[0])
        HW = feat[0].numel()
        f = feat.contiguous().float()
        g = gt.contiguous().float()
        if g.shape != f.shape:
            raise ValueError("gt must have the feature map's shape")
        m = None
        if mask is not None:
            m = mask.to(torch.float32).contiguous()
            if m.numel() != HW:
                raise ValueError("mask must be [1, H, W] (one value per pixel)")
        ab = torch.empty((2, HW), dtype=torch.float32, device=f.device)
        partial = torch.empty(L_.lsv2_cos_loss_blocks(HW), dtype=torch.float32, device=f.device)
        _abi.check(L_.lsv2_cos_loss_forward(D, HW, f.data_ptr(), g.data_ptr(), _abi.ptr(m), 1 if normalize else 0,
                                            ab.data_ptr(), partial.data_ptr(), _abi.stream_ptr(f.device)),
                   "lsv2_cos_loss_forward")
        ctx.save_for_backward(f, g, ab)
        return 1.0 - partial.sum() / HW

    @staticmethod
    def backward(ctx, grad_out):
        L_ = _abi.lib()
        f, g, ab = ctx.saved_tensors
        D, HW = int(f.shape[0]), f[0].numel()
        go = grad_out.reshape(1).contiguous().float()
        dF = torch.empty_like(f)
        _abi.check(L_.lsv2_cos_loss_backward(D, HW, f.data_ptr(), g.data_ptr(), ab.data_ptr(), go.data_ptr(),
                                             dF.data_ptr(), _abi.stream_ptr(f.device)), "lsv2_cos_loss_backward")
        return dF, None, None, None


def normalized_cos_loss(language_feature: torch.Tensor, gt: torch.Tensor, mask: Optional[torch.Tensor] = None,
                        normalize: bool = True) -> torch.Tensor:
    """``cos_loss(f * mask, gt * mask)`` with ``f = language_feature / (norm(dim=0) + 1e-10)`` when ``normalize``
    (train.py:157-164); language_feature, gt: [D, H, W]; mask: [1, H, W] or None."""
    return _CosLoss.apply(language_feature, gt, mask, normalize)


def _decode_bwd(levels, Lc, D, HW, wmap, codebooks, dF, dW, dC, dev):
    L_ = _abi.lib()
    _abi.check(L_.lsv2_decode_backward(levels, Lc, D, HW, _abi.ptr(wmap), _abi.ptr(codebooks), dF.data_ptr(),
                                       _abi.ptr(dW), _abi.ptr(dC), _abi.stream_ptr(dev)), "lsv2_decode_backward")


def column_sumsq(x: torch.Tensor) -> torch.Tensor:
    """|x_p|^2 per pixel of a [D, ...] map (one streaming kernel)."""
    L_ = _abi.lib()
    x = x.contiguous().float()
    HW = x[0].numel()
    out = torch.empty(HW, dtype=torch.float32, device=x.device)
    _abi.check(L_.lsv2_colsumsq(int(x.shape[0]), HW, x.data_ptr(), out.data_ptr(), _abi.stream_ptr(x.device)),
               "lsv2_colsumsq")
    return out


class CodebookGradSink:
    """Takes the codebook gradient of ``decode_cos_loss`` / ``decode_cos_loss_segments`` off autograd's critical path.

    The loss's backward produces two independent things: dL/dWmap (which the blend backward and the top-k backward
    are waiting for) and dL/dC (two pixel-axis GEMMs, ~0.65 ms at 1008x756, which nothing waits for until the
    optimizer).  With a sink installed the backward only records what the GEMMs need; ``compute()`` -- called by the
    trainer after ``loss.backward()`` has enqueued the rest of the backward pass -- runs them on a side stream and ADDS
    the result into ``target`` (the codebook's slot of the flat gradient buffer); ``join()`` makes the current stream
    wait for it.  At N > 1 the logits' reduce-scatter / sharded Adam / all-gather then run under the GEMMs instead of
    after them.  Autograd sees ``None`` for the codebook (no double accumulation)."""

    def __init__(self, target: torch.Tensor):
        self.target = target
        self.stream = torch.cuda.Stream(device=target.device)
        self.fork = torch.cuda.Event()
        self.done = torch.cuda.Event()
        self.pending = None

    def defer(self, fn, keep) -> None:
        if self.pending is not None:
            raise RuntimeError("CodebookGradSink: a second loss backward before compute()/join()")
        self.fork.record(torch.cuda.current_stream(self.target.device))
        self.pending = (fn, keep)           # `keep`: the operands stay referenced until the join

    def compute(self) -> None:
        if self.pending is None or self.pending[0] is None:
            return
        fn, keep = self.pending
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.fork)
            dC = fn()
            self.target.add_(dC.reshape(self.target.shape))
            keep.append(dC)
            self.done.record(self.stream)
        self.pending = (None, keep)

    def join(self) -> None:
        if self.pending is None:
            return
        if self.pending[0] is not None:
            self.compute()
        # the operands were allocated on the current stream and are released only now, after it has been ordered behind
        # their last use on the side stream; the GEMMs' own temporaries never leave the side stream
        torch.cuda.current_stream(self.target.device).wait_event(self.done)
        self.pending = None


_cb_sink = None


def set_codebook_grad_sink(sink) -> None:
    global _cb_sink
    _cb_sink = sink


class _GramCosLoss(torch.autograd.Function):
    """decode one codebook layer + normalise + cos loss without forming the [512, H, W] feature map
    (csrc/loss.cu, second half): two 64x512xHW GEMMs per iteration instead of three, no F / dF tensors."""

    @staticmethod
    def forward(ctx, wmap, codebook, gt, mask, normalize, gt_sumsq):
        L_ = _abi.lib()
        _abi.require_cuda(wmap, "language_feature_weight_map")
        dev = wmap.device
        w = wmap.contiguous().float()
        Lc, HW = int(w.shape[0]), w[0].numel()
        Cb = codebook.contiguous().float().reshape(Lc, -1)
        D = int(Cb.shape[1])
        g = gt.contiguous().float()
        if g.shape[0] != D or g[0].numel() != HW:
            raise ValueError("gt must be [D, H, W] with the weight map's H, W")
        m = None
        if mask is not None:
            m = mask.to(torch.float32).contiguous()
            if m.numel() != HW:
                raise ValueError("mask must be [1, H, W] (one value per pixel)")
        g2 = gt_sumsq if gt_sumsq is not None else column_sumsq(g)
        allow = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        G = (Cb @ Cb.t()).contiguous()
        torch.backends.cuda.matmul.allow_tf32 = allow
        h = torch.empty_like(w)
        _decode_bwd(1, Lc, D, HW, None, Cb, g, h, None, dev)                       # h = C gt
        u = torch.empty_like(w)
        ab = torch.empty((2, HW), dtype=torch.float32, device=dev)
        partial = torch.empty(L_.lsv2_cos_loss_blocks(HW), dtype=torch.float32, device=dev)
        _abi.check(L_.lsv2_gram_loss_forward(Lc, HW, w.data_ptr(), h.data_ptr(), g2.data_ptr(), _abi.ptr(m),
                                             G.data_ptr(), 1 if normalize else 0, u.data_ptr(), ab.data_ptr(),
                                             partial.data_ptr(), _abi.stream_ptr(dev)), "lsv2_gram_loss_forward")
        ctx.save_for_backward(w, h, u, ab, g, Cb)
        ctx.cb_shape = codebook.shape
        return 1.0 - partial.sum() / HW

    @staticmethod
    def backward(ctx, grad_out):
        L_ = _abi.lib()
        w, h, u, ab, g, Cb = ctx.saved_tensors
        dev = w.device
        Lc, HW, D = int(w.shape[0]), w[0].numel(), int(Cb.shape[1])
        go = grad_out.reshape(1).contiguous().float()
        dw, aw, bw = torch.empty_like(w), torch.empty_like(w), torch.empty_like(w)
        _abi.check(L_.lsv2_gram_loss_backward(Lc, HW, w.data_ptr(), h.data_ptr(), u.data_ptr(), ab.data_ptr(),
                                              go.data_ptr(), dw.data_ptr(), aw.data_ptr(), bw.data_ptr(),
                                              _abi.stream_ptr(dev)), "lsv2_gram_loss_backward")
        dC = None
        if ctx.needs_input_grad[1]:
            cb_shape = ctx.cb_shape

            def codebook_grad():
                dC_ = torch.zeros_like(Cb)
                _decode_bwd(1, Lc, D, HW, aw, None, g, None, dC_, dev)             # (alpha w) gt^T
                M = torch.zeros((Lc, Lc), dtype=torch.float32, device=dev)
                _decode_bwd(1, Lc, Lc, HW, bw, None, w, None, M, dev)              # (beta w) w^T
                allow = torch.backends.cuda.matmul.allow_tf32
                torch.backends.cuda.matmul.allow_tf32 = False
                out = (dC_ + M @ Cb).reshape(cb_shape)
                torch.backends.cuda.matmul.allow_tf32 = allow
                return out

            if _cb_sink is not None:
                _cb_sink.defer(codebook_grad, [aw, bw, w, g, Cb])
            else:
                dC = codebook_grad()
        return (dw if ctx.needs_input_grad[0] else None), dC, None, None, None, None


class _GramCosLossSeg(torch.autograd.Function):
    """``_GramCosLoss`` for the ground truth as the reference's data path holds it: a per-segment feature table gathered
    through a segmentation map (scene/cameras.py:59-96, ``point_feature = feature_map[seg]``; seg < 0 = unlabelled,
    which is also the mask).  With gt_p = E[seg_p] the two 64 x 512 x HW products collapse:
    h = C gt = (C E^T)[:, seg]  (a 64 x 512 x S product and a gather), and
    (alpha w) gt^T = (segment sums of alpha w) . E  (``lsv2_segment_sum`` + a 64 x S x 512 product)."""

    @staticmethod
    def forward(ctx, wmap, codebook, seg, table, normalize):
        L_ = _abi.lib()
        _abi.require_cuda(wmap, "language_feature_weight_map")
        dev = wmap.device
        w = wmap.contiguous().float()
        Lc, HW = int(w.shape[0]), w[0].numel()
        Cb = codebook.contiguous().float().reshape(Lc, -1)
        E = table.contiguous().float()
        sg = seg.reshape(-1).to(torch.int32).contiguous()
        if sg.numel() != HW or E.shape[1] != Cb.shape[1]:
            raise ValueError("seg must be [H, W] and table [S, D] with the codebook's D")
        S = int(E.shape[0])
        allow = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        G = (Cb @ Cb.t()).contiguous()
        h = (Cb @ E.t()).contiguous()                                # C table^T, [Lc, S]: h_p = h[:, seg_p]
        torch.backends.cuda.matmul.allow_tf32 = allow
        e2 = (E * E).sum(1).contiguous()
        u = torch.empty_like(w)
        ab = torch.empty((2, HW), dtype=torch.float32, device=dev)
        partial = torch.empty(L_.lsv2_cos_loss_blocks(HW), dtype=torch.float32, device=dev)
        _abi.check(L_.lsv2_gram_loss_forward_seg(Lc, HW, S, w.data_ptr(), h.data_ptr(), e2.data_ptr(), sg.data_ptr(),
                                                 G.data_ptr(), 1 if normalize else 0, u.data_ptr(), ab.data_ptr(),
                                                 partial.data_ptr(), _abi.stream_ptr(dev)),
                   "lsv2_gram_loss_forward_seg")
        ctx.save_for_backward(w, h, u, ab, sg, E, Cb)
        ctx.cb_shape = codebook.shape
        return 1.0 - partial.sum() / HW

    @staticmethod
    def backward(ctx, grad_out):
        L_ = _abi.lib()
        w, h, u, ab, sg, E, Cb = ctx.saved_tensors
        dev = w.device
        Lc, HW, S = int(w.shape[0]), w[0].numel(), int(E.shape[0])
        go = grad_out.reshape(1).contiguous().float()
        dw, aw, bw = torch.empty_like(w), torch.empty_like(w), torch.empty_like(w)
        _abi.check(L_.lsv2_gram_loss_backward_seg(Lc, HW, S, w.data_ptr(), h.data_ptr(), sg.data_ptr(), u.data_ptr(),
                                                  ab.data_ptr(), go.data_ptr(), dw.data_ptr(), aw.data_ptr(),
                                                  bw.data_ptr(), _abi.stream_ptr(dev)), "lsv2_gram_loss_backward_seg")
        dC = None
        if ctx.needs_input_grad[1]:
            cb_shape = ctx.cb_shape

            def codebook_grad():
                Sm = torch.zeros((Lc, S), dtype=torch.float32, device=dev)
                _abi.check(L_.lsv2_segment_sum(Lc, HW, S, aw.data_ptr(), sg.data_ptr(), Sm.data_ptr(),
                                               _abi.stream_ptr(dev)), "lsv2_segment_sum")
                M = torch.zeros((Lc, Lc), dtype=torch.float32, device=dev)
                _decode_bwd(1, Lc, Lc, HW, bw, None, w, None, M, dev)              # (beta w) w^T
                allow = torch.backends.cuda.matmul.allow_tf32
                torch.backends.cuda.matmul.allow_tf32 = False
                out = (Sm @ E + M @ Cb).reshape(cb_shape)
                torch.backends.cuda.matmul.allow_tf32 = allow
                return out

            if _cb_sink is not None:
                _cb_sink.defer(codebook_grad, [aw, bw, w, sg, E, Cb])
            else:
                dC = codebook_grad()
        return (dw if ctx.needs_input_grad[0] else None), dC, None, None, None


def decode_cos_loss_segments(weight_map: torch.Tensor, codebook: torch.Tensor, seg: torch.Tensor, table: torch.Tensor,
                             normalize: bool = True) -> torch.Tensor:
    """``decode_cos_loss`` with the ground truth given as (seg [H, W] int, table [S, 512]): gt = table[seg], mask =
    seg >= 0 -- the form the reference's loader has before it expands it (scene/cameras.py:59-96).  Same value and
    gradients as the dense call; neither 64 x 512 x HW product is computed."""
    return _GramCosLossSeg.apply(weight_map, codebook, seg, table, normalize)


def decode_cos_loss(weight_map: torch.Tensor, codebook: torch.Tensor, gt: torch.Tensor,
                    mask: Optional[torch.Tensor] = None, normalize: bool = True,
                    gt_sumsq: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``cos_loss(normalise(codebook^T @ weight_map) * mask, gt * mask)`` (train.py:156-164 with one codebook layer)
    without the feature map.  weight_map [64, H, W]; codebook [64, 512] or [1, 64, 512]; gt [512, H, W];
    ``gt_sumsq`` = ``column_sumsq(gt)`` may be passed when the same gt is reused across iterations."""
    return _GramCosLoss.apply(weight_map, codebook, gt, mask, normalize, gt_sumsq)
