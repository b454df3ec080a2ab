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
