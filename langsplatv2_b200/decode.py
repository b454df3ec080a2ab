"""Global semantic-codebook decode.

One call replaces the body of the six duplicated helpers of the reference
(/root/reference/eval_lerf.py:210-220, eval_3d_ovs.py:276-286, eval_mip_nerf360.py:168-178,
train.py:38-58, backend_renderer.py:16-36, demo_prompt.py:20-41) and
``GaussianModel.compute_*_feature_map`` (scene/gaussian_model.py:520-549):

    F_l[D, H, W] = C_l^T . W_l ;  optionally F_l /= (||F_l||_2 over D + 1e-10)
"""
from __future__ import annotations

import torch

from . import _abi

_PREC = {"fp32": _abi.DECODE_FP32, "bf16": _abi.DECODE_BF16}


def _decode_raw(wmap: torch.Tensor, codebooks: torch.Tensor, normalize: bool, precision: str,
                out_dtype: torch.dtype) -> torch.Tensor:
    L_ = _abi.lib()
    _abi.require_cuda(wmap, "weight_map")
    levels, Lc, D = (int(s) for s in codebooks.shape)
    if wmap.shape[0] != levels * Lc:
        raise ValueError(f"weight_map has {wmap.shape[0]} channels, codebooks need {levels}x{Lc}")
    pix_shape = tuple(wmap.shape[1:])
    HW = 1
    for s in pix_shape:
        HW *= int(s)
    w = wmap.contiguous().float()
    cb = codebooks.to(wmap.device).contiguous().float()
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("out_dtype must be float32 or bfloat16")
    out = torch.empty((levels, D) + pix_shape, dtype=out_dtype, device=wmap.device)
    nbytes = L_.lsv2_decode_scratch_bytes(levels, Lc, D)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=wmap.device)
    _abi.check(L_.lsv2_decode(levels, Lc, D, HW, w.data_ptr(), cb.data_ptr(), 1 if normalize else 0,
                              _PREC[precision], 0 if out_dtype == torch.float32 else 1, out.data_ptr(),
                              scratch.data_ptr(), nbytes, _abi.stream_ptr(wmap.device)), "lsv2_decode")
    return out


class _Decode(torch.autograd.Function):
    @staticmethod
    def forward(ctx, wmap, codebooks, precision):
        ctx.save_for_backward(wmap, codebooks)
        return _decode_raw(wmap, codebooks, False, precision, torch.float32)

    @staticmethod
    def backward(ctx, dF):
        L_ = _abi.lib()
        wmap, codebooks = ctx.saved_tensors
        levels, Lc, D = (int(s) for s in codebooks.shape)
        HW = wmap[0].numel()
        w = wmap.contiguous().float()
        cb = codebooks.contiguous().float()
        dF = dF.contiguous().float()
        dW = torch.empty_like(w) if ctx.needs_input_grad[0] else None
        dC = torch.zeros_like(cb) if ctx.needs_input_grad[1] else None
        _abi.check(L_.lsv2_decode_backward(levels, Lc, D, HW, w.data_ptr(), cb.data_ptr(), dF.data_ptr(),
                                           _abi.ptr(dW), _abi.ptr(dC), _abi.stream_ptr(w.device)),
                   "lsv2_decode_backward")
        return dW, dC, None


@_abi.guarded(lambda weight_map, *a, **k: weight_map.device)
def decode_feature_map(weight_map: torch.Tensor, codebooks: torch.Tensor, normalize: bool = True,
                       precision: str = "bf16", out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """weight_map [levels*Lc, H, W] (the op's 2nd return value), codebooks [levels, Lc, D]
    -> [levels, D, H, W].  ``precision='bf16'`` runs the contraction on tcgen05 tensor cores with
    bf16 operands and fp32 accumulation; ``'fp32'`` is the CUDA-core fp32-FMA variant."""
    if codebooks.dim() == 2:
        codebooks = codebooks.unsqueeze(0)
    needs_grad = torch.is_grad_enabled() and (weight_map.requires_grad or codebooks.requires_grad)
    if not needs_grad:
        return _decode_raw(weight_map, codebooks, normalize, precision, out_dtype)
    # training: gradients flow through the un-normalised product (scene/gaussian_model.py:538);
    # normalisation, when asked for, is the caller's torch expression (train.py:157-159)
    F = _Decode.apply(weight_map, codebooks, "fp32")
    if normalize:
        F = F / (F.norm(dim=1, keepdim=True) + 1e-10)
    return F.to(out_dtype)
