"""``distCUDA2``: mean squared distance of every point to its three nearest neighbours (scale initialisation of a scene,
/root/reference/scene/gaussian_model.py:194; reference: submodules/simple-knn/spatial.cu:15-25, simple_knn.cu:148-220).
One C-ABI call (``lsv2_dist2_knn3``, csrc/knn.cu), no host synchronisation."""
from __future__ import annotations

import torch

from . import _abi


@_abi.guarded(lambda points, *a, **k: points.device)
def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """points: [P, 3] CUDA tensor -> float32 [P].  As in the reference, fewer than three other points leave FLT_MAX
    in the sum: FLT_MAX / 3 or +inf (callers clamp from below only, gaussian_model.py:194)."""
    _abi.require_cuda(points, "points")
    if points.dim() != 2 or points.shape[1] != 3:
        raise ValueError("points must be [P, 3]")
    L_ = _abi.lib()
    pts = points.detach().contiguous().float()
    P = int(pts.shape[0])
    out = torch.zeros(P, dtype=torch.float32, device=pts.device)            # spatial.cu:21: full({P}, 0.0)
    if P == 0:
        return out
    scratch = torch.empty(L_.lsv2_knn_scratch_bytes(P), dtype=torch.uint8, device=pts.device)
    _abi.check(L_.lsv2_dist2_knn3(P, pts.data_ptr(), out.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                  _abi.stream_ptr(pts.device)), "lsv2_dist2_knn3")
    return out
