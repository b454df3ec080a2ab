"""TEST INFRASTRUCTURE -- CPU restatement of ``simple_knn._C.distCUDA2`` (mean squared distance to the three nearest
neighbours).  Only tests/, ``__graft_entry__.smoke()`` and bench.py's baseline legs may import this module.

Follows /root/reference/submodules/simple-knn/simple_knn.cu:
  * :135-146  ``updateKBest<3>``: d = q - p per axis in fp32, ``d.x*d.x + d.y*d.y + d.z*d.z`` -- which nvcc (no
    fast-math, setup.py:30) contracts to ``fma(dz, dz, fma(dx, dx, dy*dy))`` (checked in the SASS of that expression for
    sm_100a); the three smallest values are kept in ascending order, strict ``>`` on insertion;
  * :148-183  ``boxMeanDist``: every other point is a candidate (the Morton boxes only prune, :170-172), the result is
    ``(best[0] + best[1] + best[2]) / 3.0f`` with FLT_MAX left in unused slots (two other points -> FLT_MAX / 3, fewer -> +inf).

Parity status: PINNED on a GPU box -- the reference's own ``SimpleKNN::knn`` compiles from its source where it lies
(``make -C oracle knn_ref`` -> oracle/_ref/libsimple_knn_ref.so, CUDA, so it runs on the GPU box only);
tests/golden/knn_golden.npz holds its outputs (tests/golden/make_knn_golden.py, run there), and this restatement is
checked against them on CPU.  The fp32 fma is emulated through float64 (a*b is exact there; the sum is rounded twice,
which can differ from a true fma in the last bit in rare cases: the tests allow 1 ulp and count exact matches).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

FLT_MAX = np.float32(np.finfo(np.float32).max)


def _fma32(a: np.ndarray, b: np.ndarray, c: np.ndarray) -> np.ndarray:
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def dist2(q: np.ndarray, p: np.ndarray) -> np.ndarray:
    """simple_knn.cu:138-139 as compiled: fma(dz, dz, fma(dx, dx, dy * dy)); q, p broadcastable [..., 3] fp32."""
    d = (p.astype(np.float32) - q.astype(np.float32)).astype(np.float32)
    t = (d[..., 1] * d[..., 1]).astype(np.float32)
    t = _fma32(d[..., 0], d[..., 0], t)
    return _fma32(d[..., 2], d[..., 2], t)


def _mean3(best: np.ndarray) -> np.ndarray:
    """simple_knn.cu:182: ascending sum, then / 3.0f (FLT_MAX + FLT_MAX overflows to +inf as in fp32)."""
    with np.errstate(over="ignore"):
        s = (best[:, 0] + best[:, 1]).astype(np.float32)
        s = (s + best[:, 2]).astype(np.float32)
        return (s / np.float32(3.0)).astype(np.float32)


def dist_cuda2_bruteforce(points: np.ndarray, block: int = 512) -> np.ndarray:
    """All pairs; for P up to a few 10^4."""
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    P = pts.shape[0]
    best = np.full((P, 3), FLT_MAX, dtype=np.float32)
    for a in range(0, P, block):
        q = pts[a:a + block]
        d = dist2(q[:, None, :], pts[None, :, :])                     # [b, P]
        d[np.arange(q.shape[0]), np.arange(a, a + q.shape[0])] = np.inf     # j != i (duplicates of i stay)
        k = min(3, P - 1)
        if k > 0:
            part = np.partition(d, k - 1, axis=1)[:, :k]
            part.sort(axis=1)
            best[a:a + block, :k] = part
    return _mean3(best)


def dist_cuda2(points: np.ndarray) -> np.ndarray:
    """Exact 3-NN through a k-d tree (float64 search over the fp32 coordinates, 8 candidates per point re-evaluated in
    the reference's fp32 arithmetic); for sizes the brute force cannot reach."""
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    P = pts.shape[0]
    if P <= 4096:
        return dist_cuda2_bruteforce(pts)
    from scipy.spatial import cKDTree
    k = min(8, P)
    _, nn = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=k, workers=-1)
    d = dist2(pts[:, None, :], pts[nn])                               # [P, k]
    own = nn == np.arange(P)[:, None]
    # drop the point itself once; with > k coincident points it may not be among the candidates: all d are 0 then
    first_own = own & (np.cumsum(own, axis=1) == 1)
    d = np.where(first_own, np.float32(np.inf), d)
    d.sort(axis=1)
    return _mean3(np.ascontiguousarray(d[:, :3], dtype=np.float32))


# ---- the reference itself (CUDA; GPU box only) -------------------------------------------------------------------
_REF = None


def reference_lib():
    """oracle/_ref/libsimple_knn_ref.so (``ref_simple_knn(P, points_dev, out_dev)``: the reference's SimpleKNN::knn +
    a device synchronise), or None when it was not built."""
    global _REF
    if _REF is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libsimple_knn_ref.so")
        if not os.path.exists(path):
            return None
        lib = ctypes.CDLL(path)
        lib.ref_simple_knn.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib.ref_simple_knn.restype = ctypes.c_int
        _REF = lib
    return _REF


def reference_dist_cuda2(points_cuda):
    """The reference's distCUDA2 (spatial.cu:15-25) on a CUDA tensor [P, 3] fp32; synchronises the device."""
    import torch
    lib = reference_lib()
    if lib is None:
        raise RuntimeError("oracle/_ref/libsimple_knn_ref.so is not built (make -C oracle knn_ref)")
    pts = points_cuda.contiguous().float()
    out = torch.zeros(pts.shape[0], dtype=torch.float32, device=pts.device)
    if pts.shape[0] == 0:
        return out
    torch.cuda.synchronize(pts.device)
    with torch.cuda.device(pts.device):
        rc = lib.ref_simple_knn(int(pts.shape[0]), pts.data_ptr(), out.data_ptr())
    if rc != 0:
        raise RuntimeError(f"reference SimpleKNN::knn failed: cudaError {rc}")
    return out
