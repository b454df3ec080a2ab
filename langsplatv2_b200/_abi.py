"""ctypes binding of liblsv2.so (C ABI in include/lsv2.h).

The product path has no CPU fallback: if the shared library is missing or a call fails, this
module raises.  Nothing here imports ``oracle``.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# LSV2_LIB: developer hook to load an instrumented build (e.g. liblsv2_prof.so, csrc/Makefile target `prof`)
LIB_PATH = os.environ.get("LSV2_LIB") or os.path.join(_HERE, "liblsv2.so")

LSV2_ABI_VERSION = 2
TILE = 16
SPARSE_CAP = 16
PACKED_WORDS = 24
STATUS_WORDS = 8
STATUS_NUM_RENDERED, STATUS_OVERFLOW, STATUS_BAD_INDEX = 0, 1, 2
MODE_RGB, MODE_DENSE, MODE_QUICK = 0, 1, 2
DECODE_FP32, DECODE_BF16 = 0, 1
OUT_HWC = 2

_vp = C.c_void_p


class View(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("tanfovx", C.c_float), ("tanfovy", C.c_float),
                ("scale_modifier", C.c_float), ("sh_degree", C.c_int32),
                ("viewmatrix", _vp), ("projmatrix", _vp), ("campos", _vp), ("bg", _vp)]


class Scene(C.Structure):
    _fields_ = [("P", C.c_int32), ("sh_coeffs", C.c_int32), ("means3d", _vp), ("opacities", _vp),
                ("scales", _vp), ("rotations", _vp), ("cov3d_precomp", _vp), ("shs", _vp),
                ("colors_precomp", _vp)]


class Geom(C.Structure):
    _fields_ = [("radii", _vp), ("xy", _vp), ("depth", _vp), ("conic_opacity", _vp), ("rgb", _vp),
                ("clamped", _vp), ("tiles_touched", _vp), ("offsets", _vp), ("cov3d", _vp), ("rect", _vp), ("blend_rec", _vp)]


class Binning(C.Structure):
    _fields_ = [("capacity", C.c_uint64), ("keys", _vp), ("point_list", _vp), ("keys_alt", _vp),
                ("vals_alt", _vp), ("ranges", _vp), ("scratch", _vp), ("scratch_bytes", C.c_size_t),
                ("status", _vp)]


class Features(C.Structure):
    _fields_ = [("mode", C.c_int32), ("L", C.c_int32), ("K", C.c_int32), ("coeff_dense", _vp),
                ("weights_quick", _vp), ("indices_quick", _vp), ("sparse_count", _vp),
                ("sparse_index", _vp), ("sparse_weight", _vp), ("packed", _vp)]


class Image(C.Structure):
    _fields_ = [("color", _vp), ("wmap", _vp), ("final_T", _vp), ("n_contrib", _vp)]


class GeomGrads(C.Structure):
    _fields_ = [("means3d", _vp), ("means2d", _vp), ("opacities", _vp), ("scales", _vp), ("rotations", _vp),
                ("cov3d", _vp), ("shs", _vp), ("colors", _vp)]


GEOM_ACC = 12


EXPORTS = [
    "lsv2_abi_version", "lsv2_strerror", "lsv2_last_cuda_error", "lsv2_sort_scratch_bytes",
    "lsv2_preprocess", "lsv2_bin", "lsv2_compact_coeff", "lsv2_render_forward", "lsv2_forward",
    "lsv2_mark_visible", "lsv2_render_backward", "lsv2_decode_scratch_bytes", "lsv2_decode",
    "lsv2_decode_backward", "lsv2_topk_soft_code", "lsv2_topk_soft_code_backward", "lsv2_cos_loss_blocks", "lsv2_cos_loss_forward", "lsv2_cos_loss_backward",
    "lsv2_gram_loss_forward", "lsv2_gram_loss_backward", "lsv2_colsumsq", "lsv2_segment_sum",
    "lsv2_gram_loss_forward_seg", "lsv2_gram_loss_backward_seg",
    "lsv2_render_decode_scratch_bytes", "lsv2_render_decode", "lsv2_render_forward_tc", "lsv2_adam_step", "lsv2_pack_coeff",
    "lsv2_render_query_scratch_bytes", "lsv2_render_query",
    "lsv2_render_backward_rgb", "lsv2_preprocess_backward", "lsv2_decode_static", "lsv2_topk_soft_code_sparse", "lsv2_knn_scratch_bytes", "lsv2_dist2_knn3",
    "lsv2_render_backward_rows", "lsv2_topk_soft_code_backward_rows",
]

_lib = None


class Lsv2Error(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load liblsv2.so; raises if it has not been built (``python -c 'import __graft_entry__ as g; g.build()'``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Lsv2Error(f"{LIB_PATH} is missing: the CUDA extension has not been built "
                        "(run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.lsv2_abi_version.restype = C.c_int
    if L.lsv2_abi_version() != LSV2_ABI_VERSION:
        raise Lsv2Error("liblsv2.so ABI version mismatch; rebuild")
    L.lsv2_strerror.restype = C.c_char_p
    L.lsv2_strerror.argtypes = [C.c_int]
    L.lsv2_last_cuda_error.restype = C.c_char_p
    L.lsv2_sort_scratch_bytes.restype = C.c_size_t
    L.lsv2_sort_scratch_bytes.argtypes = [C.c_uint64, C.c_int32]
    P = C.POINTER
    L.lsv2_preprocess.argtypes = [P(View), P(Scene), P(Geom), P(Binning), _vp]
    L.lsv2_bin.argtypes = [P(View), C.c_int32, P(Geom), P(Binning), _vp]
    L.lsv2_compact_coeff.argtypes = [C.c_int32, P(Geom), P(Features), _vp]
    L.lsv2_render_forward.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp]
    L.lsv2_render_query_scratch_bytes.restype = C.c_size_t
    L.lsv2_render_query_scratch_bytes.argtypes = [C.c_int32]
    L.lsv2_render_query.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp, C.c_int32, C.c_int32,
                                    C.c_int32, _vp, C.c_int32, C.c_int32, _vp, _vp, C.c_size_t, _vp]
    L.lsv2_pack_coeff.argtypes = [C.c_int32, P(Geom), P(Features), _vp, _vp]
    L.lsv2_render_forward_tc.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp]
    L.lsv2_forward.argtypes = [P(View), P(Scene), P(Geom), P(Binning), P(Features), P(Image), _vp]
    L.lsv2_mark_visible.argtypes = [C.c_int32, _vp, _vp, _vp, _vp]
    L.lsv2_render_backward.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp, _vp,
                                       C.c_int32, _vp, _vp, _vp]
    L.lsv2_render_backward_rows.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp, _vp,
                                            C.c_int32, _vp, _vp, _vp]
    L.lsv2_render_backward_rows.restype = C.c_int
    L.lsv2_topk_soft_code_backward_rows.argtypes = [C.c_int64, _vp, _vp, _vp, _vp, _vp, C.c_int32, _vp, _vp]
    L.lsv2_topk_soft_code_backward_rows.restype = C.c_int
    L.lsv2_render_backward_rgb.argtypes = [P(View), P(Geom), P(Binning), P(Image), _vp, _vp, _vp]
    L.lsv2_preprocess_backward.argtypes = [P(View), P(Scene), P(Geom), _vp, P(GeomGrads), _vp]
    L.lsv2_decode_scratch_bytes.restype = C.c_size_t
    L.lsv2_decode_scratch_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    L.lsv2_decode.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, _vp, _vp, C.c_int32, C.c_int32,
                              C.c_int32, _vp, _vp, C.c_size_t, _vp]
    L.lsv2_decode_static.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, _vp, _vp, C.c_int32, C.c_int32, _vp, _vp]
    L.lsv2_decode_static.restype = C.c_int
    L.lsv2_decode_backward.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_topk_soft_code.argtypes = [C.c_int64, C.c_int32, C.c_int32, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_knn_scratch_bytes.argtypes = [C.c_int32]
    L.lsv2_knn_scratch_bytes.restype = C.c_size_t
    L.lsv2_dist2_knn3.argtypes = [C.c_int32, _vp, _vp, _vp, C.c_size_t, _vp]
    L.lsv2_dist2_knn3.restype = C.c_int
    L.lsv2_topk_soft_code_sparse.argtypes = [C.c_int64, C.c_int32, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_topk_soft_code_sparse.restype = C.c_int
    L.lsv2_topk_soft_code_backward.argtypes = [C.c_int64, C.c_int32, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_cos_loss_blocks.restype = C.c_int32
    L.lsv2_cos_loss_blocks.argtypes = [C.c_int64]
    L.lsv2_cos_loss_forward.argtypes = [C.c_int32, C.c_int64, _vp, _vp, _vp, C.c_int32, _vp, _vp, _vp]
    L.lsv2_cos_loss_backward.argtypes = [C.c_int32, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_gram_loss_forward.argtypes = [C.c_int32, C.c_int64, _vp, _vp, _vp, _vp, _vp, C.c_int32, _vp, _vp, _vp, _vp]
    L.lsv2_gram_loss_backward.argtypes = [C.c_int32, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
    L.lsv2_colsumsq.argtypes = [C.c_int32, C.c_int64, _vp, _vp, _vp]
    L.lsv2_segment_sum.argtypes = [C.c_int32, C.c_int64, C.c_int32, _vp, _vp, _vp, _vp]
    L.lsv2_gram_loss_forward_seg.argtypes = [C.c_int32, C.c_int64, C.c_int32, _vp, _vp, _vp, _vp, _vp, C.c_int32, _vp, _vp,
                                             _vp, _vp]
    L.lsv2_gram_loss_backward_seg.argtypes = [C.c_int32, C.c_int64, C.c_int32] + [_vp] * 10
    L.lsv2_render_decode_scratch_bytes.restype = C.c_size_t
    L.lsv2_render_decode_scratch_bytes.argtypes = [C.c_int32]
    L.lsv2_render_decode.argtypes = [P(View), P(Geom), P(Binning), P(Features), P(Image), _vp, C.c_int32,
                                     C.c_int32, C.c_int32, C.c_int32, C.c_int32, _vp, _vp, C.c_size_t, _vp]
    L.lsv2_adam_step.argtypes = [_vp, _vp, _vp, _vp, C.c_int64, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64,
                                 C.c_int32, _vp, _vp]
    for name in ("lsv2_preprocess", "lsv2_bin", "lsv2_compact_coeff", "lsv2_render_forward", "lsv2_forward",
                 "lsv2_mark_visible", "lsv2_render_backward", "lsv2_render_backward_rgb", "lsv2_preprocess_backward",
                 "lsv2_decode", "lsv2_decode_backward", "lsv2_topk_soft_code", "lsv2_topk_soft_code_backward", "lsv2_cos_loss_forward", "lsv2_cos_loss_backward",
                 "lsv2_gram_loss_forward", "lsv2_gram_loss_backward", "lsv2_colsumsq", "lsv2_segment_sum", "lsv2_render_decode",
                 "lsv2_render_forward_tc", "lsv2_gram_loss_forward_seg", "lsv2_gram_loss_backward_seg", "lsv2_adam_step", "lsv2_pack_coeff", "lsv2_render_query"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(code: int, what: str) -> None:
    if code != 0:
        L = lib()
        msg = L.lsv2_strerror(code).decode()
        if code == 2:
            msg += ": " + L.lsv2_last_cuda_error().decode()
        raise Lsv2Error(f"{what} failed: {msg}")


def ptr(t) -> int | None:
    """Device pointer of a contiguous CUDA tensor (None passes NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def guarded(get_device):
    """Decorator: run the call with the tensors' CUDA device current (liblsv2 launches on the CURRENT device; a caller
    holding several GPUs in one process need not have switched)."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*a, **k):
            dev = get_device(*a, **k)
            if dev is not None and dev.type == "cuda" and torch.cuda.is_available() and \
                    torch.cuda.current_device() != (dev.index if dev.index is not None else torch.cuda.current_device()):
                with torch.cuda.device(dev):
                    return fn(*a, **k)
            return fn(*a, **k)
        return wrapper
    return deco


def device_guard(device):
    """Context manager making ``device`` the current CUDA device for the launches inside it: the kernels,
    cudaFuncSetAttribute and the tensor-map encode of liblsv2 act on the CURRENT device, which need not be the
    one the tensors live on."""
    return torch.cuda.device(device)


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise Lsv2Error(f"{name} must be a CUDA tensor: langsplatv2_b200 has no CPU path")
