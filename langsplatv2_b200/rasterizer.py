"""Host-side mirror of the reference's operator API for the rasterization hot path.

``GaussianRasterizationSettings`` / ``GaussianRasterizer`` keep the names, argument meaning and
error behaviour of the ``diff_gaussian_rasterization`` package that the reference imports at
/root/reference/gaussian_renderer/__init__.py:15 and calls at :37-54,108-119, so ``render()``,
``train.py`` and the ``eval_*.py`` scripts can use this module as a drop-in.  Everything below the
Python signatures is new: PyTorch only owns device memory and streams, the work is done by the
hand-written sm_100a kernels in liblsv2.so, reached through the C ABI (include/lsv2.h).

There is no CPU path.  A missing extension or a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _abi
from ._abi import (Binning, Features, Geom, Image, Scene, View, MODE_DENSE, MODE_QUICK, MODE_RGB,
                   PACKED_WORDS, SPARSE_CAP, STATUS_BAD_INDEX, STATUS_NUM_RENDERED, STATUS_OVERFLOW, STATUS_WORDS, TILE)

#: Gradient of ``language_feature_precomp`` at the operator boundary: strict (every channel of every touched
#: Gaussian, what the reference op returns) by default.  The support-only form (channels whose coefficient is
#: non-zero) is exact only when the caller masks the non-top-k channels afterwards, as the reference's
#: ``torch.where`` does (utils/vq_utils.py:20); ``FeatureTrainer`` opts into it per rasterizer
#: (``GaussianRasterizer.dense_coeff_grad = False``).
DENSE_COEFF_GRAD = True


class GaussianRasterizationSettings(NamedTuple):
    """Same fields, same order as the keyword construction at gaussian_renderer/__init__.py:37-52."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    include_feature: bool = False
    quick_render: bool = False


@dataclass
class FrameState:
    """Everything one forward leaves on the device (the reference's geom/binning/img buffers,
    as named tensors instead of opaque byte blobs)."""
    P: int
    H: int
    W: int
    mode: int
    L: int
    K: int
    view: View = None
    geom: dict = field(default_factory=dict)
    binning: dict = field(default_factory=dict)
    image: dict = field(default_factory=dict)
    feat: dict = field(default_factory=dict)
    status: torch.Tensor = None
    capacity: int = 0
    keep: list = field(default_factory=list)    # tensors the View struct points into
    scene: dict = field(default_factory=dict)   # the op's geometry / colour inputs as passed to lsv2_preprocess

    @property
    def num_rendered(self) -> int:
        return int(self.status[STATUS_NUM_RENDERED].item()) & 0xFFFFFFFF

    @property
    def overflowed(self) -> bool:
        return bool(self.status[STATUS_OVERFLOW].item())

    def for_backward(self, color_copy: bool = False) -> "FrameState":
        """Shallow copy without the forward's output tensors (see _RasterizeGaussians.forward).  ``color_copy``:
        the geometry backward needs the final colour of every pixel; it gets a private copy (3 H W floats), never
        the output tensor itself."""
        b = FrameState(P=self.P, H=self.H, W=self.W, mode=self.mode, L=self.L, K=self.K, view=self.view,
                       geom=self.geom, binning=self.binning, feat=self.feat, status=self.status,
                       capacity=self.capacity, keep=self.keep, scene=self.scene)
        b.image = {k: v for k, v in self.image.items() if k in ("final_T", "n_contrib")}
        if color_copy:
            b.image["color"] = self.image["color"].detach().clone()
        return b

    def c_scene(self) -> Scene:
        s = self.scene
        shs = s.get("shs")
        return Scene(self.P, 0 if shs is None else int(shs.shape[1]), _abi.ptr(s.get("means3D")),
                     _abi.ptr(s.get("opacities")), _abi.ptr(s.get("scales")), _abi.ptr(s.get("rotations")),
                     _abi.ptr(s.get("cov3D_precomp")), _abi.ptr(shs), _abi.ptr(s.get("colors_precomp")))

    def c_geom(self) -> Geom:
        g = self.geom
        return Geom(*[_abi.ptr(g.get(k)) for k in ("radii", "xy", "depth", "conic_opacity", "rgb", "clamped",
                                                   "tiles_touched", "offsets", "cov3d", "rect", "blend_rec")])

    def c_binning(self) -> Binning:
        b = self.binning
        sc = b.get("scratch")
        return Binning(self.capacity, _abi.ptr(b.get("keys")), _abi.ptr(b.get("point_list")),
                       _abi.ptr(b.get("keys_alt")), _abi.ptr(b.get("vals_alt")), _abi.ptr(b.get("ranges")),
                       _abi.ptr(sc), 0 if sc is None else sc.numel(), _abi.ptr(self.status))

    def c_features(self) -> Features:
        f = self.feat
        return Features(self.mode, self.L, self.K, _abi.ptr(f.get("coeff_dense")), _abi.ptr(f.get("weights_quick")),
                        _abi.ptr(f.get("indices_quick")), _abi.ptr(f.get("sparse_count")),
                        _abi.ptr(f.get("sparse_index")), _abi.ptr(f.get("sparse_weight")), _abi.ptr(f.get("packed")))

    def ensure_packed(self, geom_c=None) -> None:
        """The tensor-core blend reads the coefficients from the packed table of ``lsv2_pack_coeff`` (fp16 head /
        remainder pairs + channel bytes).  Quick mode: depends on the coefficients only -- pass ``packed=`` to
        ``rasterize_forward`` to reuse one across frames.  Dense mode: built from this frame's compacted rows."""
        if "packed" in self.feat:
            return
        L_ = _abi.lib()
        dev = self.status.device
        self.feat["packed"] = torch.empty((self.P, PACKED_WORDS), dtype=torch.int32, device=dev)
        geom_c = geom_c or self.c_geom()
        feat_c = self.c_features()
        _abi.check(L_.lsv2_pack_coeff(self.P, C.byref(geom_c), C.byref(feat_c), _abi.ptr(self.status),
                                      _abi.stream_ptr(dev)), "lsv2_pack_coeff")

    def c_image(self) -> Image:
        i = self.image
        return Image(_abi.ptr(i.get("color")), _abi.ptr(i.get("wmap")), _abi.ptr(i.get("final_T")),
                     _abi.ptr(i.get("n_contrib")))


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    _abi.require_cuda(t, name)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _make_view(rs: GaussianRasterizationSettings, dev, keep: list) -> View:
    vm = _f32c(rs.viewmatrix.to(dev), "viewmatrix")
    pm = _f32c(rs.projmatrix.to(dev), "projmatrix")
    cp = _f32c(rs.campos.to(dev), "campos")
    bg = _f32c(rs.bg.to(dev), "bg")
    if vm.numel() != 16 or pm.numel() != 16 or cp.numel() != 3 or bg.numel() != 3:
        raise ValueError("viewmatrix/projmatrix must be 4x4, campos/bg 3-vectors")
    keep += [vm, pm, cp, bg]
    return View(int(rs.image_width), int(rs.image_height), float(rs.tanfovx), float(rs.tanfovy),
                float(rs.scale_modifier), int(rs.sh_degree), vm.data_ptr(), pm.data_ptr(), cp.data_ptr(),
                bg.data_ptr())


def feature_mode(rs: GaussianRasterizationSettings) -> int:
    """Mode comes from the settings flags, never from tensor shapes: the reference passes 1-element
    float placeholders for the inputs a mode does not use (gaussian_renderer/__init__.py:93-103)."""
    if rs.quick_render:
        return MODE_QUICK
    if rs.include_feature:
        return MODE_DENSE
    return MODE_RGB


def rasterize_forward(rs: GaussianRasterizationSettings, means3D, opacities, shs=None, colors_precomp=None,
                      scales=None, rotations=None, cov3D_precomp=None, language_feature_precomp=None,
                      language_feature_weights_quick=None, language_feature_indices=None,
                      capacity: Optional[int] = None, debug_buffers: bool = False,
                      render: bool = True, blend: str = "auto", packed: Optional[torch.Tensor] = None,
                      prepared: Optional[FrameState] = None, sparse_rows: Optional[dict] = None) -> FrameState:
    """Preprocess -> bin -> blend.  ``capacity=None`` reads the instance count back between the two
    halves (one host sync, as the reference does); an integer capacity runs the whole frame with no
    host involvement and flags overflow in ``state.status``.

    ``blend``: ``"tc"`` accumulates the sparse coefficients on the tensor cores (``lsv2_render_forward_tc``: weight
    map within 1e-4 of its per-pixel norm, everything else bit-identical), ``"simt"`` is the CUDA-core kernel whose
    weight map reproduces the checker's fp32 operation order bit for bit, ``"auto"`` picks ``"tc"`` where the shape
    allows it (L a multiple of 64 up to 192, K <= 16).

    ``sparse_rows`` (extension, include_feature only): the compacted and packed rows of ``language_feature_precomp``
    as ``sparse.softmax_to_topk_soft_code_rows`` emits them with the dense rows; the frame then launches neither
    ``lsv2_compact_coeff`` nor ``lsv2_pack_coeff``."""
    L_ = _abi.lib()
    dev = means3D.device
    _abi.require_cuda(means3D, "means3D")
    if torch.cuda.current_device() != dev.index:
        # liblsv2 launches on the CURRENT device: make it the tensors' device for the duration of the call
        with _abi.device_guard(dev):
            return rasterize_forward(rs, means3D, opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                                     rotations=rotations, cov3D_precomp=cov3D_precomp,
                                     language_feature_precomp=language_feature_precomp,
                                     language_feature_weights_quick=language_feature_weights_quick,
                                     language_feature_indices=language_feature_indices, capacity=capacity,
                                     debug_buffers=debug_buffers, render=render, blend=blend, packed=packed,
                                     prepared=prepared, sparse_rows=sparse_rows)
    P = int(means3D.shape[0])
    H, W = int(rs.image_height), int(rs.image_width)
    mode = feature_mode(rs)
    means3D = _f32c(means3D, "means3D")
    opacities = _f32c(opacities, "opacities").reshape(-1)
    if opacities.numel() != P:
        raise ValueError("opacities must have one entry per Gaussian")

    sh_coeffs = 0
    if shs is not None:
        shs = _f32c(shs, "shs")
        if shs.dim() != 3 or shs.shape[0] != P or shs.shape[2] != 3:
            raise ValueError("shs must be [P, M, 3]")
        sh_coeffs = int(shs.shape[1])
    if colors_precomp is not None:
        colors_precomp = _f32c(colors_precomp, "colors_precomp")
    if scales is not None:
        scales = _f32c(scales, "scales")
    if rotations is not None:
        rotations = _f32c(rotations, "rotations")
    if cov3D_precomp is not None:
        cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp")

    L = K = 0
    feat = {}
    if mode == MODE_DENSE:
        if language_feature_precomp is None or language_feature_precomp.dim() != 2 or \
                language_feature_precomp.shape[0] != P:
            raise ValueError("include_feature needs language_feature_precomp of shape [P, L]")
        cd = _f32c(language_feature_precomp, "language_feature_precomp")
        L = int(cd.shape[1])
        feat["coeff_dense"] = cd
        if sparse_rows is not None:
            want = dict(sparse_count=((P,), torch.int32), sparse_index=((P, SPARSE_CAP), torch.int16),
                        sparse_weight=((P, SPARSE_CAP), torch.float32), packed=((P, PACKED_WORDS), torch.int32))
            for key, (shape, dt) in want.items():
                t = sparse_rows.get(key)
                if t is None or tuple(t.shape) != shape or t.dtype != dt or not t.is_contiguous() or t.device != dev:
                    raise ValueError(f"sparse_rows[{key!r}] must be a contiguous {dt} tensor of shape {shape} on {dev}")
                feat[key] = t
        else:
            feat["sparse_count"] = torch.empty(P, dtype=torch.int32, device=dev)
            feat["sparse_index"] = torch.empty((P, SPARSE_CAP), dtype=torch.int16, device=dev)
            feat["sparse_weight"] = torch.empty((P, SPARSE_CAP), dtype=torch.float32, device=dev)
    elif mode == MODE_QUICK:
        wq, iq = language_feature_weights_quick, language_feature_indices
        if wq is None or iq is None or wq.dim() != 2 or wq.shape != iq.shape or wq.shape[0] != P:
            raise ValueError("quick_render needs language_feature_weights_quick / language_feature_indices "
                             "of shape [P, K]")
        feat["weights_quick"] = _f32c(wq, "language_feature_weights_quick")
        feat["indices_quick"] = _f32c(iq, "language_feature_indices")
        K = int(wq.shape[1])
        # callers hard-wire 3 levels x 64 codes (eval_lerf.py:215); tolerate other level counts
        L = int(getattr(rs, "quick_channels", 0) or QUICK_CHANNELS)

    if sparse_rows is not None and mode != MODE_DENSE:
        raise ValueError("sparse_rows goes with include_feature (language_feature_precomp)")
    if packed is not None and mode == MODE_QUICK:
        if packed.shape != (P, PACKED_WORDS) or packed.dtype != torch.int32 or not packed.is_contiguous():
            raise ValueError("packed must be the [P, 24] int32 table of a previous frame's state.feat['packed']")
        feat["packed"] = packed
    st = FrameState(P=P, H=H, W=W, mode=mode, L=L, K=K)
    st.feat = feat
    if prepared is not None:
        if (prepared.P, prepared.H, prepared.W) != (P, H, W) or prepared.capacity is None or not prepared.binning:
            raise ValueError("prepared must be the state of rasterize_forward(..., render=False) for this view and scene")
        if not render:
            raise ValueError("prepared and render=False together do nothing")
        st.view, st.keep, st.scene = prepared.view, prepared.keep, prepared.scene
        st.geom, st.binning, st.status, st.capacity = prepared.geom, prepared.binning, prepared.status, prepared.capacity
        f32_ = dict(dtype=torch.float32, device=dev)
        st.image["color"] = torch.empty((3, H, W), **f32_)
        st.image["final_T"] = torch.empty((H, W), **f32_)
        st.image["n_contrib"] = torch.empty((H, W), dtype=torch.int32, device=dev)
        if mode != MODE_RGB:
            st.image["wmap"] = torch.empty((L, H, W), **f32_)
        if blend not in ("auto", "tc", "simt"):
            raise ValueError("blend must be 'auto', 'tc' or 'simt'")
        tc_ok_ = mode != MODE_RGB and L > 0 and L % 64 == 0 and L <= 192 and (mode == MODE_DENSE or 0 < K <= SPARSE_CAP)
        if blend == "tc" and not tc_ok_:
            raise _abi.Lsv2Error("blend='tc' needs L a multiple of 64 up to 192 and at most 16 entries per Gaussian")
        geom_c, bin_c = st.c_geom(), st.c_binning()
        feat_c, img_c = st.c_features(), st.c_image()
        stream_ = _abi.stream_ptr(dev)
        if mode == MODE_DENSE and sparse_rows is None:
            _abi.check(L_.lsv2_compact_coeff(P, C.byref(geom_c), C.byref(feat_c), stream_), "lsv2_compact_coeff")
        if tc_ok_ and blend != "simt":
            st.ensure_packed(geom_c)
            feat_c = st.c_features()
            _abi.check(L_.lsv2_render_forward_tc(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                                 C.byref(img_c), stream_), "lsv2_render_forward_tc")
        else:
            _abi.check(L_.lsv2_render_forward(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                              C.byref(img_c), stream_), "lsv2_render_forward")
        return st
    st.view = _make_view(rs, dev, st.keep)
    st.keep += [means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp]
    st.scene = dict(means3D=means3D, opacities=opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                    rotations=rotations, cov3D_precomp=cov3D_precomp)
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE

    i32 = dict(dtype=torch.int32, device=dev)
    f32 = dict(dtype=torch.float32, device=dev)
    g = st.geom
    g["radii"] = torch.empty(P, **i32)
    g["xy"] = torch.empty((P, 2), **f32)
    g["depth"] = torch.empty(P, **f32)
    g["conic_opacity"] = torch.empty((P, 4), **f32)
    g["rgb"] = torch.empty((P, 3), **f32)
    g["clamped"] = torch.empty((P, 3), dtype=torch.uint8, device=dev)
    g["tiles_touched"] = torch.empty(P, **i32)
    g["offsets"] = torch.empty(P, **i32)
    g["blend_rec"] = torch.empty((P, 12), **f32)
    if debug_buffers:
        g["cov3d"] = torch.empty((P, 6), **f32)
        g["rect"] = torch.empty((P, 4), **i32)
    st.status = torch.zeros(STATUS_WORDS, **i32)

    scene = Scene(P, sh_coeffs, _abi.ptr(means3D), _abi.ptr(opacities), _abi.ptr(scales), _abi.ptr(rotations),
                  _abi.ptr(cov3D_precomp), _abi.ptr(shs), _abi.ptr(colors_precomp))
    stream = _abi.stream_ptr(dev)

    img = st.image
    img["color"] = torch.empty((3, H, W), **f32)
    img["final_T"] = torch.empty((H, W), **f32)
    img["n_contrib"] = torch.empty((H, W), **i32)
    if mode != MODE_RGB and render:
        img["wmap"] = torch.empty((L, H, W), **f32)

    def alloc_binning(cap: int):
        b = st.binning
        st.capacity = cap
        n = max(cap, 1)
        b["keys"] = torch.empty(n, dtype=torch.int64, device=dev)
        b["point_list"] = torch.empty(n, **i32)
        b["keys_alt"] = torch.empty(n, dtype=torch.int64, device=dev)
        b["ranges"] = torch.empty((gx * gy, 2), **i32)
        b["scratch"] = torch.empty(L_.lsv2_sort_scratch_bytes(cap, P), dtype=torch.uint8, device=dev)

    if blend not in ("auto", "tc", "simt"):
        raise ValueError("blend must be 'auto', 'tc' or 'simt'")
    tc_ok = mode != MODE_RGB and L > 0 and L % 64 == 0 and L <= 192 and (mode == MODE_DENSE or 0 < K <= SPARSE_CAP)
    if blend == "tc" and not tc_ok:
        raise _abi.Lsv2Error("blend='tc' needs L a multiple of 64 up to 192 and at most 16 entries per Gaussian")
    use_tc = tc_ok and blend != "simt"

    def run_blend(geom_c, bin_c):
        feat_c, img_c = st.c_features(), st.c_image()
        if mode == MODE_DENSE and sparse_rows is None:
            _abi.check(L_.lsv2_compact_coeff(P, C.byref(geom_c), C.byref(feat_c), stream), "lsv2_compact_coeff")
        if use_tc:
            st.ensure_packed(geom_c)
            feat_c = st.c_features()
            _abi.check(L_.lsv2_render_forward_tc(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                                 C.byref(img_c), stream), "lsv2_render_forward_tc")
        else:
            _abi.check(L_.lsv2_render_forward(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                              C.byref(img_c), stream), "lsv2_render_forward")

    if capacity is None:
        st.binning["scratch"] = torch.empty(L_.lsv2_sort_scratch_bytes(0, P), dtype=torch.uint8, device=dev)
        geom_c, bin_c = st.c_geom(), st.c_binning()
        _abi.check(L_.lsv2_preprocess(C.byref(st.view), C.byref(scene), C.byref(geom_c), C.byref(bin_c), stream),
                   "lsv2_preprocess")
        R = st.num_rendered                                   # the one host sync of the frame
        alloc_binning(R)
        bin_c = st.c_binning()
        _abi.check(L_.lsv2_bin(C.byref(st.view), P, C.byref(geom_c), C.byref(bin_c), stream), "lsv2_bin")
        if render:
            run_blend(geom_c, bin_c)
    else:
        alloc_binning(int(capacity))
        geom_c, bin_c = st.c_geom(), st.c_binning()
        _abi.check(L_.lsv2_preprocess(C.byref(st.view), C.byref(scene), C.byref(geom_c), C.byref(bin_c), stream),
                   "lsv2_preprocess")
        _abi.check(L_.lsv2_bin(C.byref(st.view), P, C.byref(geom_c), C.byref(bin_c), stream), "lsv2_bin")
        if render:
            run_blend(geom_c, bin_c)
    if rs.debug:
        # the reference's debug mode: synchronise after the launches, and on failure dump the inputs before
        # re-raising (snapshot_fw.dump, [LINEAGE] diff_gaussian_rasterization/__init__.py)
        try:
            torch.cuda.synchronize(dev)
            if st.overflowed:
                raise _abi.Lsv2Error(f"instance capacity {st.capacity} too small for {st.num_rendered} instances")
            if mode == MODE_QUICK and int(st.status[STATUS_BAD_INDEX].item()):
                raise _abi.Lsv2Error(f"language_feature_indices outside [0, {L})")
        except Exception:
            _dump_snapshot("snapshot_fw.dump", dict(
                raster_settings=rs._asdict(), means3D=means3D, opacities=opacities, shs=shs,
                colors_precomp=colors_precomp, scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp,
                language_feature_precomp=language_feature_precomp,
                language_feature_weights_quick=language_feature_weights_quick,
                language_feature_indices=language_feature_indices))
            raise
    return st


def _dump_snapshot(path: str, items: dict) -> None:
    """torch.save of the op's inputs (CPU copies), as the reference's debug mode does before re-raising."""
    def cpu(v):
        if torch.is_tensor(v):
            try:
                return v.detach().cpu()
            except Exception:
                return None
        if isinstance(v, dict):
            return {k: cpu(x) for k, x in v.items()}
        return v
    try:
        torch.save({k: cpu(v) for k, v in items.items()}, path)
        print(f"\nAn error occured in the rasterizer. Writing {path} for debugging.")
    except Exception:
        pass


@_abi.guarded(lambda st, *a, **k: st.status.device)
def render_decode(st: FrameState, codebooks: torch.Tensor, normalize: bool = True,
                  out_dtype: torch.dtype = torch.float32, want_wmap: bool = False,
                  out_layout: str = "CHW") -> torch.Tensor:
    """Fused blend + codebook decode on a frame that has been preprocessed and binned
    (``rasterize_forward(..., render=False)``): -> ``[levels, 512, H, W]`` (``out_layout="HWC"``: ``[levels, H, W,
    512]``, the layout ``OpenCLIPNetwork.get_max_across_quick`` takes); fills ``st.image`` (RGB, final_T, n_contrib,
    and the weight map when ``want_wmap``).  The L-channel weight map never round-trips HBM."""
    L_ = _abi.lib()
    if st.mode == MODE_RGB:
        raise _abi.Lsv2Error("render_decode needs include_feature / quick_render settings")
    dev = st.status.device
    cb = _f32c(codebooks.to(dev), "codebooks")
    if cb.dim() == 2:
        cb = cb.unsqueeze(0)
    levels, Lc, D = (int(x) for x in cb.shape)
    if levels * Lc != st.L:
        raise ValueError(f"codebooks give {levels}x{Lc} channels, the frame has L={st.L}")
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("out_dtype must be float32 or bfloat16")
    if not want_wmap:
        st.image.pop("wmap", None)
    elif "wmap" not in st.image:
        st.image["wmap"] = torch.empty((st.L, st.H, st.W), dtype=torch.float32, device=dev)
    if out_layout not in ("CHW", "HWC"):
        raise ValueError("out_layout must be 'CHW' or 'HWC'")
    hwc = out_layout == "HWC"
    out = torch.empty((levels, st.H, st.W, D) if hwc else (levels, D, st.H, st.W), dtype=out_dtype, device=dev)
    nbytes = L_.lsv2_render_decode_scratch_bytes(levels)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    geom_c, bin_c, feat_c, img_c = st.c_geom(), st.c_binning(), st.c_features(), st.c_image()
    if st.mode == MODE_DENSE:
        _abi.check(L_.lsv2_compact_coeff(st.P, C.byref(geom_c), C.byref(feat_c), _abi.stream_ptr(dev)),
                   "lsv2_compact_coeff")
    st.ensure_packed(geom_c)
    feat_c = st.c_features()
    _abi.check(L_.lsv2_render_decode(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                     C.byref(img_c), cb.data_ptr(), levels, Lc, D, 1 if normalize else 0,
                                     (0 if out_dtype == torch.float32 else 1) | (_abi.OUT_HWC if hwc else 0),
                                     out.data_ptr(), scratch.data_ptr(), nbytes, _abi.stream_ptr(dev)),
               "lsv2_render_decode")
    st.keep += [cb, scratch]
    return out


@_abi.guarded(lambda st, *a, **k: st.status.device)
def render_query(st: FrameState, codebooks: torch.Tensor, phrase_embeds: torch.Tensor, n_pos: int) -> torch.Tensor:
    """Fused blend + relevancy query on a preprocessed and binned frame: ``phrase_embeds [P + N, 512]`` (positives
    first, as ``torch.cat([pos_embeds, neg_embeds])`` at eval/openclip_encoder.py:127) -> relevancy ``[levels, P, H, W]``
    = ``OpenCLIPNetwork.get_max_across_quick`` of the decoded map (:114-138), without ever forming that map."""
    L_ = _abi.lib()
    if st.mode == MODE_RGB:
        raise _abi.Lsv2Error("render_query needs include_feature / quick_render settings")
    dev = st.status.device
    cb = _f32c(codebooks.to(dev), "codebooks")
    if cb.dim() == 2:
        cb = cb.unsqueeze(0)
    levels, Lc, D = (int(x) for x in cb.shape)
    if levels * Lc != st.L:
        raise ValueError(f"codebooks give {levels}x{Lc} channels, the frame has L={st.L}")
    e = _f32c(phrase_embeds.to(dev), "phrase_embeds")
    if e.dim() != 2 or e.shape[1] != D or not (0 < n_pos < e.shape[0]):
        raise ValueError("phrase_embeds must be [P + N, D] with 0 < n_pos < P + N")
    n_neg = int(e.shape[0]) - int(n_pos)
    st.image.pop("wmap", None)
    out = torch.empty((levels, n_pos, st.H, st.W), dtype=torch.float32, device=dev)
    nbytes = L_.lsv2_render_query_scratch_bytes(levels)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    geom_c, bin_c, feat_c, img_c = st.c_geom(), st.c_binning(), st.c_features(), st.c_image()
    if st.mode == MODE_DENSE:
        _abi.check(L_.lsv2_compact_coeff(st.P, C.byref(geom_c), C.byref(feat_c), _abi.stream_ptr(dev)),
                   "lsv2_compact_coeff")
    st.ensure_packed(geom_c)
    feat_c = st.c_features()
    _abi.check(L_.lsv2_render_query(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                    C.byref(img_c), cb.data_ptr(), levels, Lc, D, e.data_ptr(), int(n_pos), n_neg,
                                    out.data_ptr(), scratch.data_ptr(), nbytes, _abi.stream_ptr(dev)),
               "lsv2_render_query")
    st.keep += [cb, e, scratch]
    return out


#: weight-map channels in quick mode: the callers view the result as (3, 64, H, W)
#: (eval_lerf.py:215, eval_3d_ovs.py:281, eval_mip_nerf360.py:173).
QUICK_CHANNELS = 192


@_abi.guarded(lambda st, *a, **k: st.status.device)
def rasterize_backward(st: FrameState, grad_wmap: Optional[torch.Tensor], grad_color: Optional[torch.Tensor] = None,
                       dense_grad: Optional[bool] = None, rows_out: Optional[torch.Tensor] = None):
    """dL/d language_feature_precomp [P, L] (and dL/d colours [P, 3] when grad_color is given).

    ``rows_out`` (extension): a zero-initialised float32 [P, stride] tensor; the gradient is ADDED to it in the order
    of the compacted rows (entry e of row g = channel ``sparse_index[g, e]``; ``lsv2_render_backward_rows``) and the
    first return value is None -- 16 B per Gaussian for top-4 instead of a 256-B dense row."""
    L_ = _abi.lib()
    if st.mode != MODE_DENSE:
        raise _abi.Lsv2Error("backward is defined for include_feature (dense coefficient) mode only")
    dev = st.status.device
    dense = DENSE_COEFF_GRAD if dense_grad is None else dense_grad
    gw = _f32c(grad_wmap, "grad_wmap")
    gc = _f32c(grad_color, "grad_color") if grad_color is not None else None
    d_colors = torch.zeros((st.P, 3), dtype=torch.float32, device=dev) if gc is not None else None
    geom_c, bin_c, feat_c, img_c = st.c_geom(), st.c_binning(), st.c_features(), st.c_image()
    if rows_out is not None:
        if rows_out.dim() != 2 or rows_out.shape[0] != st.P or rows_out.dtype != torch.float32 or \
                not rows_out.is_contiguous() or rows_out.device != dev:
            raise ValueError("rows_out must be a contiguous float32 [P, stride] tensor on the frame's device")
        _abi.check(L_.lsv2_render_backward_rows(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                                C.byref(img_c), _abi.ptr(gw), _abi.ptr(gc), int(rows_out.shape[1]),
                                                _abi.ptr(rows_out), _abi.ptr(d_colors), _abi.stream_ptr(dev)),
                   "lsv2_render_backward_rows")
        return None, d_colors
    d_coeff = torch.zeros((st.P, st.L), dtype=torch.float32, device=dev)
    _abi.check(L_.lsv2_render_backward(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(feat_c),
                                       C.byref(img_c), _abi.ptr(gw), _abi.ptr(gc), 1 if dense else 0,
                                       _abi.ptr(d_coeff), _abi.ptr(d_colors), _abi.stream_ptr(dev)),
               "lsv2_render_backward")
    return d_coeff, d_colors


def rasterize_backward_geometry(st: FrameState, grad_color: torch.Tensor, want: Optional[dict] = None) -> dict:
    """The RGB stage's backward (train.py:169-172 without --include_feature): dL/d(means3D, means2D, opacities,
    scales, rotations | cov3D_precomp, shs | colors_precomp) from dL/d rendered_image.  ``st.image['color']`` must
    hold the forward's rendered image.  Two launches: ``lsv2_render_backward_rgb`` (per-tile replay, per-Gaussian
    sums) and ``lsv2_preprocess_backward`` (per-Gaussian chain rule).  Returns a dict of the gradients, shaped like
    the inputs."""
    L_ = _abi.lib()
    dev = st.status.device
    if "color" not in st.image:
        raise _abi.Lsv2Error("the geometry backward needs the forward's rendered image in state.image['color']")
    gc = _f32c(grad_color, "grad_color")
    if gc.shape != (3, st.H, st.W):
        raise ValueError("grad_color must be [3, H, W]")
    sc = st.scene
    f32 = dict(dtype=torch.float32, device=dev)
    acc = torch.zeros((st.P, _abi.GEOM_ACC), **f32)
    out = {"means3D": torch.empty((st.P, 3), **f32), "means2D": torch.empty((st.P, 3), **f32),
           "opacities": torch.empty((st.P,), **f32)}
    if sc.get("cov3D_precomp") is not None:
        out["cov3D_precomp"] = torch.empty((st.P, 6), **f32)
    else:
        out["scales"] = torch.empty((st.P, 3), **f32)
        out["rotations"] = torch.empty((st.P, 4), **f32)
    if sc.get("shs") is not None:
        out["shs"] = torch.empty_like(sc["shs"])
    else:
        out["colors_precomp"] = torch.empty((st.P, 3), **f32)
    geom_c, bin_c, img_c, scene_c = st.c_geom(), st.c_binning(), st.c_image(), st.c_scene()
    stream = _abi.stream_ptr(dev)
    with _abi.device_guard(dev):
        _abi.check(L_.lsv2_render_backward_rgb(C.byref(st.view), C.byref(geom_c), C.byref(bin_c), C.byref(img_c),
                                               _abi.ptr(gc), _abi.ptr(acc), stream), "lsv2_render_backward_rgb")
        grads = _abi.GeomGrads(_abi.ptr(out["means3D"]), _abi.ptr(out["means2D"]), _abi.ptr(out["opacities"]),
                               _abi.ptr(out.get("scales")), _abi.ptr(out.get("rotations")),
                               _abi.ptr(out.get("cov3D_precomp")), _abi.ptr(out.get("shs")),
                               _abi.ptr(out.get("colors_precomp")))
        _abi.check(L_.lsv2_preprocess_backward(C.byref(st.view), C.byref(scene_c), C.byref(geom_c), _abi.ptr(acc),
                                               C.byref(grads), stream), "lsv2_preprocess_backward")
    return out


class _RasterizeGaussians(torch.autograd.Function):
    last_state = None          # side channel forward -> GaussianRasterizer.forward (never referenced from ctx)
    last_rows = None           # side channel backward -> caller: the sparse_rows dict whose "grad_rows" was just written

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, language_feature_precomp, weights_quick, indices_quick,
                opacities, scales, rotations, cov3Ds_precomp, raster_settings, capacity=None, dense_coeff_grad=None,
                prepared=None, sparse_rows=None):
        st = rasterize_forward(raster_settings, means3D, opacities, shs=sh, colors_precomp=colors_precomp,
                               scales=scales, rotations=rotations, cov3D_precomp=cov3Ds_precomp,
                               language_feature_precomp=language_feature_precomp,
                               language_feature_weights_quick=weights_quick, language_feature_indices=indices_quick,
                               capacity=capacity, prepared=prepared, sparse_rows=sparse_rows)
        # what backward needs, WITHOUT the forward's outputs: a state holding `color` / `wmap` on ctx closes the cycle
        # output -> grad_fn -> ctx -> state -> output through C++ shared_ptrs, which Python's GC cannot collect --
        # every grad-enabled forward would leak its whole frame
        # inputs: 0 means3D, 1 means2D, 2 sh, 3 colors_precomp, ..., 7 opacities, 8 scales, 9 rotations, 10 cov3D
        need = ctx.needs_input_grad
        # (means2D alone does not count outside the RGB stage: render() always passes it with requires_grad=True,
        # gaussian_renderer/__init__.py:27-31, and only the RGB stage's densification reads its gradient)
        ctx.geometry = bool(need[0] or need[2] or need[7] or need[8] or need[9] or need[10] or
                            ((need[1] or need[3]) and st.mode == MODE_RGB))
        ctx.state = st.for_backward(color_copy=ctx.geometry)
        ctx.opacity_shape = opacities.shape
        ctx.color_is_precomp = colors_precomp is not None
        ctx.dense_grad = dense_coeff_grad
        # extension: the coefficient gradient in row form, written into the caller's buffer (None flows back instead)
        ctx.grad_rows = sparse_rows.get("grad_rows") if sparse_rows is not None else None
        ctx.sparse_rows = sparse_rows if ctx.grad_rows is not None else None    # dies with the graph if never used
        ctx.debug = bool(raster_settings.debug)
        _RasterizeGaussians.last_state = st
        H, W = st.H, st.W
        wmap = st.image.get("wmap")
        if wmap is None:
            wmap = torch.zeros((0, H, W), dtype=torch.float32, device=means3D.device)
        radii = st.geom["radii"]
        ctx.mark_non_differentiable(radii)
        return st.image["color"], wmap, radii

    @staticmethod
    def backward(ctx, grad_color, grad_wmap, _grad_radii):
        st = ctx.state
        ctx.state = None                     # one backward per forward; drop the frame buffers now
        if st is None:
            raise RuntimeError("the rasterizer's backward ran twice (retain_graph is not supported: the frame "
                               "buffers are released after the first pass)")
        need = ctx.needs_input_grad
        # inputs: 0 means3D, 1 means2D, 2 sh, 3 colors_precomp, 4 language_feature_precomp, 5 weights_quick,
        #         6 indices_quick, 7 opacities, 8 scales, 9 rotations, 10 cov3Ds_precomp, 11 settings
        if ctx.geometry:
            # the RGB stage (train.py without --include_feature): everything the rendered image depends on
            if st.mode != MODE_RGB:
                raise NotImplementedError(
                    "geometry gradients through the language-feature weight map: the reference never asks for them "
                    "(the feature stage freezes geometry, scene/gaussian_model.py:238-243; the RGB stage renders "
                    "with include_feature=False)")
            try:
                g = rasterize_backward_geometry(st, grad_color)
                if ctx.debug:
                    torch.cuda.synchronize(grad_color.device)
            except Exception:
                if ctx.debug:
                    _dump_snapshot("snapshot_bw.dump", dict(grad_color=grad_color, radii=st.geom.get("radii"),
                                                            num_rendered=st.capacity))
                raise
            pick = lambda i, k: g.get(k) if need[i] else None
            d_op = g["opacities"].reshape(ctx.opacity_shape) if need[7] else None
            return (pick(0, "means3D"), pick(1, "means2D"), pick(2, "shs"), pick(3, "colors_precomp"), None, None,
                    None, d_op, pick(8, "scales"), pick(9, "rotations"), pick(10, "cov3D_precomp"), None, None, None,
                    None, None)
        d_coeff = d_colors = None
        if st.mode == MODE_DENSE and need[4]:
            gc = grad_color if (need[3] and ctx.color_is_precomp) else None
            try:
                d_coeff, d_colors = rasterize_backward(st, grad_wmap, gc, dense_grad=ctx.dense_grad,
                                                       rows_out=ctx.grad_rows)
                if ctx.grad_rows is not None:
                    _RasterizeGaussians.last_rows, ctx.sparse_rows = ctx.sparse_rows, None
                if ctx.debug:
                    torch.cuda.synchronize(grad_wmap.device)
            except Exception:
                if ctx.debug:
                    _dump_snapshot("snapshot_bw.dump", dict(grad_color=grad_color, grad_wmap=grad_wmap,
                                                            radii=st.geom.get("radii"), num_rendered=st.capacity))
                raise
        elif need[3]:
            raise NotImplementedError("colour gradients are produced together with the feature backward only")
        return (None, None, None, d_colors, d_coeff, None, None, None, None, None, None, None, None, None, None,
                None)


class GaussianRasterizer(nn.Module):
    """Drop-in for ``diff_gaussian_rasterization.GaussianRasterizer``."""

    def __init__(self, raster_settings: GaussianRasterizationSettings, capacity: Optional[int] = None):
        """``capacity`` (extension, default off): an upper bound on the number of (Gaussian, tile) instances.  With it
        the frame runs without the reference's one blocking device->host read of ``num_rendered``; the true count and
        an overflow flag stay on the device in ``last_state.status``."""
        super().__init__()
        self.raster_settings = raster_settings
        self.capacity = capacity
        self.dense_coeff_grad: Optional[bool] = None     # None: module default DENSE_COEFF_GRAD (strict)
        self.prepared: Optional[FrameState] = None       # extension: geometry stage done earlier (rasterize_forward)
        self.sparse_rows: Optional[dict] = None          # extension: rows emitted by the top-k producer (rasterize_forward)
        self.last_state: Optional[FrameState] = None

    def markVisible(self, positions: torch.Tensor) -> torch.Tensor:
        L_ = _abi.lib()
        with torch.no_grad():
            pos = _f32c(positions, "positions")
            vm = _f32c(self.raster_settings.viewmatrix.to(pos.device), "viewmatrix")
            out = torch.empty(pos.shape[0], dtype=torch.uint8, device=pos.device)
            _abi.check(L_.lsv2_mark_visible(pos.shape[0], pos.data_ptr(), vm.data_ptr(), out.data_ptr(),
                                            _abi.stream_ptr(pos.device)), "lsv2_mark_visible")
        return out.bool()

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, language_feature_precomp=None,
                language_feature_weights_quick=None, language_feature_indices=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        out = _RasterizeGaussians.apply(means3D, means2D, shs, colors_precomp, language_feature_precomp,
                                        language_feature_weights_quick, language_feature_indices, opacities,
                                        scales, rotations, cov3D_precomp, rs, self.capacity, self.dense_coeff_grad,
                                        self.prepared, self.sparse_rows)
        # the full frame state (outputs included) lives on the module, outside the autograd graph
        self.last_state = _RasterizeGaussians.last_state
        _RasterizeGaussians.last_state = None
        return out
