"""CPU oracle for the LangSplatV2 rasterization hot path (TEST INFRASTRUCTURE).

ctypes/numpy wrapper around ``lsv2_oracle.c`` (see that file's header for the
parity status: rasterizer core "parity unpinned", Python-present parts pinned by
``tests/golden``).  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package; the product
(``langsplatv2_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblsv2_oracle.so")
_lib = None

TILE = 16


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "lsv2_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "liblsv2_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.lsv2o_scan.restype = C.c_int64
        _lib.lsv2o_expf.restype = C.c_float
        _lib.lsv2o_expf.argtypes = [C.c_float]
    return _lib


def num_threads() -> int:
    return int(lib().lsv2o_num_threads())


def _p(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle wants contiguous arrays"
    return a.ctypes.data_as(C.c_void_p)


def _f32(a) -> Optional[np.ndarray]:
    if a is None:
        return None
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def expf(x) -> np.ndarray:
    x = _f32(x).ravel()
    y = np.empty_like(x)
    lib().lsv2o_expf_array(C.c_int(x.size), _p(x), _p(y))
    return y


def grid(W: int, H: int):
    return (W + TILE - 1) // TILE, (H + TILE - 1) // TILE


def preprocess(means3d, opacities, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
               scales=None, rotations=None, cov3d_precomp=None, shs=None, colors_precomp=None,
               sh_degree=0, scale_modifier=1.0):
    means3d = _f32(means3d)
    P = means3d.shape[0]
    opacities = _f32(opacities).reshape(-1)
    scales, rotations, cov3d_precomp = _f32(scales), _f32(rotations), _f32(cov3d_precomp)
    shs, colors_precomp = _f32(shs), _f32(colors_precomp)
    assert (shs is None) != (colors_precomp is None)
    assert (cov3d_precomp is None) != (scales is None or rotations is None)
    M = 0 if shs is None else shs.shape[1]
    vm, pm, cp = _f32(viewmatrix).reshape(16), _f32(projmatrix).reshape(16), _f32(campos).reshape(3)
    out = dict(
        radii=np.zeros(P, np.int32), xy=np.zeros((P, 2), np.float32), depth=np.zeros(P, np.float32),
        cov3d=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
        rgb=np.zeros((P, 3), np.float32), clamped=np.zeros((P, 3), np.uint8),
        tiles_touched=np.zeros(P, np.uint32), rect=np.zeros((P, 4), np.int32))
    lib().lsv2o_preprocess(
        C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3d), _p(scales), C.c_float(scale_modifier),
        _p(rotations), _p(opacities), _p(shs), _p(colors_precomp), _p(cov3d_precomp),
        _p(vm), _p(pm), _p(cp), C.c_int(W), C.c_int(H), C.c_float(tanfovx), C.c_float(tanfovy),
        _p(out["radii"]), _p(out["xy"]), _p(out["depth"]), _p(out["cov3d"]), _p(out["conic_opacity"]),
        _p(out["rgb"]), _p(out["clamped"]), _p(out["tiles_touched"]), _p(out["rect"]))
    return out


def bin_instances(pre: dict, W: int, H: int):
    P = pre["radii"].shape[0]
    offsets = np.zeros(P, np.uint32)
    R = int(lib().lsv2o_scan(C.c_int(P), _p(pre["tiles_touched"]), _p(offsets)))
    gx, gy = grid(W, H)
    n = max(R, 1)
    out = dict(R=R, offsets=offsets,
               keys_unsorted=np.zeros(n, np.uint64), vals_unsorted=np.zeros(n, np.uint32),
               keys=np.zeros(n, np.uint64), point_list=np.zeros(n, np.uint32),
               ranges=np.zeros((gx * gy, 2), np.uint32))
    sk, sv = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    lib().lsv2o_bin(C.c_int(P), C.c_int(W), C.c_int(H), _p(pre["radii"]), _p(pre["rect"]),
                    _p(pre["depth"]), _p(offsets), C.c_int64(R),
                    _p(out["keys_unsorted"]), _p(out["vals_unsorted"]), _p(out["keys"]),
                    _p(out["point_list"]), _p(sk), _p(sv), _p(out["ranges"]))
    for k in ("keys_unsorted", "vals_unsorted", "keys", "point_list"):
        out[k] = out[k][:R]
    return out


def render_forward(pre: dict, binned: dict, W: int, H: int, bg, mode: int = 0, L: int = 0,
                   coeff_dense=None, weights_quick=None, indices_quick=None, tile_stride: int = 1,
                   tile_phase: int = 0):
    bg = _f32(bg).reshape(3)
    coeff_dense, wq, iq = _f32(coeff_dense), _f32(weights_quick), _f32(indices_quick)
    K = 0 if wq is None else wq.shape[1]
    out = dict(color=np.zeros((3, H, W), np.float32), wmap=np.zeros((max(L, 0), H, W), np.float32),
               final_T=np.zeros((H, W), np.float32), n_contrib=np.zeros((H, W), np.uint32))
    pl = np.ascontiguousarray(binned["point_list"]) if binned["R"] > 0 else np.zeros(1, np.uint32)
    lib().lsv2o_render_forward_tiles(
        C.c_int(W), C.c_int(H), _p(binned["ranges"]), _p(pl), _p(pre["xy"]),
        _p(pre["conic_opacity"]), _p(pre["rgb"]), C.c_int(mode), C.c_int(L), C.c_int(K),
        _p(coeff_dense), _p(wq), _p(iq), _p(bg),
        _p(out["color"]), _p(out["wmap"]), _p(out["final_T"]), _p(out["n_contrib"]),
        C.c_int(tile_stride), C.c_int(tile_phase))
    return out


def tile_pixel_mask(W: int, H: int, tile_stride: int, tile_phase: int = 0) -> np.ndarray:
    """bool [H, W]: pixels of the tiles that render_forward(tile_stride, tile_phase) renders."""
    gx, gy = grid(W, H)
    ty, tx = np.meshgrid(np.arange(H) // TILE, np.arange(W) // TILE, indexing="ij")
    return ((ty * gx + tx) % tile_stride) == tile_phase


def render_backward_coeff(pre: dict, binned: dict, fwd: dict, W: int, H: int, L: int, dL_dwmap,
                          dense: bool = True, coeff_dense=None, dL_dcolor=None):
    P = pre["radii"].shape[0]
    dL_dwmap = _f32(dL_dwmap).reshape(L, H, W)
    coeff_dense, dL_dcolor = _f32(coeff_dense), _f32(dL_dcolor)
    g = np.zeros((P, L), np.float64)
    gc = np.zeros((P, 3), np.float64) if dL_dcolor is not None else None
    pl = np.ascontiguousarray(binned["point_list"]) if binned["R"] > 0 else np.zeros(1, np.uint32)
    lib().lsv2o_render_backward_coeff(
        C.c_int(W), C.c_int(H), _p(binned["ranges"]), _p(pl), _p(pre["xy"]),
        _p(pre["conic_opacity"]), _p(fwd["n_contrib"]), C.c_int(L), C.c_int(1 if dense else 0),
        _p(coeff_dense), _p(dL_dwmap), _p(dL_dcolor), _p(g), _p(gc))
    return (g, gc) if gc is not None else g


def decode(wmap, codebooks, normalize: bool = True):
    """wmap [levels*Lc, ...pixels], codebooks [levels, Lc, D] -> [levels, D, ...pixels]."""
    wmap, codebooks = _f32(wmap), _f32(codebooks)
    levels, Lc, D = codebooks.shape
    pix_shape = wmap.shape[1:]
    HW = int(np.prod(pix_shape))
    assert wmap.shape[0] == levels * Lc
    out = np.zeros((levels, D) + tuple(pix_shape), np.float32)
    lib().lsv2o_decode(C.c_int(levels), C.c_int(Lc), C.c_int(D), C.c_int64(HW), _p(wmap),
                       _p(codebooks), C.c_int(1 if normalize else 0), _p(out))
    return out


def decode_backward(wmap, codebooks, dF):
    wmap, codebooks, dF = _f32(wmap), _f32(codebooks), _f32(dF)
    levels, Lc, D = codebooks.shape
    HW = int(np.prod(wmap.shape[1:]))
    dW = np.zeros(wmap.shape, np.float64)
    dCb = np.zeros(codebooks.shape, np.float64)
    lib().lsv2o_decode_backward(C.c_int(levels), C.c_int(Lc), C.c_int(D), C.c_int64(HW), _p(wmap),
                                _p(codebooks), _p(dF), _p(dW), _p(dCb))
    return dW, dCb


def topk_soft_code(logits, k: int):
    """utils/vq_utils.py:9-24 -> dense [N, C] with k non-zeros."""
    logits = _f32(logits)
    N, Cc = logits.shape
    dense = np.zeros((N, Cc), np.float32)
    lib().lsv2o_topk_soft_code(C.c_int64(N), C.c_int(Cc), C.c_int(k), _p(logits), _p(dense), None, None)
    return dense


def weights_and_indices(logits, k: int):
    """utils/vq_utils.py:26-40 -> (weights [N,k], indices [N,k] float32, ascending channel order)."""
    logits = _f32(logits)
    N, Cc = logits.shape
    w = np.zeros((N, k), np.float32)
    i = np.zeros((N, k), np.float32)
    lib().lsv2o_topk_soft_code(C.c_int64(N), C.c_int(Cc), C.c_int(k), _p(logits), None, _p(w), _p(i))
    return w, i


def forward(scene: dict, view: dict, mode: int = 0, L: int = 0, coeff_dense=None,
            weights_quick=None, indices_quick=None, bg=(0.0, 0.0, 0.0)):
    """Whole forward (R6+R7+R8).  scene: means3d, opacities, scales, rotations, shs|colors,
    sh_degree; view: viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy."""
    W, H = int(view["W"]), int(view["H"])
    pre = preprocess(scene["means3d"], scene["opacities"], view["viewmatrix"], view["projmatrix"],
                     view["campos"], W, H, view["tanfovx"], view["tanfovy"],
                     scales=scene.get("scales"), rotations=scene.get("rotations"),
                     cov3d_precomp=scene.get("cov3d_precomp"), shs=scene.get("shs"),
                     colors_precomp=scene.get("colors_precomp"),
                     sh_degree=int(scene.get("sh_degree", 0)),
                     scale_modifier=float(view.get("scale_modifier", 1.0)))
    binned = bin_instances(pre, W, H)
    fwd = render_forward(pre, binned, W, H, bg, mode=mode, L=L, coeff_dense=coeff_dense,
                         weights_quick=weights_quick, indices_quick=indices_quick)
    return pre, binned, fwd


def relevancy(feature_map, phrase_embeds, n_pos: int):
    """Restatement of OpenCLIPNetwork.get_max_across_quick (/root/reference/eval/openclip_encoder.py:114-138):
    feature_map [levels, 512, H, W] (the decoded, normalised map), phrase_embeds [P+N, 512] (positives first) ->
    relevancy [levels, P, H, W] = min over the negatives of softmax(10 * [pos, neg])[0], in float32 like the
    reference (einsum, pairwise softmax over the stacked pair, min)."""
    f = _f32(feature_map)
    e = _f32(phrase_embeds)
    levels, D, H, W = f.shape
    sim = np.einsum("ldn,pd->lnp", f.reshape(levels, D, H * W), e).astype(np.float32)        # [levels, HW, P+N]
    pos, neg = sim[:, :, :n_pos], sim[:, :, n_pos:]
    a = (np.float32(10.0) * pos)[:, :, :, None]                                              # [levels, HW, P, 1]
    b = (np.float32(10.0) * neg)[:, :, None, :]                                              # [levels, HW, 1, N]
    m = np.maximum(a, b)
    ea, eb = np.exp(a - m, dtype=np.float32), np.exp(b - m, dtype=np.float32)
    p0 = ea / (ea + eb)                                                                      # softmax(...)[..., 0]
    return np.ascontiguousarray(p0.min(axis=-1).transpose(0, 2, 1).reshape(levels, n_pos, H, W))
