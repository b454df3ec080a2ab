"""Scene readers / writers either side of the hot path: the formats a LangSplatV2 run leaves on disk,
turned into the scene dictionary the renderer / trainer of this package consume (``synth.make_scene`` documents its
keys: raw parameters as the checkpoint tuple holds them + the activated ones the op reads).

  * ``read_ply`` / ``write_ply``        -- the trained-Gaussian PLY of ``GaussianModel.save_ply`` / ``load_ply``
                                           (/root/reference/scene/gaussian_model.py:269-301, 308-350)
  * ``read_point_cloud`` / ``write_point_cloud`` -- the SfM point cloud of ``fetchPly`` / ``storePly``
                                           (scene/dataset_readers.py:120-143)
  * ``scene_from_point_cloud``          -- ``GaussianModel.create_from_pcd`` (gaussian_model.py:184-210): RGB -> SH DC,
                                           scales from ``distCUDA2`` (this package's kernel), identity rotations,
                                           opacity 0.1
  * ``read_checkpoint`` / ``checkpoint_tuple`` -- ``(GaussianModel.capture(...), iteration)`` as ``train.py:267`` saves it
                                           and ``restore`` unpacks it (gaussian_model.py:67-138)
  * ``combine_levels``                  -- the three per-level checkpoints of an evaluation merged into one quick-render
                                           scene (eval_lerf.py:333-346, backend_renderer.py:110-128)

The reference reads PLY through the ``plyfile`` package; this module parses the (binary little-endian or ASCII) format
itself with numpy, so it has no dependency outside the image.  Plain host-side IO: nothing here is on the timed path.
"""
from __future__ import annotations

import math
import os
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

SH_C0 = 0.28209479177387814            # utils/sh_utils.py:24

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


# ---- PLY container ------------------------------------------------------------------------------------------------
def _read_ply_vertices(path: str) -> np.ndarray:
    """The ``vertex`` element of a PLY file as a numpy structured array (scalar properties only -- which is all the
    reference writes; a list property in the vertex element is rejected)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt = None
        elements = []                    # (name, count, [(prop, dtype)])
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: truncated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append((tok[1], int(tok[2]), []))
            elif tok[0] == "property":
                if tok[1] == "list":
                    elements[-1][2].append((tok[-1], None))
                else:
                    if tok[1] not in _PLY_TYPES:
                        raise ValueError(f"{path}: unknown PLY type {tok[1]!r}")
                    elements[-1][2].append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt not in ("binary_little_endian", "binary_big_endian", "ascii"):
            raise ValueError(f"{path}: unsupported PLY format {fmt!r}")
        for name, count, props in elements:
            if name != "vertex":
                if any(t is None for _, t in props):
                    raise ValueError(f"{path}: element {name!r} precedes 'vertex' and holds a list property")
                # skip a fixed-size element in front of the vertices
                if fmt == "ascii":
                    for _ in range(count):
                        f.readline()
                else:
                    f.seek(count * sum(np.dtype(t).itemsize for _, t in props), os.SEEK_CUR)
                continue
            if any(t is None for _, t in props):
                raise ValueError(f"{path}: list property in the vertex element")
            order = "<" if fmt != "binary_big_endian" else ">"
            dt = np.dtype([(p, order + t) for p, t in props])
            if fmt == "ascii":
                rows = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2)
                if rows.shape != (count, len(props)):
                    raise ValueError(f"{path}: expected {count} x {len(props)} vertex values, got {rows.shape}")
                out = np.empty(count, dtype=dt)
                for j, (p, _) in enumerate(props):
                    out[p] = rows[:, j]
                return out
            raw = f.read(count * dt.itemsize)
            if len(raw) != count * dt.itemsize:
                raise ValueError(f"{path}: truncated vertex data ({len(raw)} of {count * dt.itemsize} bytes)")
            return np.frombuffer(raw, dtype=dt, count=count)
    raise ValueError(f"{path}: no vertex element")


def _write_ply_vertices(path: str, names: Sequence[str], types: Sequence[str], columns: np.ndarray) -> None:
    """binary_little_endian, one ``vertex`` element -- byte-compatible with what plyfile writes for the reference."""
    inv = {"f4": "float", "u1": "uchar"}
    dt = np.dtype([(n, "<" + t) for n, t in zip(names, types)])
    rec = np.empty(columns.shape[0], dtype=dt)
    for j, n in enumerate(names):
        rec[n] = columns[:, j]
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    with open(path, "wb") as f:
        head = ["ply", "format binary_little_endian 1.0", f"element vertex {columns.shape[0]}"]
        head += [f"property {inv[t]} {n}" for n, t in zip(names, types)]
        head.append("end_header")
        f.write(("\n".join(head) + "\n").encode("ascii"))
        f.write(rec.tobytes())


def _numbered(names: Sequence[str], prefix: str):
    sel = [n for n in names if n.startswith(prefix)]
    return sorted(sel, key=lambda n: int(n.split("_")[-1]))           # gaussian_model.py:323, 333, 339


# ---- scene dictionary ---------------------------------------------------------------------------------------------
def _scene(xyz, features, log_scale, rot_raw, opacity_logit, sh_degree: int, device="cpu", **extra) -> Dict:
    f32 = lambda a: torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a).detach() \
        .to(device=device, dtype=torch.float32).contiguous()
    xyz, features, log_scale, rot_raw, opacity_logit = map(f32, (xyz, features, log_scale, rot_raw, opacity_logit))
    out = dict(xyz=xyz, features=features, log_scale=log_scale, rot_raw=rot_raw, opacity_logit=opacity_logit,
               scaling=torch.exp(log_scale), rotation=torch.nn.functional.normalize(rot_raw),
               opacity=torch.sigmoid(opacity_logit), sh_degree=int(sh_degree))
    out.update(extra)
    return out


def read_ply(path: str, max_sh_degree: int = 3, device="cpu") -> Dict:
    """``GaussianModel.load_ply`` (gaussian_model.py:308-350): x y z, f_dc_0..2, f_rest_* (channel-major: the file holds
    [P, 3, M-1] flattened), opacity (logit), scale_* (log), rot_* (unnormalised quaternion w x y z)."""
    v = _read_ply_vertices(path)
    names = v.dtype.names
    need = ["x", "y", "z", "opacity", "f_dc_0", "f_dc_1", "f_dc_2"]
    missing = [n for n in need if n not in names]
    if missing:
        raise ValueError(f"{path}: not a trained-Gaussian PLY (missing {missing})")
    P = v.shape[0]
    col = lambda n: np.asarray(v[n], dtype=np.float32)
    xyz = np.stack([col("x"), col("y"), col("z")], 1)
    dc = np.stack([col("f_dc_0"), col("f_dc_1"), col("f_dc_2")], 1)[:, None, :]          # [P, 1, 3]
    rest_names = _numbered(names, "f_rest_")
    M = (max_sh_degree + 1) ** 2
    if len(rest_names) != 3 * M - 3:                                                      # gaussian_model.py:324
        raise ValueError(f"{path}: {len(rest_names)} f_rest_* properties, sh degree {max_sh_degree} needs {3 * M - 3}")
    rest = np.stack([col(n) for n in rest_names], 1).reshape(P, 3, M - 1).transpose(0, 2, 1) if M > 1 \
        else np.zeros((P, 0, 3), np.float32)
    scale_names, rot_names = _numbered(names, "scale_"), _numbered(names, "rot")
    if len(scale_names) != 3 or len(rot_names) != 4:
        raise ValueError(f"{path}: expected scale_0..2 and rot_0..3")
    scales = np.stack([col(n) for n in scale_names], 1)
    rots = np.stack([col(n) for n in rot_names], 1)
    return _scene(xyz, np.concatenate([dc, rest], 1), scales, rots, col("opacity")[:, None], max_sh_degree, device)


def write_ply(path: str, scene: Dict) -> None:
    """``GaussianModel.save_ply`` (gaussian_model.py:269-301): property order x y z nx ny nz f_dc_* f_rest_* opacity
    scale_* rot_*, all float32, normals zero."""
    g = lambda k: scene[k].detach().cpu().float().numpy()
    xyz, feat = g("xyz"), g("features")
    P, M = feat.shape[0], feat.shape[1]
    f_dc = feat[:, :1, :].transpose(0, 2, 1).reshape(P, -1)
    f_rest = feat[:, 1:, :].transpose(0, 2, 1).reshape(P, -1)
    cols = np.concatenate([xyz, np.zeros_like(xyz), f_dc, f_rest, g("opacity_logit").reshape(P, 1), g("log_scale"),
                           g("rot_raw")], 1)
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + \
            [f"f_rest_{i}" for i in range(3 * (M - 1))] + ["opacity"] + [f"scale_{i}" for i in range(3)] + \
            [f"rot_{i}" for i in range(4)]
    _write_ply_vertices(path, names, ["f4"] * len(names), cols)


def read_point_cloud(path: str) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """``fetchPly`` (dataset_readers.py:120-126): positions [P, 3], colours [P, 3] in [0, 1] (stored as uchar), normals."""
    v = _read_ply_vertices(path)
    names = v.dtype.names
    for n in ("x", "y", "z", "red", "green", "blue"):
        if n not in names:
            raise ValueError(f"{path}: not a point-cloud PLY (missing {n})")
    pos = np.stack([v["x"], v["y"], v["z"]], 1).astype(np.float64)
    col = np.stack([v["red"], v["green"], v["blue"]], 1).astype(np.float64) / 255.0
    nrm = np.stack([v[n] for n in ("nx", "ny", "nz")], 1).astype(np.float64) if "nx" in names else np.zeros_like(pos)
    return pos, col, nrm


def write_point_cloud(path: str, xyz: np.ndarray, rgb: np.ndarray) -> None:
    """``storePly`` (dataset_readers.py:128-143): x y z nx ny nz float + red green blue uchar (0..255)."""
    xyz = np.asarray(xyz, np.float32)
    cols = np.concatenate([xyz, np.zeros_like(xyz), np.asarray(rgb, np.float32)], 1)
    _write_ply_vertices(path, ["x", "y", "z", "nx", "ny", "nz", "red", "green", "blue"], ["f4"] * 6 + ["u1"] * 3, cols)


def scene_from_point_cloud(points, colors, max_sh_degree: int = 3, device="cuda") -> Dict:
    """``GaussianModel.create_from_pcd`` (gaussian_model.py:184-210).  ``distCUDA2`` is this package's kernel
    (``lsv2_dist2_knn3``), so this needs the CUDA device, as the reference's does."""
    from .knn import distCUDA2
    pts = torch.as_tensor(np.asarray(points)).float().to(device)
    col = torch.as_tensor(np.asarray(colors)).float().to(device)
    P, M = pts.shape[0], (max_sh_degree + 1) ** 2
    feats = torch.zeros((P, M, 3), dtype=torch.float32, device=device)
    feats[:, 0, :] = (col - 0.5) / SH_C0                                               # RGB2SH, sh_utils.py:114-115
    dist2 = torch.clamp_min(distCUDA2(pts), 0.0000001)
    log_scale = torch.log(torch.sqrt(dist2))[..., None].repeat(1, 3)
    rots = torch.zeros((P, 4), device=device)
    rots[:, 0] = 1
    op = torch.full((P, 1), math.log(0.1 / 0.9), dtype=torch.float32, device=device)   # inverse_sigmoid(0.1)
    return _scene(pts, feats, log_scale, rots, op, max_sh_degree, device)


# ---- checkpoints --------------------------------------------------------------------------------------------------
def read_checkpoint(source, device="cpu") -> Dict:
    """``(model_params, first_iter) = torch.load(chkpnt)`` + ``GaussianModel.restore`` (gaussian_model.py:103-138).
    ``source``: a path or the loaded pair.  The 14-tuple of a language-feature stage carries ``logits`` [P, layers*64]
    and ``codebooks`` [layers, 64, 512]; the 12-tuple of the RGB stage does not.  Extra keys: ``iteration``,
    ``active_sh_degree``, ``max_radii2D``, ``xyz_gradient_accum``, ``denom``, ``optimizer_state``, ``spatial_lr_scale``."""
    if isinstance(source, (str, os.PathLike)):
        source = torch.load(source, map_location="cpu", weights_only=False)
    params, iteration = source
    if len(params) == 14:
        (sh_deg, xyz, f_dc, f_rest, scaling, rotation, opacity, logits, codebooks, max_r, grad_accum, denom, opt,
         lr_scale) = params
    elif len(params) == 12:
        (sh_deg, xyz, f_dc, f_rest, scaling, rotation, opacity, max_r, grad_accum, denom, opt, lr_scale) = params
        logits = codebooks = None
    else:
        raise ValueError(f"checkpoint tuple of length {len(params)}: expected 12 (RGB stage) or 14 (language features)")
    feats = torch.cat([f_dc.detach(), f_rest.detach()], dim=1)                          # get_features, :166-169
    M = feats.shape[1]
    max_deg = int(round(math.sqrt(M))) - 1
    extra = dict(iteration=int(iteration), active_sh_degree=int(sh_deg), max_radii2D=max_r,
                 xyz_gradient_accum=grad_accum, denom=denom, optimizer_state=opt, spatial_lr_scale=lr_scale)
    if logits is not None:
        lg = logits.detach().to(device=device, dtype=torch.float32).contiguous()
        cb = codebooks.detach().to(device=device, dtype=torch.float32).contiguous()
        extra.update(logits=lg, codebooks=cb, levels=int(cb.shape[0]), codebook_size=int(cb.shape[1]),
                     feat_dim=int(cb.shape[2]))
    return _scene(xyz, feats, scaling, rotation, opacity, max_deg, device, **extra)


def checkpoint_tuple(scene: Dict, iteration: int, include_feature: Optional[bool] = None, optimizer_state=None):
    """The pair ``train.py:267`` saves -- ``torch.save(checkpoint_tuple(...), path)`` gives a file the reference's
    ``restore`` reads (gaussian_model.py:67-101 for the field order)."""
    if include_feature is None:
        include_feature = "logits" in scene and scene.get("logits") is not None
    P = scene["xyz"].shape[0]
    feat = scene["features"]
    dev = scene["xyz"].device
    zeros = lambda *s: torch.zeros(s, device=dev)
    head = (int(scene.get("active_sh_degree", scene.get("sh_degree", 3))), scene["xyz"], feat[:, :1, :].contiguous(),
            feat[:, 1:, :].contiguous(), scene["log_scale"], scene["rot_raw"], scene["opacity_logit"])
    tail = (scene.get("max_radii2D", zeros(P)), scene.get("xyz_gradient_accum", zeros(P, 1)),
            scene.get("denom", zeros(P, 1)), optimizer_state if optimizer_state is not None
            else scene.get("optimizer_state", {}), scene.get("spatial_lr_scale", 1.0))
    if include_feature:
        return (head + (scene["logits"], scene["codebooks"]) + tail, int(iteration))
    return (head + tail, int(iteration))


def combine_levels(level_scenes: Sequence[Dict], topk: int = 4) -> Dict:
    """Quick-render scene from one checkpoint per level (eval_lerf.py:333-346): geometry of the first, per level the
    top-k weights / indices (indices offset by level * codebook size, float-typed as the reference passes them) and the
    stacked codebooks [levels, 64, 512].  The weights come from this package's top-k kernel (CUDA tensors needed)."""
    from .sparse import get_weights_and_indices
    if not level_scenes:
        raise ValueError("no levels")
    out = dict(level_scenes[0])
    ws, idx, cbs = [], [], []
    for lvl, s in enumerate(level_scenes):
        if s["xyz"].shape[0] != out["xyz"].shape[0]:
            raise ValueError("the levels of one scene share their geometry: Gaussian counts differ")
        cb = s["codebooks"].reshape(-1, s["codebooks"].shape[-1])                       # .view(-1, 512)
        w, i = get_weights_and_indices(s["logits"], topk)
        ws.append(w)
        idx.append(i + float(lvl * s["codebooks"].shape[1]))
        cbs.append(cb)
    out.update(codebooks=torch.stack(cbs, 0), weights_quick=torch.cat(ws, 1), indices_quick=torch.cat(idx, 1),
               levels=len(level_scenes), logits=torch.cat([s["logits"] for s in level_scenes], 1))
    return out
