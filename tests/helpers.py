"""Shared test plumbing: seeded scenes, the oracle (checker) and the CUDA path side by side."""
import math

import numpy as np
import torch

import oracle
from langsplatv2_b200 import synth
from langsplatv2_b200.rasterizer import GaussianRasterizationSettings, rasterize_forward

MODE_RGB, MODE_DENSE, MODE_QUICK = 0, 1, 2


def make_case(N, W, H, seed=0, levels=3, view=0, views=8, sh_degree=3, scale_boost=1.0):
    scene = synth.make_scene(N, seed=seed, levels=levels, sh_degree=sh_degree)
    if scale_boost != 1.0:
        scene["scaling"] = scene["scaling"] * scale_boost
    cam = synth.ring_camera(view, views, W, H)
    return scene, cam


def oracle_scene(scene):
    return dict(means3d=scene["xyz"].numpy(), opacities=scene["opacity"].numpy(),
                scales=scene["scaling"].numpy(), rotations=scene["rotation"].numpy(),
                shs=scene["features"].numpy(), sh_degree=scene["sh_degree"])


def oracle_view(cam):
    return dict(viewmatrix=cam["world_view_transform"].numpy(), projmatrix=cam["full_proj_transform"].numpy(),
                campos=cam["camera_center"].numpy(), W=cam["image_width"], H=cam["image_height"],
                tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])


def settings(cam, scene, bg, include_feature=False, quick_render=False, device="cuda", debug=False):
    return GaussianRasterizationSettings(
        image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
        tanfovy=cam["tanfovy"], bg=torch.tensor(bg, dtype=torch.float32, device=device), scale_modifier=1.0,
        viewmatrix=cam["world_view_transform"].to(device), projmatrix=cam["full_proj_transform"].to(device),
        sh_degree=scene["sh_degree"], campos=cam["camera_center"].to(device), prefiltered=False, debug=debug,
        include_feature=include_feature, quick_render=quick_render)


def coeffs(scene, k=4):
    """(dense [N, levels*64], weights [N, levels*k], indices [N, levels*k]) from the checker's restatement
    of utils/vq_utils.py, merged across levels as eval_lerf.py:333-345 does."""
    levels, cs = scene["levels"], scene["codebook_size"]
    logits = scene["logits"].numpy()
    dense, ws, idx = [], [], []
    for l in range(levels):
        lg = logits[:, l * cs:(l + 1) * cs]
        dense.append(oracle.topk_soft_code(lg, k))
        w, i = oracle.weights_and_indices(lg, k)
        ws.append(w)
        idx.append(i + l * cs)
    return np.concatenate(dense, 1), np.concatenate(ws, 1), np.concatenate(idx, 1)


def run_cuda(scene, cam, bg, mode, dense=None, wq=None, iq=None, capacity=None, debug_buffers=True, device="cuda",
             render=True, blend="auto"):
    rs = settings(cam, scene, bg, include_feature=(mode != MODE_RGB), quick_render=(mode == MODE_QUICK), device=device)
    t = lambda a: None if a is None else torch.as_tensor(a).to(device)
    return rasterize_forward(rs, scene["xyz"].to(device), scene["opacity"].to(device), shs=scene["features"].to(device),
                             scales=scene["scaling"].to(device), rotations=scene["rotation"].to(device),
                             language_feature_precomp=t(dense), language_feature_weights_quick=t(wq),
                             language_feature_indices=t(iq), capacity=capacity, debug_buffers=debug_buffers,
                             render=render, blend=blend)


def u32(t):
    return t.detach().cpu().numpy().view(np.uint32)


def u64(t):
    return t.detach().cpu().numpy().view(np.uint64)


def ref_keys(st, R=None):
    """The reference's sort keys (tile << 32 | depth bits) of the sorted instance list, rebuilt from the packed words
    (tile << 32 | Gaussian id) the CUDA sort leaves in binning["keys"] (include/lsv2.h, lsv2_binning.keys)."""
    R = st.num_rendered if R is None else R
    packed = u64(st.binning["keys"])[:R]
    depth = st.geom["depth"].detach().cpu().numpy().view(np.uint32)
    ids = (packed & np.uint64(0xFFFFFFFF)).astype(np.int64)
    return (packed & np.uint64(0xFFFFFFFF00000000)) | depth[ids].astype(np.uint64)
