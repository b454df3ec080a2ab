#!/usr/bin/env python
"""Generates tests/golden/reference_python.npz by RUNNING the reference's own Python.

Run in the build container only (needs /root/reference; the GPU box has no such path):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The rasterizer core of the reference (submodule efficient-langsplat-rasterization) is absent from the
scrape, so nothing here pins it.  What IS present -- and what the hot path's inputs/outputs are defined
by -- is imported unmodified and evaluated on seeded inputs:

  vq_*      utils/vq_utils.py:9-24 (softmax_to_topk_soft_code), :26-40 (get_weights_and_indices)
  cam_*     utils/graphics_utils.py:38-49 (getWorld2View2), :51-71 (getProjectionMatrix), composed
            exactly as scene/cameras.py:55-58 does (without the .cuda())
  cov_*     utils/general_utils.py:78-110 (build_rotation / build_scaling_rotation) + :64-76
            (strip_symmetric), composed as scene/gaussian_model.py:28-32
  sh_*      utils/sh_utils.py:57-112 (eval_sh) with the +0.5 / clamp_min(0) of
            gaussian_renderer/__init__.py:77-81
  loss_*    utils/loss_utils.py:18-19 (l1_loss), :24-25 (cos_loss)
  dec_*     eval_lerf.py:214-218 (einsum decode + L2 normalise) and scene/gaussian_model.py:545-549
            (compute_final_feature_map).  Those two live in modules whose imports (open_clip, plyfile,
            simple_knn) are absent here, so their two tensor expressions are evaluated by this script
            with the same torch calls on the same shapes.

``general_utils`` hard-codes device="cuda" in torch.zeros; the script substitutes a torch.zeros that
drops the device argument while those functions run (the arithmetic is untouched).
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.dont_write_bytecode = True
sys.path.insert(0, REF)

from utils import general_utils, graphics_utils, loss_utils, sh_utils, vq_utils  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_python.npz")


class _CpuZeros:
    def __enter__(self):
        self.orig = torch.zeros

        def zeros(*a, **k):
            k.pop("device", None)
            return self.orig(*a, **k)
        torch.zeros = zeros

    def __exit__(self, *exc):
        torch.zeros = self.orig


def main():
    g = torch.Generator().manual_seed(1234)
    out = {}

    # ---- sparse coefficients -----------------------------------------------------------------
    logits = torch.randn(301, 64, generator=g)
    logits[0] = 0.0                       # ties: every probability equal
    logits[1, :8] = 5.0                   # ties inside the top-k
    logits[2] = torch.linspace(-30, 30, 64)  # near one-hot
    out["vq_logits"] = logits.numpy()
    for k in (1, 3, 4):
        out[f"vq_dense_k{k}"] = vq_utils.softmax_to_topk_soft_code(logits, k).numpy()
        # rows 0/1 have ties: which tied channels survive is torch.topk's choice; they are kept in the
        # file but the tests compare only the tie-free rows (3:) bit-tightly
        w, i = vq_utils.get_weights_and_indices(logits[3:], k)
        out[f"vq_w_k{k}"] = w.numpy()
        out[f"vq_i_k{k}"] = i.numpy()

    # ---- cameras ---------------------------------------------------------------------------------
    poses = []
    for v in range(5):
        az, el = 2 * math.pi * v / 5 + 0.3, 0.2 + 0.1 * v
        c = 4.0 * np.array([math.cos(el) * math.cos(az), math.sin(el), math.cos(el) * math.sin(az)])
        z = -c / np.linalg.norm(c)
        x = np.cross(z, [0.0, 1.0, 0.0])
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        R = np.stack([x, y, z], axis=1)
        t = -R.T @ c
        fovx = math.radians(50 + 5 * v)
        W, H = 640 + 16 * v, 360 + 7 * v
        fovy = 2 * math.atan(math.tan(fovx / 2) * H / W)
        wv = torch.tensor(graphics_utils.getWorld2View2(R, t, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
        proj = graphics_utils.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
        full = (wv.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        center = wv.inverse()[3, :3]
        poses.append((R, t, fovx, fovy, W, H, wv, proj, full, center))
    out["cam_R"] = np.stack([p[0] for p in poses])
    out["cam_t"] = np.stack([p[1] for p in poses])
    out["cam_fov"] = np.array([[p[2], p[3]] for p in poses])
    out["cam_wh"] = np.array([[p[4], p[5]] for p in poses], dtype=np.int32)
    out["cam_world_view"] = np.stack([p[6].numpy() for p in poses])
    out["cam_proj"] = np.stack([p[7].numpy() for p in poses])
    out["cam_full_proj"] = np.stack([p[8].numpy() for p in poses])
    out["cam_center"] = np.stack([p[9].numpy() for p in poses])

    # ---- cov3D ----------------------------------------------------------------------------------
    P = 200
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.6 + math.log(0.05))
    rots = torch.nn.functional.normalize(torch.randn(P, 4, generator=g))
    with _CpuZeros():
        for mod in (1.0, 0.7):
            Lm = general_utils.build_scaling_rotation(mod * scales, rots)
            cov = general_utils.strip_symmetric(Lm @ Lm.transpose(1, 2))
            out[f"cov_cov3d_mod{mod}"] = cov.numpy()
    out["cov_scales"] = scales.numpy()
    out["cov_rots"] = rots.numpy()

    # ---- SH -> RGB ------------------------------------------------------------------------------
    xyz = torch.rand(P, 3, generator=g) * 4 - 2
    sh = torch.randn(P, 16, 3, generator=g) * 0.3
    campos = poses[0][9]
    for deg in (0, 1, 2, 3):
        shs_view = sh.transpose(1, 2).reshape(-1, 3, 16)
        dir_pp = xyz - campos.repeat(P, 1)
        dir_n = dir_pp / dir_pp.norm(dim=1, keepdim=True)
        rgb = torch.clamp_min(sh_utils.eval_sh(deg, shs_view, dir_n) + 0.5, 0.0)
        out[f"sh_rgb_deg{deg}"] = rgb.numpy()
    out["sh_xyz"] = xyz.numpy()
    out["sh_coeffs"] = sh.numpy()
    out["sh_campos"] = campos.numpy()

    # ---- decode ---------------------------------------------------------------------------------
    Hh, Ww = 4, 6
    cb = torch.randn(3, 64, 512, generator=g)
    wmap = torch.rand(192, Hh, Ww, generator=g) * (torch.rand(192, Hh, Ww, generator=g) < 0.1)
    wmap[:, 0, 0] = 0.0                   # an empty pixel: 0 / (0 + 1e-10)
    lw = wmap.view(3, 64, Hh, Ww).view(3, 64, Hh * Ww)
    fm = torch.einsum('ldk,lkn->ldn', cb.permute(0, 2, 1), lw).view(3, 512, Hh, Ww)
    out["dec_quick_raw"] = fm.numpy()
    out["dec_quick_norm"] = (fm / (fm.norm(dim=1, keepdim=True) + 1e-10)).numpy()
    final = (cb.view(-1, 512).T @ wmap.view(192, -1)).view(512, Hh, Ww)       # compute_final_feature_map
    out["dec_final"] = final.numpy()
    layer0 = (cb[0].T @ wmap.view(192, -1)[0:64]).view(512, Hh, Ww)           # compute_layer_feature_map(.., 0)
    out["dec_layer0"] = layer0.numpy()
    out["dec_codebooks"] = cb.numpy()
    out["dec_wmap"] = wmap.numpy()

    # ---- losses ---------------------------------------------------------------------------------
    a = torch.randn(512, 6, 7, generator=g)
    b = torch.randn(512, 6, 7, generator=g)
    out["loss_a"], out["loss_b"] = a.numpy(), b.numpy()
    out["loss_cos"] = np.float32(loss_utils.cos_loss(a, b).item())
    out["loss_l1"] = np.float32(loss_utils.l1_loss(a, b).item())

    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
