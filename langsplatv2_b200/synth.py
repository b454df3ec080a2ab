"""Seeded synthetic scenes and cameras for the BASELINE.json configs.

Camera matrices follow the reference's conventions exactly:
  * ``world_view_transform = getWorld2View2(R, T).T``  (utils/graphics_utils.py:38-49,
    scene/cameras.py:55) -- row-vector (transposed) form,
  * ``projection_matrix = getProjectionMatrix(znear=0.01, zfar=100, ...).T``
    (utils/graphics_utils.py:51-71, scene/cameras.py:49-56),
  * ``full_proj_transform = world_view_transform @ projection_matrix`` (cameras.py:57),
  * ``camera_center = inverse(world_view_transform)[3, :3]`` (cameras.py:58).
tests/test_golden.py checks these against golden vectors generated from the
reference's own functions.
"""
from __future__ import annotations

import math
from typing import Dict

import numpy as np
import torch

ZNEAR, ZFAR = 0.01, 100.0


def world_to_view(R: np.ndarray, t: np.ndarray, translate=(0.0, 0.0, 0.0), scale: float = 1.0) -> np.ndarray:
    """getWorld2View2 (utils/graphics_utils.py:38-49): R is camera-to-world rotation, t the
    world-to-camera translation."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    C2W[:3, 3] = (C2W[:3, 3] + np.asarray(translate)) * scale
    return np.float32(np.linalg.inv(C2W))


def projection_matrix(znear: float, zfar: float, fovX: float, fovY: float) -> torch.Tensor:
    """getProjectionMatrix (utils/graphics_utils.py:51-71)."""
    tanY, tanX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tanY * znear, tanX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def camera_from_pose(R: np.ndarray, t: np.ndarray, W: int, H: int, fovx: float, fovy: float,
                     device="cpu") -> Dict:
    wv = torch.tensor(world_to_view(R, t)).transpose(0, 1)
    proj = projection_matrix(ZNEAR, ZFAR, fovx, fovy).transpose(0, 1)
    full = wv.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
    center = wv.inverse()[3, :3]
    return dict(image_width=W, image_height=H, FoVx=fovx, FoVy=fovy,
                tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5),
                world_view_transform=wv.contiguous().to(device),
                full_proj_transform=full.contiguous().to(device),
                camera_center=center.contiguous().to(device))


def ring_camera(v: int, V: int, W: int, H: int, fovx_deg: float = 60.0, radius: float = 4.0,
                elev_deg: float = 15.0, device="cpu") -> Dict:
    """View v of V on a ring around the origin: azimuth 2*pi*v/V, fixed elevation,
    looking at the origin; camera axes x right, y down, z forward (COLMAP convention)."""
    az, el = 2.0 * math.pi * v / V, math.radians(elev_deg)
    c = radius * np.array([math.cos(el) * math.cos(az), math.sin(el), math.cos(el) * math.sin(az)])
    z = -c / np.linalg.norm(c)
    up = np.array([0.0, 1.0, 0.0])
    x = np.cross(z, up)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)          # camera-to-world rotation (columns = camera axes)
    t = -R.T @ c                              # world-to-camera translation
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * H / W)
    return camera_from_pose(R, t, W, H, fovx, fovy, device)


def make_scene(N: int, seed: int = 0, levels: int = 3, codebook_size: int = 64, feat_dim: int = 512,
               sh_degree: int = 3, device="cpu") -> Dict:
    """Random Gaussians, distributions of SURVEY.md §8(d).  Holds RAW parameters (as the reference's
    checkpoint tuple does, scene/gaussian_model.py:67-101) and the activated ones the op consumes
    (get_scaling = exp, get_rotation = normalize, get_opacity = sigmoid, gaussian_model.py:28-40)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    xyz = torch.rand(N, 3, generator=g) * 4.0 - 2.0
    log_scale = (torch.randn(N, 3, generator=g) * 0.6 + math.log(0.012)).clamp(math.log(1e-3), math.log(0.15))
    rot = torch.randn(N, 4, generator=g)
    op_logit = torch.randn(N, 1, generator=g) * 1.5
    M = (sh_degree + 1) ** 2
    sh = torch.randn(N, M, 3, generator=g) * 0.1
    sh[:, 0, :] = torch.randn(N, 3, generator=g)
    logits = torch.randn(N, levels * codebook_size, generator=g)
    codebooks = torch.randn(levels, codebook_size, feat_dim, generator=g)
    s = dict(xyz=xyz, log_scale=log_scale, rot_raw=rot, opacity_logit=op_logit, features=sh,
             logits=logits, codebooks=codebooks,
             scaling=torch.exp(log_scale), rotation=torch.nn.functional.normalize(rot),
             opacity=torch.sigmoid(op_logit))
    out = {k: v.contiguous().to(device) for k, v in s.items()}
    out.update(sh_degree=sh_degree, levels=levels, codebook_size=codebook_size, feat_dim=feat_dim)
    return out


def make_gt_features(W: int, H: int, seed: int = 2, feat_dim: int = 512, device="cpu"):
    """Piecewise-constant GT feature map of SURVEY.md §8(d): seg = (x//32 + 7*(y//32)) % 256,
    gt[:, y, x] = E[seg], E ~ N(0,1)[256,512] row-normalised; mask all true."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    E = torch.nn.functional.normalize(torch.randn(256, feat_dim, generator=g), dim=1)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    seg = (xs // 32 + 7 * (ys // 32)) % 256
    return E.to(device), seg.to(device)
