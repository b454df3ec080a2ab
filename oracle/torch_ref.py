"""Pure-PyTorch CPU evaluation of the rasterizer math (TEST INFRASTRUCTURE).

This is BASELINE.json config 1's "pure-PyTorch CPU evaluation" and the autograd
cross-check for the C oracle's analytic backward.  Same algorithm as
``lsv2_oracle.c`` (rows R6-R10 of SURVEY.md §8a; [LINEAGE] for the rasterizer core,
"parity unpinned"), vectorised per tile; it uses ``torch.exp`` so thresholds
(alpha < 1/255, T < 1e-4) can flip on the last ulp relative to the C oracle --
comparisons against it are tolerance-based, never bit-exact.
"""
from __future__ import annotations

import math

import torch

TILE = 16


def _xform(m, p, rows):
    # m: [4,4] transposed (row-vector) matrix as the reference passes it; p: [N,3]
    ph = torch.cat([p, torch.ones_like(p[:, :1])], dim=1)
    return (ph @ m)[:, :rows]


def cov3d(scales, rotations, mod=1.0):
    r, x, y, z = rotations.unbind(-1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=-1).view(-1, 3, 3)
    Lm = R * (mod * scales)[:, None, :]
    return Lm @ Lm.transpose(1, 2)


def eval_sh_rgb(deg, sh, pos, campos):
    """utils/sh_utils.py:57-112 + gaussian_renderer/__init__.py:77-81.  sh: [N, M, 3]."""
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
          -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
    d = pos - campos[None]
    d = d / d.norm(dim=1, keepdim=True)
    x, y, z = d[:, 0:1], d[:, 1:2], d[:, 2:3]
    res = C0 * sh[:, 0]
    if deg > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6]
                   + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
            if deg > 2:
                res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10]
                       + C3[2] * y * (4 * zz - xx - yy) * sh[:, 11]
                       + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12]
                       + C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14]
                       + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return torch.clamp_min(res + 0.5, 0.0)


def preprocess(means3d, opacities, scales, rotations, viewmatrix, projmatrix, campos, W, H,
               tanfovx, tanfovy, shs=None, colors_precomp=None, sh_degree=0, scale_modifier=1.0,
               cov3d_precomp=None):
    dt = means3d.dtype
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    pv = _xform(viewmatrix.to(dt), means3d, 3)
    ph = _xform(projmatrix.to(dt), means3d, 4)
    pw = 1.0 / (ph[:, 3] + 1e-7)
    ppx, ppy = ph[:, 0] * pw, ph[:, 1] * pw
    if cov3d_precomp is not None:
        c = cov3d_precomp
        Sig = torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4],
                           c[:, 2], c[:, 4], c[:, 5]], dim=-1).view(-1, 3, 3)
    else:
        Sig = cov3d(scales, rotations, scale_modifier)
    tz = pv[:, 2]
    tz_safe = torch.where(tz > 0.2, tz, torch.ones_like(tz))
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tx = torch.clamp(pv[:, 0] / tz_safe, -limx, limx) * tz_safe
    ty = torch.clamp(pv[:, 1] / tz_safe, -limy, limy) * tz_safe
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz_safe, zero, -(fx * tx) / (tz_safe * tz_safe),
                     zero, fy / tz_safe, -(fy * ty) / (tz_safe * tz_safe)], dim=-1).view(-1, 2, 3)
    Rv = viewmatrix.to(dt)[:3, :3].t()            # W2C rotation
    Mm = J @ Rv
    cov = Mm @ Sig @ Mm.transpose(1, 2)
    a, b, c_ = cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3
    det = a * c_ - b * b
    det_safe = torch.where(det != 0, det, torch.ones_like(det))
    conic = torch.stack([c_ / det_safe, -b / det_safe, a / det_safe], dim=-1)
    mid = 0.5 * (a + c_)
    lam = mid + torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
    rad = torch.ceil(3.0 * torch.sqrt(lam))
    px = ((ppx + 1.0) * W - 1.0) * 0.5
    py = ((ppy + 1.0) * H - 1.0) * 0.5
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    rminx = torch.clamp(torch.trunc((px - rad) / TILE), 0, gx).long()
    rminy = torch.clamp(torch.trunc((py - rad) / TILE), 0, gy).long()
    rmaxx = torch.clamp(torch.trunc((px + rad + TILE - 1) / TILE), 0, gx).long()
    rmaxy = torch.clamp(torch.trunc((py + rad + TILE - 1) / TILE), 0, gy).long()
    ntiles = (rmaxx - rminx) * (rmaxy - rminy)
    vis = (tz > 0.2) & (det != 0) & (ntiles > 0)
    if colors_precomp is not None:
        rgb = colors_precomp
    else:
        rgb = eval_sh_rgb(sh_degree, shs, means3d, campos.to(dt))
    return dict(vis=vis, radii=torch.where(vis, rad, torch.zeros_like(rad)).int(), xy=torch.stack([px, py], -1),
                depth=tz, conic=conic, opacity=opacities.reshape(-1), rgb=rgb,
                rect=torch.stack([rminx, rminy, rmaxx, rmaxy], -1), ntiles=torch.where(vis, ntiles, 0 * ntiles))


def bin_instances(pre, W, H):
    gx = (W + TILE - 1) // TILE
    vis_idx = torch.nonzero(pre["vis"]).squeeze(1)
    rect = pre["rect"][vis_idx]
    nt = pre["ntiles"][vis_idx]
    gid = torch.repeat_interleave(vis_idx, nt)
    start = torch.cumsum(nt, 0) - nt
    local = torch.arange(gid.numel()) - torch.repeat_interleave(start, nt)
    rw = torch.repeat_interleave(rect[:, 2] - rect[:, 0], nt)
    tx = torch.repeat_interleave(rect[:, 0], nt) + local % rw
    ty = torch.repeat_interleave(rect[:, 1], nt) + local // rw
    tile = ty * gx + tx
    dbits = pre["depth"].detach().float().contiguous().view(torch.int32).long()[gid]
    key = (tile << 32) | dbits
    key_sorted, order = torch.sort(key, stable=True)
    return dict(keys=key_sorted, point_list=gid[order], tile=tile[order])


def render(pre, binned, W, H, bg, coeff=None):
    """coeff: dense [N, L] (scatter quick-mode weights into it first) or None."""
    dt = pre["xy"].dtype
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    L = 0 if coeff is None else coeff.shape[1]
    color = torch.zeros(3, H, W, dtype=dt)
    wmap = torch.zeros(L, H, W, dtype=dt)
    final_T = torch.ones(H, W, dtype=dt)
    n_contrib = torch.zeros(H, W, dtype=torch.int32)
    tile_sorted = binned["tile"]
    bounds = torch.searchsorted(tile_sorted, torch.arange(gx * gy + 1))
    outs = {"color": [], "wmap": []}
    ly, lx = torch.meshgrid(torch.arange(TILE), torch.arange(TILE), indexing="ij")
    for t in range(gx * gy):
        s, e = int(bounds[t]), int(bounds[t + 1])
        ty_, tx_ = t // gx, t % gx
        ys = (ty_ * TILE + ly).reshape(-1)
        xs = (tx_ * TILE + lx).reshape(-1)
        inside = (ys < H) & (xs < W)
        ys, xs = ys[inside], xs[inside]
        if e > s:
            g = binned["point_list"][s:e]
            dx = pre["xy"][g, 0:1] - xs[None].to(dt)
            dy = pre["xy"][g, 1:2] - ys[None].to(dt)
            con = pre["conic"][g]
            power = -0.5 * (con[:, 0:1] * dx * dx + con[:, 2:3] * dy * dy) - con[:, 1:2] * dx * dy
            alpha_raw = pre["opacity"][g][:, None] * torch.exp(torch.clamp_max(power, 0.0))
            # the clamp at 0.99 is the identity in the derivative ([LINEAGE]: the 3DGS backward multiplies dL/dalpha
            # by o and G without looking at the clamp); same forward value
            alpha = alpha_raw + (torch.clamp_max(alpha_raw, 0.99) - alpha_raw).detach()
            valid = (power <= 0) & (alpha >= 1.0 / 255.0)
            a = torch.where(valid, alpha, torch.zeros_like(alpha))
            Tincl = torch.cumprod(1.0 - a, dim=0)
            Texcl = torch.cat([torch.ones_like(Tincl[:1]), Tincl[:-1]], 0)
            stop = torch.cumsum((valid & (Tincl < 1e-4)).int(), 0) > 0
            contrib = valid & ~stop
            w = torch.where(contrib, a * Texcl, torch.zeros_like(a))            # [G, px]
            Tfin = torch.where(contrib, Tincl, torch.ones_like(Tincl) * 2).amin(0).clamp_max(1.0)
            Tfin = torch.where(contrib.any(0), Tfin, torch.ones_like(Tfin))
            idx = torch.arange(1, e - s + 1)[:, None] * contrib
            ncon = idx.amax(0).int()
            col = pre["rgb"][g].t() @ w + Tfin[None] * bg.to(dt)[:, None]
            color[:, ys, xs] = col
            if L:
                wmap[:, ys, xs] = coeff[g].t() @ w
            final_T[ys, xs] = Tfin
            n_contrib[ys, xs] = ncon
        else:
            color[:, ys, xs] = bg.to(dt)[:, None].expand(3, ys.numel())
    return dict(color=color, wmap=wmap, final_T=final_T, n_contrib=n_contrib)


def decode(wmap, codebooks, normalize=True):
    """eval_lerf.py:214-218.  wmap [levels*Lc, H, W], codebooks [levels, Lc, D] -> [levels, D, H, W]."""
    levels, Lc, D = codebooks.shape
    H, W = wmap.shape[1:]
    f = torch.einsum('ldk,lkn->ldn', codebooks.permute(0, 2, 1), wmap.view(levels, Lc, H * W)).view(levels, D, H, W)
    if normalize:
        f = f / (f.norm(dim=1, keepdim=True) + 1e-10)
    return f


def topk_soft_code(logits, k):
    """utils/vq_utils.py:9-24 restated."""
    y = logits.softmax(dim=1)
    _, idx = torch.topk(y, k, dim=1)
    mask = torch.zeros_like(y, dtype=torch.bool).scatter_(1, idx, True)
    yt = torch.where(mask, y, torch.zeros_like(y))
    return yt / (yt.sum(dim=1, keepdim=True) + 1e-10)
