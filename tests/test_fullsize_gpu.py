"""The hot path at BASELINE.json's full sizes (1M Gaussians; 1008x756 = config 2, 1920x1080 = the metric's size).

The checker cannot blend a whole frame of this size in seconds, so the full-size parity is:
  * preprocess + binning of the WHOLE frame against the checker, bit for bit (radii, tiles_touched, instance count,
    sorted Gaussian ids, tile ranges);
  * blend + decode of every `stride`-th tile against the checker (n_contrib bit-exact, features within the fused
    path's stated tolerance 1e-3 on the unit-norm feature);
  * size-independent properties over the whole frame: sorted keys, ranges partition the instance list, a checksum of
    the Gaussian ids survives the sort, n_contrib <= list length, final_T in [0, 1], unit feature norms where a
    splat contributed and exact zeros elsewhere, and the fused kernel's blend state equals the unfused kernel's.
"""
import numpy as np
import pytest
import torch

import oracle
from langsplatv2_b200 import render_decode
from tests import helpers as hp

pytestmark = pytest.mark.gpu
BG = (0.0, 0.0, 0.0)


@pytest.mark.parametrize("W,H,stride", [(1008, 756, 97), (1920, 1080, 211)])
def test_full_size_frame(W, H, stride):
    N = 1_000_000
    scene, cam = hp.make_case(N, W, H, seed=0, levels=3, view=0, views=64)
    _, wq, iq = hp.coeffs(scene, k=4)
    osc, view = hp.oracle_scene(scene), hp.oracle_view(cam)

    # ---- checker: whole-frame front end, sampled tiles for the blend ----
    pre = oracle.preprocess(osc["means3d"], osc["opacities"], view["viewmatrix"], view["projmatrix"], view["campos"],
                            W, H, view["tanfovx"], view["tanfovy"], scales=osc["scales"], rotations=osc["rotations"],
                            shs=osc["shs"], sh_degree=3)
    binned = oracle.bin_instances(pre, W, H)
    R = binned["R"]
    fwd = oracle.render_forward(pre, binned, W, H, BG, mode=2, L=192, weights_quick=wq, indices_quick=iq,
                                tile_stride=stride, tile_phase=3)
    mask = oracle.tile_pixel_mask(W, H, stride, 3)

    # ---- CUDA path ----
    st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
    out = render_decode(st, scene["codebooks"].cuda(), normalize=True)
    torch.cuda.synchronize()
    assert st.num_rendered == R
    np.testing.assert_array_equal(st.geom["radii"].cpu().numpy(), pre["radii"])
    np.testing.assert_array_equal(hp.u32(st.geom["tiles_touched"]), pre["tiles_touched"])
    pl = hp.u32(st.binning["point_list"])[:R]
    np.testing.assert_array_equal(pl, binned["point_list"])
    ranges = hp.u32(st.binning["ranges"])
    np.testing.assert_array_equal(ranges, binned["ranges"])

    # ---- properties over the whole frame ----
    keys = hp.ref_keys(st, R)
    assert np.all(keys[1:] >= keys[:-1]), "keys not sorted"
    assert int(pre["tiles_touched"].astype(np.int64).sum()) == R
    vis = np.nonzero(pre["tiles_touched"])[0]
    assert int(pl.astype(np.int64).sum()) == int((vis.astype(np.int64) * pre["tiles_touched"][vis]).sum()), \
        "id checksum changed across the sort"
    lens = ranges[:, 1].astype(np.int64) - ranges[:, 0]
    nz = lens > 0
    assert lens.sum() == R and np.all(ranges[nz][1:, 0] == ranges[nz][:-1, 1]), "ranges do not partition the list"
    tile_of_key = (keys >> np.uint64(32)).astype(np.int64)
    assert np.array_equal(np.bincount(tile_of_key, minlength=lens.size), lens)
    ncon = hp.u32(st.image["n_contrib"]).astype(np.int64)
    gx = (W + 15) // 16
    ty, tx = np.meshgrid(np.arange(H) // 16, np.arange(W) // 16, indexing="ij")
    assert np.all(ncon <= lens[ty * gx + tx])
    T = st.image["final_T"].cpu().numpy()
    assert T.min() >= 0.0 and T.max() <= 1.0 and np.all(T[ncon == 0] == 1.0)
    nrm = out.float().norm(dim=1).cpu().numpy()                       # [3, H, W]
    live = ncon > 0
    assert np.all(np.abs(nrm[:, live] - 1.0) <= 2e-3)
    assert np.all(nrm[:, ~live] == 0.0)

    # ---- sampled tiles against the checker ----
    np.testing.assert_array_equal(ncon[mask], fwd["n_contrib"][mask].astype(np.int64))
    np.testing.assert_array_equal(T[mask].view(np.uint32), fwd["final_T"][mask].view(np.uint32))
    ref = oracle.decode(np.ascontiguousarray(fwd["wmap"][:, mask]), scene["codebooks"].numpy(), normalize=True)
    got = out[:, :, torch.as_tensor(mask, device=out.device)].float().cpu().numpy()
    assert np.max(np.abs(got - ref)) <= 1e-3

    # ---- the unfused kernel on the same frame: identical blend state ----
    del out
    un = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq)
    for k in ("color", "final_T", "n_contrib"):
        assert torch.equal(st.image[k], un.image[k]), k
    wm = un.image["wmap"][:, torch.as_tensor(mask, device="cuda")].cpu().numpy()
    assert np.max(np.abs(wm - fwd["wmap"][:, mask])) <= 1e-5
    # conservation over the whole frame: each level's top-k coefficients sum to 1 per Gaussian, so the 64 channels of
    # a level sum to the blended opacity sum_i alpha_i T_i = 1 - T_final
    lvl = un.image["wmap"].view(3, 64, H, W).sum(dim=1)
    assert float((lvl - (1.0 - un.image["final_T"]).unsqueeze(0)).abs().max()) <= 2e-4


def test_config4_forward_backward_adjoint_identity():
    """BASELINE.json config 4 (4M Gaussians, 1600x1066, forward + backward).  The blend is linear in the coefficients:
    wmap = J c, so the backward must satisfy <J c, dW> == <c, J^T dW> -- checked in float64 on the device, together with
    linearity of the backward in dW."""
    from langsplatv2_b200.rasterizer import rasterize_backward
    N, W, H = 4_000_000, 1600, 1066
    scene, cam = hp.make_case(N, W, H, seed=4, levels=1, view=5, views=64)
    dense = torch.zeros((N, 64), dtype=torch.float32)
    g = torch.Generator(device="cpu").manual_seed(9)
    idx = torch.randint(0, 64, (N, 4), generator=g)
    dense.scatter_(1, idx, torch.rand((N, 4), generator=g))
    st = hp.run_cuda(scene, cam, BG, hp.MODE_DENSE, dense=dense, debug_buffers=False)
    assert not st.overflowed
    wmap = st.image["wmap"]
    dW = torch.randn(wmap.shape, generator=torch.Generator(device="cuda").manual_seed(1), device="cuda")
    d_coeff, _ = rasterize_backward(st, dW, dense_grad=True)
    lhs = (wmap.double() * dW.double()).sum()
    rhs = (d_coeff.double() * dense.cuda().double()).sum()
    assert abs(float(lhs - rhs)) <= 1e-4 * abs(float(lhs)) + 1e-3, (float(lhs), float(rhs))
    d2, _ = rasterize_backward(st, 2.0 * dW, dense_grad=True)
    assert float((d2 - 2.0 * d_coeff).abs().max()) <= 2e-4 * float(d_coeff.abs().max())
    keys = torch.as_tensor(hp.ref_keys(st).view(np.int64))
    assert bool((keys[1:] >= keys[:-1]).all())


def test_config5_4k_render_properties():
    """BASELINE.json config 5 (camera-sharded sweep up to 4K): one 3840x2160 frame of 2M Gaussians through the fused
    kernel; unit norms where a splat contributed, zeros elsewhere, sorted keys, ranges partition the list."""
    N, W, H = 2_000_000, 3840, 2160
    scene, cam = hp.make_case(N, W, H, seed=6, levels=3, view=9, views=64)
    logits = scene["logits"].cuda()
    from langsplatv2_b200 import get_weights_and_indices
    ws, idx = [], []
    for l in range(3):
        w, i = get_weights_and_indices(logits[:, l * 64:(l + 1) * 64], 4)
        ws.append(w)
        idx.append(i + l * 64)
    st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=torch.cat(ws, 1), iq=torch.cat(idx, 1), render=False,
                     debug_buffers=False)
    out = render_decode(st, scene["codebooks"].cuda(), normalize=True)
    R = st.num_rendered
    keys = torch.as_tensor(hp.ref_keys(st, R).view(np.int64))
    assert bool((keys[1:] >= keys[:-1]).all())
    ranges = st.binning["ranges"].to(torch.int64)
    lens = ranges[:, 1] - ranges[:, 0]
    assert int(lens.sum()) == R
    ncon = st.image["n_contrib"].to(torch.int64)
    live = ncon > 0
    for l in range(3):
        nrm = out[l].float().pow(2).sum(0).sqrt()
        assert float((nrm[live] - 1.0).abs().max()) <= 2e-3
        assert float(nrm[~live].abs().sum()) == 0.0
