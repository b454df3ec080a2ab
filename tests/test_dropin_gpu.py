"""The drop-in claim, replayed: every way the reference's UNMODIFIED ``gaussian_renderer/__init__.py::render`` calls
the op -- recorded by ``tests/golden/make_render_calls.py`` from the file as it lies in /root/reference (which does
not exist on the GPU box) into ``tests/golden/render_calls.json`` -- goes through ``diff_gaussian_rasterization``
(the shim package of this repo) with the same keyword set, None-ness, placeholder shapes, dtypes and requires_grad
flags, and the three return values are checked against the CPU checker.  On top of the quick branch the body of
``eval_lerf.py:210-220`` (view / einsum / norm / divide) is run as the reference writes it and compared with the
checker's decode and with the fused path.
"""
import json
import os

import numpy as np
import pytest
import torch

import oracle
from tests import helpers as hp

pytestmark = pytest.mark.gpu

BG = (0.1, 0.2, 0.3)
CALLS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "render_calls.json")))


def _build_kwargs(spec, order, rec_n, scene, dense, wq, iq, cov3d, colors):
    """Tensors for one recorded call: same keyword set / None-ness / placeholder shapes as the recording."""
    N = scene["xyz"].shape[0]
    src = dict(means3D=scene["xyz"], means2D=torch.zeros_like(scene["xyz"]), shs=scene["features"],
               colors_precomp=colors, language_feature_precomp=dense, language_feature_weights_quick=wq,
               language_feature_indices=iq, opacities=scene["opacity"].reshape(N, 1), scales=scene["scaling"],
               rotations=scene["rotation"], cov3D_precomp=cov3d)
    out = {}
    for name in order:                               # the keyword order of gaussian_renderer/__init__.py:108-119
        s = spec[name]
        if s is None:
            out[name] = None
            continue
        if s["shape"] == [1]:                       # the reference's 1-element float placeholder
            t = torch.zeros((1,), dtype=getattr(torch, s["dtype"]))
        else:
            t = torch.as_tensor(src[name]).to(getattr(torch, s["dtype"]))
            want = [N if d == rec_n else d for d in s["shape"]]
            assert list(t.shape) == want, (name, list(t.shape), want)
        t = t.cuda()
        if s["requires_grad"]:
            t.requires_grad_(True)
        out[name] = t
    return out


@pytest.mark.parametrize("call", CALLS["calls"], ids=[c["branch"] for c in CALLS["calls"]])
def test_reference_render_call_replayed_through_the_shim(call):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer   # the import of :15
    N, W, H = 6000, 176, 112
    scene, cam = hp.make_case(N, W, H, seed=31, levels=3, scale_boost=2.5)
    dense_all, wq, iq = hp.coeffs(scene, k=4)
    dense = np.ascontiguousarray(dense_all[:, :64])
    osc, ov = hp.oracle_scene(scene), hp.oracle_view(cam)
    kwspec = call["kwargs"]
    # the python-side fallbacks of render() :65-85 hand the op precomputed covariances / colours
    cov3d = colors = None
    if kwspec["cov3D_precomp"] is not None:          # what pc.get_covariance() returns: the symmetric 3x3 as 6 floats
        from oracle import torch_ref
        S = torch_ref.cov3d(scene["scaling"].float(), scene["rotation"].float(), 1.0)
        cov3d = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).numpy()
    if kwspec["colors_precomp"] is not None:
        colors = np.random.default_rng(3).random((N, 3), dtype=np.float32)
    kw = _build_kwargs(kwspec, call["kwargs_order"], CALLS["N"], scene, dense, wq, iq, cov3d, colors)
    assert list(kw.keys()) == call["kwargs_order"]
    s = call["settings"]
    vals = dict(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                bg=torch.tensor(BG, dtype=torch.float32, device="cuda"), scale_modifier=s["scale_modifier"],
                viewmatrix=cam["world_view_transform"].cuda(), projmatrix=cam["full_proj_transform"].cuda(),
                sh_degree=s["sh_degree"], campos=cam["camera_center"].cuda(), prefiltered=s["prefiltered"],
                debug=s["debug"], include_feature=s["include_feature"], quick_render=s["quick_render"])
    rs = GaussianRasterizationSettings(**{k: vals[k] for k in call["settings_order"]})   # keyword-constructed, :37-52
    rasterizer = GaussianRasterizer(raster_settings=rs)
    with torch.no_grad() if s["quick_render"] else torch.enable_grad():
        rendered_image, wmap, radii = rasterizer(**kw)                                   # :108-119
    mode = 2 if s["quick_render"] else (1 if s["include_feature"] else 0)
    okw = dict(mode=mode, bg=BG)
    if mode == 2:
        okw.update(L=192, weights_quick=wq, indices_quick=iq)
    elif mode == 1:
        okw.update(L=64, coeff_dense=dense)
    if cov3d is not None:
        osc = dict(osc, scales=None, rotations=None, cov3d_precomp=cov3d)
    if colors is not None:
        osc = dict(osc, shs=None, colors_precomp=colors)
    pre, binned, fwd = oracle.forward(osc, ov, **okw)
    np.testing.assert_array_equal(radii.cpu().numpy(), pre["radii"])
    assert radii.dtype == torch.int32
    assert bool(((radii > 0).cpu().numpy() == (pre["radii"] > 0)).all())                 # visibility_filter, :126
    np.testing.assert_allclose(rendered_image.detach().cpu().numpy(), fwd["color"], rtol=0, atol=1e-5)
    if mode:
        w = wmap.detach().cpu().numpy()
        assert w.shape == (192 if mode == 2 else 64, H, W)
        norm = np.sqrt((fwd["wmap"] ** 2).sum(0, keepdims=True))
        assert np.max(np.abs(w - fwd["wmap"]) / np.maximum(norm, 1e-3)) <= 1e-4
    if mode == 1:
        # train.py:173: gradients reach language_feature_precomp (strict per-channel gradient at the op boundary)
        dW = np.random.default_rng(1).standard_normal(w.shape).astype(np.float32)
        (wmap * torch.tensor(dW).cuda()).sum().backward()
        ref = oracle.render_backward_coeff(pre, binned, fwd, W, H, 64, dW, dense=True, coeff_dense=dense)
        g = kw["language_feature_precomp"].grad.cpu().numpy()
        assert np.max(np.abs(g - ref)) <= 1e-4 * np.abs(ref).max() + 1e-6
    if mode == 2:
        # eval_lerf.py:214-218, as written there
        language_feature_weight_map = wmap
        D_, H_, W_ = language_feature_weight_map.shape
        language_feature_weight_map = language_feature_weight_map.view(3, 64, H_, W_).view(3, 64, H_ * W_)
        language_codebooks = scene["codebooks"].cuda().permute(0, 2, 1)
        language_feature_map = torch.einsum('ldk,lkn->ldn', language_codebooks, language_feature_weight_map).view(3, 512, H_, W_)
        language_feature_map = language_feature_map / (language_feature_map.norm(dim=1, keepdim=True) + 1e-10)
        ref = oracle.decode(fwd["wmap"].reshape(192, -1), scene["codebooks"].numpy(), normalize=True).reshape(3, 512, H, W)
        assert np.max(np.abs(language_feature_map.cpu().numpy() - ref)) <= 2e-4
        # ... and the one-kernel replacement of those five lines
        from langsplatv2_b200.renderer import FeatureRenderer
        fr = FeatureRenderer(scene, topk=4)
        fused = fr.render({k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()})
        assert float((fused - language_feature_map).abs().max()) <= 1e-3
