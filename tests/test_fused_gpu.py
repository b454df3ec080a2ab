"""Fused blend + decode (lsv2_render_decode) vs the CPU checker.

Integer / fp32-blend outputs (n_contrib, final_T, RGB) must match the checker exactly as the unfused kernel does:
the per-pixel alpha / T arithmetic is the same.  The weight map is accumulated on the tensor cores from alpha*T and
coefficients that are each split into an fp16 head + fp16 remainder (three MMAs, fp32 accumulate in TMEM): stated
tolerance max-abs <= 1e-4 of the per-pixel norm of the weight vector (north_star's fp32-blend tolerance).  The
512-d features come from an fp16 x fp16 -> fp32 contraction of the Gram-normalised weights: stated tolerance
max-abs <= 1e-3 on the unit-norm feature (components ~ 0.04; SURVEY.md 8c allows 4e-3 for the 16-bit decode),
cosine >= 0.99999 against the checker's fp32 decode of its own fp32 weight map.
"""
import numpy as np
import pytest
import torch

import oracle
from langsplatv2_b200 import render_decode
from langsplatv2_b200.renderer import FeatureRenderer
from tests import helpers as hp

pytestmark = pytest.mark.gpu


BG = (0.1, 0.2, 0.3)
W_TOL = 1e-4          # relative to the per-pixel norm of the weight vector (floored at 1e-3)
F_TOL = 1e-3


def _wmap_err(w, ref):
    norm = np.sqrt((ref ** 2).sum(0, keepdims=True))
    return float(np.max(np.abs(w - ref) / np.maximum(norm, 1e-3)))


def _check_blend_state(st, fwd):
    img = st.image
    np.testing.assert_array_equal(hp.u32(img["n_contrib"]), fwd["n_contrib"])
    np.testing.assert_array_equal(img["final_T"].cpu().numpy().view(np.uint32), fwd["final_T"].view(np.uint32))
    np.testing.assert_allclose(img["color"].cpu().numpy(), fwd["color"], rtol=0, atol=1e-5)


def _check_features(out, fwd_wmap, cb, H, W, normalize=True, tol=F_TOL):
    levels = cb.shape[0]
    ref = oracle.decode(fwd_wmap.reshape(levels * 64, -1), cb, normalize=normalize).reshape(levels, 512, H, W)
    got = out.float().cpu().numpy()
    scale = 1.0 if normalize else max(1.0, float(np.abs(ref).max()))
    assert np.max(np.abs(got - ref)) <= tol * scale
    if normalize:
        live = (np.abs(ref).sum(1) > 0)
        cos = (got * ref).sum(1)[live]
        assert cos.min() >= 0.99999
        assert np.all(got.transpose(0, 2, 3, 1)[~live] == 0.0)         # empty pixels: 0 / (0 + 1e-10)


@pytest.mark.parametrize("N,W,H,seed", [(4000, 144, 100, 5), (30000, 320, 240, 6), (60000, 250, 131, 7)])
def test_fused_quick_matches_checker(N, W, H, seed):
    scene, cam = hp.make_case(N, W, H, seed=seed, levels=3, scale_boost=2.0)
    _, wq, iq = hp.coeffs(scene, k=4)
    pre, binned, fwd = oracle.forward(hp.oracle_scene(scene), hp.oracle_view(cam), mode=2, L=192,
                                      weights_quick=wq, indices_quick=iq, bg=BG)
    st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
    out = render_decode(st, scene["codebooks"].cuda(), normalize=True, want_wmap=True)
    torch.cuda.synchronize()
    _check_blend_state(st, fwd)
    assert _wmap_err(st.image["wmap"].cpu().numpy(), fwd["wmap"]) <= W_TOL
    _check_features(out, fwd["wmap"], scene["codebooks"].numpy(), H, W)


def test_fused_equals_unfused_blend_state_and_is_deterministic():
    scene, cam = hp.make_case(20000, 256, 192, seed=1, levels=3, scale_boost=2.0)
    _, wq, iq = hp.coeffs(scene, k=4)
    un = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq)
    outs = []
    for _ in range(2):
        st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
        outs.append(render_decode(st, scene["codebooks"].cuda()))
        assert "wmap" not in st.image                                  # the weight map never exists in HBM
        for k in ("color", "final_T", "n_contrib"):
            assert torch.equal(st.image[k], un.image[k]), k
    assert torch.equal(outs[0], outs[1])


def test_fused_many_frames_bit_identical():
    """The kernel's hand-offs are mbarrier phases between six warp roles; a hand-off that can be overtaken shows up as
    a rare wrong pixel, not as a wrong frame (scripts/fused_stress.py found one in round 2 at 1 frame in 10).  300
    frames of a scene with long and short tile lists must reproduce the first bit for bit."""
    scene, cam = hp.make_case(150000, 640, 360, seed=4, levels=3, scale_boost=1.5)
    _, wq, iq = hp.coeffs(scene, k=4)
    st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
    cb = scene["codebooks"].cuda()
    first = render_decode(st, cb).clone()
    bad = 0
    for _ in range(300):
        bad += int(not torch.equal(render_decode(st, cb), first))
    assert bad == 0, f"{bad} of 300 frames differ from the first"


def test_fused_sparse_scene_with_empty_tiles_and_no_normalize():
    scene, cam = hp.make_case(300, 200, 120, seed=3, levels=3, scale_boost=1.0)
    _, wq, iq = hp.coeffs(scene, k=4)
    pre, binned, fwd = oracle.forward(hp.oracle_scene(scene), hp.oracle_view(cam), mode=2, L=192,
                                      weights_quick=wq, indices_quick=iq, bg=BG)
    lens = binned["ranges"][:, 1] - binned["ranges"][:, 0]
    assert (lens == 0).any() and (lens > 0).any()
    for normalize in (True, False):
        st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
        out = render_decode(st, scene["codebooks"].cuda(), normalize=normalize, want_wmap=True)
        _check_blend_state(st, fwd)
        assert _wmap_err(st.image["wmap"].cpu().numpy(), fwd["wmap"]) <= W_TOL
        _check_features(out, fwd["wmap"], scene["codebooks"].numpy(), 120, 200, normalize=normalize, tol=2e-3)


@pytest.mark.parametrize("levels", [1, 2])
def test_fused_dense_mode_levels(levels):
    """include_feature (dense [P, L] coefficients, L = levels*64) through the compacted rows."""
    N, W, H = 12000, 208, 144
    scene, cam = hp.make_case(N, W, H, seed=9, levels=3, scale_boost=2.0)
    dense, _, _ = hp.coeffs(scene, k=4)
    d = np.ascontiguousarray(dense[:, :64 * levels])
    pre, binned, fwd = oracle.forward(hp.oracle_scene(scene), hp.oracle_view(cam), mode=1, L=64 * levels,
                                      coeff_dense=d, bg=BG)
    st = hp.run_cuda(scene, cam, BG, hp.MODE_DENSE, dense=d, render=False)
    cb = scene["codebooks"][:levels].contiguous()
    out = render_decode(st, cb.cuda(), normalize=True, want_wmap=True)
    _check_blend_state(st, fwd)
    assert _wmap_err(st.image["wmap"].cpu().numpy(), fwd["wmap"]) <= W_TOL
    _check_features(out, fwd["wmap"], cb.numpy(), H, W)


def test_fused_duplicate_and_bad_indices():
    """quick-mode entries that repeat a channel add up; indices outside [0, L) are skipped and flagged."""
    scene, cam = hp.make_case(5000, 160, 112, seed=4, levels=3, scale_boost=2.0)
    _, wq, iq = hp.coeffs(scene, k=4)
    iq = iq.copy()
    iq[:, 1] = iq[:, 0]                      # duplicate channel in every row
    iq[::7, 5] = 500.0                       # out of range
    iq[::11, 6] = -3.0
    ref_iq = iq.copy()
    pre, binned, fwd = oracle.forward(hp.oracle_scene(scene), hp.oracle_view(cam), mode=2, L=192,
                                      weights_quick=wq, indices_quick=ref_iq, bg=BG)
    st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
    out = render_decode(st, scene["codebooks"].cuda(), want_wmap=True)
    _check_blend_state(st, fwd)
    assert _wmap_err(st.image["wmap"].cpu().numpy(), fwd["wmap"]) <= W_TOL
    assert int(st.status[2].item()) == 1
    _check_features(out, fwd["wmap"], scene["codebooks"].numpy(), 112, 160, tol=2e-3)


def test_fused_bf16_output_and_renderer_paths_agree():
    scene, cam = hp.make_case(20000, 256, 160, seed=8, levels=3, scale_boost=2.0)
    cam_d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()}
    r_f = FeatureRenderer(scene, topk=4, fused=True)
    r_u = FeatureRenderer(scene, topk=4, fused=False, precision="fp32")
    assert r_f.fused
    a = r_f.render(cam_d)
    b = r_u.render(cam_d)
    assert a.shape == b.shape == (3, 512, 160, 256)
    assert float((a - b).abs().max()) <= F_TOL
    r_h = FeatureRenderer(scene, topk=4, fused=True, out_dtype=torch.bfloat16)
    h = r_h.render(cam_d)
    assert h.dtype == torch.bfloat16
    assert float((h.float() - a).abs().max()) <= 1e-3          # bf16 rounding of |x| < 0.25
    assert torch.equal(r_f.last_state.image["color"], r_u.last_state.image["color"])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_entry_points_follow_the_tensors_device():
    """liblsv2 launches on the CURRENT device; every Python entry point switches to the device its tensors live on.
    Render on cuda:1 while cuda:0 is current and compare with the same frame rendered on cuda:0."""
    from langsplatv2_b200 import decode_feature_map
    scene, cam = hp.make_case(8000, 160, 96, seed=5, levels=3, scale_boost=2.0)
    _, wq, iq = hp.coeffs(scene, k=4)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        torch.cuda.set_device(0)
        st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, device=dev)
        two = decode_feature_map(st.image["wmap"], scene["codebooks"].to(dev), precision="fp32")
        fused = render_decode(st, scene["codebooks"].to(dev))
        torch.cuda.synchronize(dev)
        assert fused.device == torch.device(dev) and torch.cuda.current_device() == 0
        outs.append((fused.cpu(), two.cpu(), st.image["n_contrib"].cpu()))
    assert torch.equal(outs[0][2], outs[1][2]) and torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_renderer_overflow_is_sticky_and_capacity_grows():
    """FeatureRenderer with bounded instance buffers: a frame that overflows is flagged on the device and the flag
    survives later frames that fit, until check_overflow() reads it once, grows the capacity and raises."""
    from langsplatv2_b200 import synth, _abi
    from langsplatv2_b200.renderer import FeatureRenderer
    scene = synth.make_scene(30000, seed=0, levels=3)
    scene["scaling"] = scene["scaling"] * 2.0
    small = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.ring_camera(0, 8, 160, 96).items()}
    big = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.ring_camera(0, 8, 640, 384).items()}
    r = FeatureRenderer(scene, topk=4, fused=True)
    r.calibrate([small])
    cap0 = r.capacity
    r.render(small)
    r.check_overflow()                                   # fits: silent
    r.render(big)                                        # needs far more instances than the small view's capacity
    r.render(small)                                      # a later frame that fits must not hide it
    with pytest.raises(_abi.Lsv2Error):
        r.check_overflow()
    assert r.capacity > cap0 and r.overflow_events == 1
    r.render(small)
    r.check_overflow()                                   # the flag was cleared by the read
