"""distCUDA2: lsv2_dist2_knn3 through the C-ABI against the reference's outputs (tests/golden/knn_golden.npz,
generated on a GPU box from the reference's own compiled SimpleKNN::knn), the compiled reference itself where
oracle/_ref/libsimple_knn_ref.so travelled with the tree, and the CPU restatement oracle/knn_ref.py."""
import os

import numpy as np
import pytest
import torch

from oracle import knn_ref

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "knn_golden.npz")


def _ours(pts: np.ndarray) -> np.ndarray:
    from langsplatv2_b200.knn import distCUDA2
    return distCUDA2(torch.from_numpy(pts).cuda()).cpu().numpy()


def _bit_equal(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_golden_vectors_of_the_reference_bit_for_bit():
    z = np.load(GOLDEN)
    names = sorted({k.split("/")[0] for k in z.files})
    assert len(names) >= 10
    for n in names:
        pts, want = z[n + "/points"], z[n + "/dist2"]
        got = _ours(pts)
        assert got.shape == want.shape and got.dtype == np.float32
        assert _bit_equal(got, want), (n, int((got != want).sum()))
    # fewer than three other points: the reference's FLT_MAX stays in the sum (FLT_MAX / 3 with two others, +inf below)
    assert np.all(z["tiny_3/dist2"] > 1e38) and np.all(np.isinf(z["tiny_2/dist2"])) and np.all(z["tiny_4/dist2"] < 1e3)


def test_against_the_compiled_reference_when_it_travelled():
    if knn_ref.reference_lib() is None:
        pytest.skip("oracle/_ref/libsimple_knn_ref.so not built")
    g = np.random.default_rng(11)
    c = g.normal(0, 2, (40, 3))
    pts = (c[g.integers(0, 40, 300_001)] + g.normal(0, 0.2, (300_001, 3)) *
           g.uniform(0.05, 1.0, (300_001, 1))).astype(np.float32)
    want = knn_ref.reference_dist_cuda2(torch.from_numpy(pts).cuda()).cpu().numpy()
    got = _ours(pts)
    assert _bit_equal(got, want), int((got != want).sum())


def test_against_the_cpu_restatement_and_edge_cases():
    g = np.random.default_rng(3)
    pts = g.uniform(-2, 2, (20_000, 3)).astype(np.float32)
    pts[::50] = pts[1::50]                                   # duplicates
    want = knn_ref.dist_cuda2_bruteforce(pts)
    got = _ours(pts)
    assert np.all(np.abs(got - want) <= np.spacing(want))    # the checker's fma goes through float64: 1 ulp
    assert (got == want).mean() > 0.999
    # empty, non-contiguous, float64 and wrong shapes
    from langsplatv2_b200.knn import distCUDA2
    assert distCUDA2(torch.zeros((0, 3), device="cuda")).shape == (0,)
    wide = torch.from_numpy(np.concatenate([pts, pts], 1)).cuda()
    assert torch.equal(distCUDA2(wide[:, 3:]), torch.from_numpy(got).cuda())
    assert torch.equal(distCUDA2(torch.from_numpy(pts).double().cuda()), torch.from_numpy(got).cuda())
    with pytest.raises(ValueError):
        distCUDA2(torch.zeros((5, 2), device="cuda"))
    with pytest.raises(Exception):
        distCUDA2(torch.zeros((5, 3)))
    # the reference's import line resolves to the same function
    from simple_knn._C import distCUDA2 as shim
    assert shim is distCUDA2


def test_full_size_scene_matches_kdtree_and_is_permutation_invariant():
    """1M points of the bench's scene: exact 3-NN by a k-d tree (CPU, float64 search, fp32 re-evaluation), and the
    size-independent property that reordering the input reorders the output bit for bit."""
    from langsplatv2_b200 import synth
    pts = synth.make_scene(1_000_000, seed=0, levels=1)["xyz"].float().numpy()
    got = _ours(pts)
    want = knn_ref.dist_cuda2(pts)
    assert np.all(np.abs(got - want) <= np.spacing(want))
    assert (got == want).mean() > 0.999
    perm = np.random.default_rng(0).permutation(pts.shape[0])
    assert _bit_equal(_ours(np.ascontiguousarray(pts[perm])), got[perm])
    # what the caller makes of it (gaussian_model.py:194-195): finite log-scales
    s = np.log(np.sqrt(np.maximum(got, 1e-7)))
    assert np.isfinite(s).all()


def test_scene_from_point_cloud_and_level_merge(tmp_path):
    """create_from_pcd (gaussian_model.py:184-210) on this package's distCUDA2, rendered through the op; and the
    per-level checkpoints of an evaluation merged as eval_lerf.py:333-346 does, against that expression in torch."""
    from langsplatv2_b200 import scene_io, synth, sparse
    from langsplatv2_b200.rasterizer import rasterize_forward
    from tests import helpers as hp
    g = np.random.default_rng(9)
    xyz = g.uniform(-1.5, 1.5, (5000, 3)).astype(np.float32)
    rgb = g.integers(0, 256, (5000, 3))
    p = tmp_path / "points3D.ply"
    scene_io.write_point_cloud(str(p), xyz, rgb)
    pos, col, _ = scene_io.read_point_cloud(str(p))
    s = scene_io.scene_from_point_cloud(pos, col)
    d2 = np.maximum(knn_ref.dist_cuda2(xyz), np.float32(1e-7))
    want = np.log(np.sqrt(d2))
    np.testing.assert_allclose(s["log_scale"][:, 0].cpu().numpy(), want, rtol=0, atol=2e-6)
    assert torch.equal(s["log_scale"][:, 0], s["log_scale"][:, 2])
    assert torch.allclose(s["opacity"], torch.full_like(s["opacity"], 0.1), atol=1e-6)
    assert torch.equal(s["rotation"][:, 0], torch.ones(5000, device="cuda")) and not s["features"][:, 1:].any()
    np.testing.assert_allclose(s["features"][:, 0].cpu().numpy(), (rgb / 255.0 - 0.5) / 0.28209479177387814, atol=1e-6)
    # the scene renders: SH DC alone reproduces the point colours where a splat covers a pixel
    _, cam = hp.make_case(10, 128, 96, seed=1, levels=1)
    rs = hp.settings(cam, dict(sh_degree=3), (0.0, 0.0, 0.0), include_feature=False, quick_render=False)
    st = rasterize_forward(rs, s["xyz"], s["opacity"], shs=s["features"], scales=s["scaling"], rotations=s["rotation"])
    img = st.image["color"]
    assert torch.isfinite(img).all() and float(img.max()) > 0.05 and int((st.geom["radii"] > 0).sum()) > 1000

    # three per-level checkpoints -> one quick-render scene
    levels = []
    for lvl in range(3):
        sc = synth.make_scene(5000, seed=20 + lvl, levels=1, device="cuda")
        path = tmp_path / f"lvl{lvl}.pth"
        torch.save(scene_io.checkpoint_tuple(sc, 10000), str(path))
        levels.append(scene_io.read_checkpoint(str(path), device="cuda"))
    merged = scene_io.combine_levels(levels, topk=4)
    assert merged["codebooks"].shape == (3, 64, 512) and merged["weights_quick"].shape == (5000, 12)
    for lvl in range(3):
        w, i = sparse.get_weights_and_indices(levels[lvl]["logits"], 4)
        assert torch.equal(merged["weights_quick"][:, 4 * lvl:4 * lvl + 4], w)
        assert torch.equal(merged["indices_quick"][:, 4 * lvl:4 * lvl + 4], i + 64.0 * lvl)
        assert torch.equal(merged["codebooks"][lvl], levels[lvl]["codebooks"][0])
    assert merged["indices_quick"].dtype == torch.float32 and float(merged["indices_quick"].max()) <= 191
    # ... which is a scene the resident renderer takes as it is: checkpoints on disk -> normalised 512-d map
    from langsplatv2_b200.renderer import FeatureRenderer
    r = FeatureRenderer(merged)
    assert torch.equal(r.weights, merged["weights_quick"]) and torch.equal(r.indices, merged["indices_quick"])
    cam_gpu = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()}
    feat = r.render(cam_gpu)
    assert feat.shape == (3, 512, 96, 128) and torch.isfinite(feat).all()
    norms = feat.norm(dim=1)
    covered = norms > 0.5
    assert int(covered.sum()) > 100 and torch.allclose(norms[covered], torch.ones_like(norms[covered]), atol=5e-3)
