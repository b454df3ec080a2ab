"""CPU suite (`-m "not gpu"`): the checker (oracle/) against the golden vectors generated from the
reference's own Python (tests/golden/make_golden.py), internal consistency of the checker, the host-side
mirror of the reference operator API, and that liblsv2.so loads and exports the whole C ABI."""
import math
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

import oracle
from langsplatv2_b200 import _abi, dist as lsdist, synth
from langsplatv2_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, MODE_DENSE, MODE_QUICK,
                                         MODE_RGB, feature_mode)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "reference_python.npz"))


# ------------------------------------------------------------------------------------------------
# golden vectors: utils/vq_utils.py
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("k", [1, 3, 4])
def test_golden_sparse_coefficients(k):
    logits = G["vq_logits"]
    dense = oracle.topk_soft_code(logits, k)
    ref = G[f"vq_dense_k{k}"]
    # rows 0/1 hold ties (which tied channel survives is torch.topk's choice): compare the tie-free rows
    np.testing.assert_allclose(dense[3:], ref[3:], rtol=2e-6, atol=1e-9)
    assert np.array_equal(dense[3:] != 0, ref[3:] != 0)
    assert ((dense != 0).sum(1) == k).all()
    np.testing.assert_allclose(dense.sum(1), 1.0, rtol=1e-5)
    w, i = oracle.weights_and_indices(logits[3:], k)
    assert np.array_equal(i, G[f"vq_i_k{k}"])                      # ascending channel order, float-typed
    np.testing.assert_allclose(w, G[f"vq_w_k{k}"], rtol=2e-6, atol=1e-9)


def test_sparse_coefficient_ties_are_valid_topk():
    logits = G["vq_logits"][:3]
    for k in (1, 3, 4):
        d = oracle.topk_soft_code(logits, k)
        assert ((d != 0).sum(1) == k).all()
        np.testing.assert_allclose(d.sum(1), 1.0, rtol=1e-5)
        # row 1: the 8 tied maxima hold all the survivors
        assert (d[1, 8:] == 0).all()


# ------------------------------------------------------------------------------------------------
# golden vectors: decode (eval_lerf.py:214-218, scene/gaussian_model.py:533-549)
# ------------------------------------------------------------------------------------------------
def test_golden_decode():
    cb, wmap = G["dec_codebooks"], G["dec_wmap"]
    raw = oracle.decode(wmap, cb, normalize=False)
    nrm = oracle.decode(wmap, cb, normalize=True)
    scale = np.abs(G["dec_quick_raw"]).max()
    assert np.abs(raw - G["dec_quick_raw"]).max() <= 2e-6 * scale
    assert np.abs(nrm - G["dec_quick_norm"]).max() <= 2e-6
    assert np.all(nrm[:, :, 0, 0] == 0.0)                           # empty pixel: 0 / (0 + 1e-10)
    # compute_final_feature_map = sum over levels of the raw decode; layer 0 = level 0
    assert np.abs(raw.sum(0) - G["dec_final"]).max() <= 4e-6 * scale
    assert np.abs(raw[0] - G["dec_layer0"]).max() <= 2e-6 * scale


def test_decode_backward_matches_autograd():
    rng = np.random.default_rng(0)
    cb = rng.standard_normal((2, 64, 512)).astype(np.float32)
    w = (rng.random((128, 37)) * (rng.random((128, 37)) < 0.1)).astype(np.float32)
    dF = rng.standard_normal((2, 512, 37)).astype(np.float32)
    dW, dC = oracle.decode_backward(w, cb, dF)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    ct = torch.tensor(cb, dtype=torch.float64, requires_grad=True)
    F = torch.einsum('ldk,lkn->ldn', ct.permute(0, 2, 1), wt.view(2, 64, 37))
    (F * torch.tensor(dF, dtype=torch.float64)).sum().backward()
    np.testing.assert_allclose(dW, wt.grad.numpy(), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(dC, ct.grad.numpy(), rtol=1e-6, atol=1e-6)


# ------------------------------------------------------------------------------------------------
# golden vectors: cameras (utils/graphics_utils.py:38-71, scene/cameras.py:55-58)
# ------------------------------------------------------------------------------------------------
def test_golden_cameras():
    for v in range(5):
        W, H = (int(x) for x in G["cam_wh"][v])
        fovx, fovy = (float(x) for x in G["cam_fov"][v])
        cam = synth.camera_from_pose(G["cam_R"][v], G["cam_t"][v], W, H, fovx, fovy)
        np.testing.assert_allclose(cam["world_view_transform"].numpy(), G["cam_world_view"][v], rtol=0, atol=1e-6)
        np.testing.assert_allclose(cam["full_proj_transform"].numpy(), G["cam_full_proj"][v], rtol=0, atol=2e-6)
        np.testing.assert_allclose(cam["camera_center"].numpy(), G["cam_center"][v], rtol=0, atol=2e-6)
        proj = synth.projection_matrix(synth.ZNEAR, synth.ZFAR, fovx, fovy).T.numpy()
        np.testing.assert_allclose(proj, G["cam_proj"][v], rtol=1e-6, atol=1e-7)
        assert cam["tanfovx"] == math.tan(fovx * 0.5)               # gaussian_renderer/__init__.py:34-35


def _wide_view():
    """A camera that sees every golden point: golden pose 0, very wide image so no tile rectangle is empty."""
    v = 0
    fovx, fovy = (float(x) for x in G["cam_fov"][v])
    return synth.camera_from_pose(G["cam_R"][v], G["cam_t"][v], 4096, 4096, math.radians(150), math.radians(150))


# ------------------------------------------------------------------------------------------------
# golden vectors: cov3D (utils/general_utils.py:78-110 + scene/gaussian_model.py:28-32)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mod", [1.0, 0.7])
def test_golden_cov3d(mod):
    scales, rots = G["cov_scales"], G["cov_rots"]
    P = scales.shape[0]
    cam = _wide_view()
    rng = np.random.default_rng(1)
    xyz = (rng.random((P, 3)) * 2 - 1).astype(np.float32)
    pre = oracle.preprocess(xyz, np.full(P, 0.5, np.float32), cam["world_view_transform"].numpy(),
                            cam["full_proj_transform"].numpy(), cam["camera_center"].numpy(), 4096, 4096,
                            cam["tanfovx"], cam["tanfovy"], scales=scales, rotations=rots,
                            colors_precomp=np.zeros((P, 3), np.float32), scale_modifier=mod)
    vis = pre["depth"] > 0
    assert vis.sum() >= P // 2
    ref = G[f"cov_cov3d_mod{mod}"]
    np.testing.assert_allclose(pre["cov3d"][vis], ref[vis], rtol=2e-5, atol=1e-9)


# ------------------------------------------------------------------------------------------------
# golden vectors: SH -> RGB (utils/sh_utils.py:57-112; +0.5 / clamp of gaussian_renderer/__init__.py:77-81)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_golden_sh(deg):
    xyz, sh, campos = G["sh_xyz"], G["sh_coeffs"], G["sh_campos"]
    P = xyz.shape[0]
    cam = _wide_view()
    pre = oracle.preprocess(xyz, np.full(P, 0.5, np.float32), cam["world_view_transform"].numpy(),
                            cam["full_proj_transform"].numpy(), campos, 4096, 4096, cam["tanfovx"], cam["tanfovy"],
                            scales=np.full((P, 3), 0.05, np.float32),
                            rotations=np.tile(np.array([1, 0, 0, 0], np.float32), (P, 1)), shs=sh, sh_degree=deg)
    vis = pre["radii"] > 0
    assert vis.sum() >= P // 2
    ref = G[f"sh_rgb_deg{deg}"]
    np.testing.assert_allclose(pre["rgb"][vis], ref[vis], rtol=0, atol=3e-6)
    assert np.array_equal(pre["clamped"][vis].astype(bool), (ref[vis] == 0.0))


def test_golden_losses_numpy_restatement():
    """cos_loss / l1_loss (utils/loss_utils.py:18-25) are what the feature stage differentiates; the training
    harness computes them in torch -- this pins the formula against the reference's value."""
    a, b = torch.tensor(G["loss_a"]), torch.tensor(G["loss_b"])
    cos = 1 - torch.nn.functional.cosine_similarity(a, b, dim=0).mean()
    assert abs(float(cos) - float(G["loss_cos"])) <= 1e-6
    assert abs(float((a - b).abs().mean()) - float(G["loss_l1"])) <= 1e-6


# ------------------------------------------------------------------------------------------------
# checker self-consistency (rasterizer core; "parity unpinned" rows R6-R8, R10)
# ------------------------------------------------------------------------------------------------
def _small_case(N=3000, W=128, H=96, seed=3):
    scene = synth.make_scene(N, seed=seed, levels=3)
    scene["scaling"] = scene["scaling"] * 2.0
    cam = synth.ring_camera(1, 8, W, H)
    osc = dict(means3d=scene["xyz"].numpy(), opacities=scene["opacity"].numpy(), scales=scene["scaling"].numpy(),
               rotations=scene["rotation"].numpy(), shs=scene["features"].numpy(), sh_degree=3)
    view = dict(viewmatrix=cam["world_view_transform"].numpy(), projmatrix=cam["full_proj_transform"].numpy(),
                campos=cam["camera_center"].numpy(), W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])
    return scene, osc, view


def test_oracle_binning_invariants():
    scene, osc, view = _small_case()
    W, H = view["W"], view["H"]
    pre, b, fwd = oracle.forward(osc, view, mode=0)
    R = b["R"]
    assert R == int(pre["tiles_touched"].sum()) and R > 0
    keys = b["keys"]
    assert np.all(keys[1:] >= keys[:-1])                            # sortedness
    # stable: equal keys keep emission order (emission index grows with the Gaussian id)
    order = np.argsort(b["keys_unsorted"], kind="stable")
    assert np.array_equal(b["point_list"], b["vals_unsorted"][order])
    gx, gy = oracle.grid(W, H)
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    for t in range(gx * gy):
        s, e = b["ranges"][t]
        assert np.all(tiles[s:e] == t)
        assert e - s == int((tiles == t).sum())
    # depth bits of the key are the fp32 depth of the Gaussian
    dbits = pre["depth"].view(np.uint32)[b["point_list"]]
    assert np.array_equal((keys & np.uint64(0xFFFFFFFF)).astype(np.uint32), dbits)
    # n_contrib never exceeds the tile's list length
    nc = fwd["n_contrib"]
    for ty in range(gy):
        for tx in range(gx):
            s, e = b["ranges"][ty * gx + tx]
            assert nc[ty * 16:(ty + 1) * 16, tx * 16:(tx + 1) * 16].max() <= e - s


def test_oracle_quick_equals_dense_and_linearity():
    scene, osc, view = _small_case()
    logits = scene["logits"].numpy()
    dense, ws, idx = [], [], []
    for l in range(3):
        lg = logits[:, l * 64:(l + 1) * 64]
        dense.append(oracle.topk_soft_code(lg, 4))
        w, i = oracle.weights_and_indices(lg, 4)
        ws.append(w)
        idx.append(i + 64 * l)
    dense, wq, iq = np.concatenate(dense, 1), np.concatenate(ws, 1), np.concatenate(idx, 1)
    _, _, fq = oracle.forward(osc, view, mode=2, L=192, weights_quick=wq, indices_quick=iq)
    _, _, fd = oracle.forward(osc, view, mode=1, L=192, coeff_dense=dense)
    assert np.array_equal(fq["n_contrib"], fd["n_contrib"])
    np.testing.assert_allclose(fq["wmap"], fd["wmap"], rtol=0, atol=2e-6)
    # per level the weights of a pixel sum to its accumulated opacity 1 - T (rows of coefficients sum to 1)
    acc = 1.0 - fq["final_T"]
    for l in range(3):
        np.testing.assert_allclose(fq["wmap"][l * 64:(l + 1) * 64].sum(0), acc, rtol=0, atol=2e-4)
    # linearity of the weight map in the coefficients
    _, _, f2 = oracle.forward(osc, view, mode=2, L=192, weights_quick=2.0 * wq, indices_quick=iq)
    np.testing.assert_allclose(f2["wmap"], 2.0 * fq["wmap"], rtol=1e-6, atol=1e-7)


def test_oracle_matches_torch_fp64_restatement():
    """oracle/torch_ref.py is an independent fp64 restatement (pure PyTorch, BASELINE.json config 1's form)."""
    from oracle import torch_ref
    scene, osc, view = _small_case(N=800, W=64, H=48)
    dense = np.concatenate([oracle.topk_soft_code(scene["logits"].numpy()[:, :64], 4)], 1)
    pre, b, fwd = oracle.forward(osc, view, mode=1, L=64, coeff_dense=dense)
    W, H = view["W"], view["H"]
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    tp = torch_ref.preprocess(t64(osc["means3d"]), t64(osc["opacities"]), t64(osc["scales"]), t64(osc["rotations"]),
                              t64(view["viewmatrix"]), t64(view["projmatrix"]), t64(view["campos"]), W, H,
                              view["tanfovx"], view["tanfovy"], shs=t64(osc["shs"]), sh_degree=3)
    assert np.array_equal(tp["radii"].numpy(), pre["radii"])
    tb = torch_ref.bin_instances(tp, W, H)
    assert np.array_equal(tb["point_list"].numpy().astype(np.uint32), b["point_list"])
    out = torch_ref.render(tp, tb, W, H, torch.zeros(3), coeff=t64(dense))
    np.testing.assert_allclose(out["color"].numpy(), fwd["color"], rtol=0, atol=3e-5)
    np.testing.assert_allclose(out["wmap"].numpy(), fwd["wmap"], rtol=0, atol=3e-5)
    # thresholds may flip on the last ulp between fp64 torch.exp and the fp32 polynomial: allow a few pixels
    assert (out["n_contrib"].numpy() != fwd["n_contrib"].astype(np.int32)).mean() < 0.01


def test_oracle_expf_accuracy():
    x = -np.abs(np.random.default_rng(0).standard_normal(20000).astype(np.float32)) * 20
    y = oracle.expf(x)
    ref = np.exp(x.astype(np.float64))
    ok = ref > 1e-37
    assert np.max(np.abs(y[ok] - ref[ok]) / ref[ok]) < 3e-7


# ------------------------------------------------------------------------------------------------
# C ABI: the library loads and exports what include/lsv2.h declares (no compute without a GPU)
# ------------------------------------------------------------------------------------------------
def test_abi_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "lsv2.h")).read()
    declared = sorted(set(re.findall(r"\b(lsv2_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 15
    lib = _abi.lib()
    for name in declared:
        assert hasattr(lib, name), f"liblsv2.so does not export {name}"
    assert set(_abi.EXPORTS) <= set(declared)
    assert lib.lsv2_abi_version() == _abi.LSV2_ABI_VERSION
    m = re.search(r"#define LSV2_ABI_VERSION (\d+)", header)
    assert int(m.group(1)) == _abi.LSV2_ABI_VERSION


def test_abi_host_only_entry_points():
    lib = _abi.lib()
    assert lib.lsv2_strerror(0) == b"ok"
    assert b"invalid" in lib.lsv2_strerror(1)
    assert lib.lsv2_sort_scratch_bytes(0, 0) > 0
    assert lib.lsv2_sort_scratch_bytes(10_000_000, 1_000_000) >= lib.lsv2_sort_scratch_bytes(1000, 1000)
    assert lib.lsv2_decode_scratch_bytes(3, 64, 512) >= 3 * 512 * 128
    # argument validation happens before any CUDA call
    assert lib.lsv2_preprocess(None, None, None, None, None) == 1
    assert lib.lsv2_bin(None, 0, None, None, None) == 1
    assert lib.lsv2_render_forward(None, None, None, None, None, None) == 1
    assert lib.lsv2_decode(3, 64, 512, 100, None, None, 1, 1, 0, None, None, 0, None) == 1
    assert lib.lsv2_topk_soft_code(-1, 64, 4, None, None, None, None, None) == 1


def test_abi_struct_layout_matches_header():
    """ctypes mirrors of the C structs: sizes under the LP64 ABI the header implies."""
    import ctypes as C
    assert C.sizeof(_abi.View) == 24 + 4 * 8
    assert C.sizeof(_abi.Scene) == 8 + 7 * 8
    assert C.sizeof(_abi.Geom) == 11 * 8
    assert C.sizeof(_abi.Binning) == 9 * 8
    assert C.sizeof(_abi.Image) == 4 * 8


# ------------------------------------------------------------------------------------------------
# host-side mirror of the reference operator API
# ------------------------------------------------------------------------------------------------
def _settings(**kw):
    base = dict(image_height=32, image_width=48, tanfovx=0.5, tanfovy=0.4, bg=torch.zeros(3), scale_modifier=1.0,
                viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=3, campos=torch.zeros(3),
                prefiltered=False, debug=False, include_feature=False, quick_render=False)
    base.update(kw)
    return GaussianRasterizationSettings(**base)


def test_settings_keywords_match_reference_call_site():
    """gaussian_renderer/__init__.py:37-52 constructs the settings with exactly these keywords."""
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "include_feature", "quick_render")
    import diff_gaussian_rasterization as shim
    assert shim.GaussianRasterizer is GaussianRasterizer


def test_mode_comes_from_flags_not_shapes():
    assert feature_mode(_settings()) == MODE_RGB
    assert feature_mode(_settings(include_feature=True)) == MODE_DENSE
    assert feature_mode(_settings(include_feature=True, quick_render=True)) == MODE_QUICK   # eval_lerf.sh:23-25
    assert feature_mode(_settings(quick_render=True)) == MODE_QUICK


def test_argument_combination_errors_match_upstream():
    r = GaussianRasterizer(_settings())
    P = 4
    m, o = torch.zeros(P, 3), torch.zeros(P, 1)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=o, scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=o, shs=torch.zeros(P, 16, 3), colors_precomp=torch.zeros(P, 3),
          scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=o, shs=torch.zeros(P, 16, 3))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=o, shs=torch.zeros(P, 16, 3), scales=torch.ones(P, 3),
          rotations=torch.ones(P, 4), cov3D_precomp=torch.zeros(P, 6))


def test_no_cpu_fallback():
    """CPU tensors must raise: the product has no CPU path."""
    r = GaussianRasterizer(_settings())
    P = 4
    with pytest.raises(_abi.Lsv2Error, match="no CPU path"):
        r(means3D=torch.zeros(P, 3), means2D=torch.zeros(P, 3), opacities=torch.zeros(P, 1),
          shs=torch.zeros(P, 16, 3), scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
    from langsplatv2_b200 import decode_feature_map, get_weights_and_indices
    with pytest.raises(_abi.Lsv2Error):
        decode_feature_map(torch.zeros(64, 4, 4), torch.zeros(1, 64, 512))
    with pytest.raises(_abi.Lsv2Error):
        get_weights_and_indices(torch.zeros(8, 64), 4)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "langsplatv2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, re.M), f"{f} imports the oracle"
                assert "lsv2_oracle" not in src, f"{f} references the oracle library"


# ------------------------------------------------------------------------------------------------
# view sharding + gradient bucket (world_size 2, gloo)
# ------------------------------------------------------------------------------------------------
def test_shard_views_partition():
    for V, world in ((64, 1), (64, 2), (7, 4), (3, 8)):
        owned = [lsdist.shard_views(V, r, world) for r in range(world)]
        flat = sorted(v for o in owned for v in o)
        assert flat == list(range(V))
    with pytest.raises(ValueError):
        lsdist.shard_views(4, 2, 2)
    # a world of B ranks covers per step what accum_iter=B covers in B iterations
    assert [lsdist.step_views(5, 64, r, 4) for r in range(4)] == [20, 21, 22, 23]


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
from langsplatv2_b200 import dist as lsdist
rank, world, _ = lsdist.world_info()
dist.init_process_group("gloo", rank=rank, world_size=world)
g = torch.Generator().manual_seed(7)
logits = torch.randn(50, 64, generator=g, requires_grad=True)
codebook = torch.randn(1, 64, 16, generator=g, requires_grad=True)
targets = torch.randn(4, 50, 16, generator=g)
def loss_for(view):
    w = torch.softmax(logits, 1)
    return ((w @ codebook[0] - targets[view]) ** 2).mean()
# data parallel: this rank's view of step 0
view = lsdist.step_views(0, 4, rank, world)
loss_for(view).backward()
bucket = lsdist.GradBucket([logits, codebook])
bucket.sync()
got = [logits.grad.clone(), codebook.grad.clone()]
# reference: single process accumulating `world` views (train.py --accum_iter semantics: grads SUM)
logits.grad = None; codebook.grad = None
for v in range(world):
    loss_for(v).backward()
ok = all(torch.allclose(a, b.grad, rtol=1e-5, atol=1e-7) for a, b in zip(got, [logits, codebook]))
assert bucket.nbytes == (50 * 64 + 64 * 16) * 4
# attached mode (what FeatureTrainer uses): .grad tensors are views of the bucket, sync() is the all-reduce alone
logits.grad = None; codebook.grad = None
b2 = lsdist.GradBucket([logits, codebook]); b2.attach()
loss_for(view).backward()
assert logits.grad.data_ptr() == b2.flat.data_ptr()
b2.sync()
ok = ok and all(torch.allclose(a, b.grad, rtol=1e-5, atol=1e-7) for a, b in zip(got, [logits, codebook]))
b2.zero()
ok = ok and float(logits.grad.abs().sum()) == 0.0
# sliced reduction (FeatureTrainer at N > 1: the top-k backward hands d logits over in row blocks, each block's
# all-reduce starts as soon as it is written; the codebook part goes with finish_overlap)
loss_for(view).backward()
b2.begin_overlap()
n_log = logits.numel()
for a in range(0, n_log, 1024):
    b2.reduce_slice(a, min(n_log, a + 1024))
b2.finish_overlap(n_log)
ok = ok and all(torch.allclose(a, b.grad, rtol=1e-5, atol=1e-7) for a, b in zip(got, [logits, codebook]))
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 3)
"""


def test_grad_bucket_allreduce_equals_accum_iter_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=180)
        assert p.returncode == 0, out.decode()[-2000:]


# ------------------------------------------------------------------------------------------------
# drop-in boundary: the reference's unmodified render() against this repo's operator signature
# ------------------------------------------------------------------------------------------------
RENDER_CALLS = os.path.join(ROOT, "tests", "golden", "render_calls.json")


def test_recorded_render_calls_bind_to_our_operator():
    """Every call recorded from the reference's gaussian_renderer/__init__.py::render (tests/golden/
    make_render_calls.py) binds to GaussianRasterizer.forward / GaussianRasterizationSettings of the shim package the
    reference imports (gaussian_renderer/__init__.py:15), keyword for keyword; the GPU replay is tests/test_dropin_gpu.py."""
    import inspect
    import json
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    rec = json.load(open(RENDER_CALLS))
    assert {c["branch"] for c in rec["calls"]} >= {"rgb", "include_feature", "quick"}
    sig = inspect.signature(GaussianRasterizer.forward)
    for c in rec["calls"]:
        sig.bind(None, **{k: None for k in c["kwargs_order"]})
        assert tuple(c["settings_order"]) == GaussianRasterizationSettings._fields
        assert c["returned_keys"] == ["language_feature_weight_map", "radii", "render", "viewspace_points", "visibility_filter"]
        # placeholders are 1-element float tensors, never None (gaussian_renderer/__init__.py:93-103)
        for k in ("language_feature_precomp", "language_feature_weights_quick", "language_feature_indices"):
            assert c["kwargs"][k] is not None


@pytest.mark.skipif(not os.path.isdir("/root/reference/gaussian_renderer"), reason="the reference is only mounted in the build container")
def test_recorded_render_calls_are_current():
    """Re-runs the recorder on the reference file as it lies on disk and compares with the committed fixture."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("make_render_calls", os.path.join(ROOT, "tests", "golden", "make_render_calls.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    live = json.loads(json.dumps(m.record_calls(), sort_keys=True))
    assert live == json.load(open(RENDER_CALLS))


_WORKER_SHARDED = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
from langsplatv2_b200 import dist as lsdist
rank, world, _ = lsdist.world_info()
dist.init_process_group("gloo", rank=rank, world_size=world)
g = torch.Generator().manual_seed(3)
n_log, n_cb = 64 * 96, 64 * 16
flat_p0 = torch.randn(n_log + n_cb, generator=g)
grads = torch.randn(world, n_log + n_cb, generator=g)          # rank r's local gradient = grads[r]
blocks = [(0, 64 * 32), (64 * 32, 64 * 64), (64 * 64, n_log), (n_log, n_log + n_cb)]
lr = 0.1
# scheme A (round 1): all-reduce, every rank steps everything
pa = flat_p0.clone(); ga = grads[rank].clone()
dist.all_reduce(ga, op=dist.ReduceOp.SUM)
pa -= lr * ga
# scheme B: reduce-scatter -> step this rank's shard -> all-gather the parameters
pb = flat_p0.clone(); gb = grads[rank].clone()
ex = lsdist.ShardedExchange(gb, pb, blocks)
for i in range(len(blocks)):
    ex.reduce_scatter(i)
for i in range(len(blocks)):
    ex.wait_grad(i)
    lo, hi = ex.shard_range(i)
    pb[lo:hi] -= lr * ex.shard_grad[i]
    ex.all_gather_params(i)
ex.finish()
ok = torch.allclose(pa, pb, rtol=1e-6, atol=1e-7)
ok = ok and torch.allclose(ex.full_reduced_grad(), grads.sum(0), rtol=1e-6, atol=1e-6)
# every rank holds the same parameters bit for bit (each element was computed by exactly one rank)
ref = pb.clone(); dist.broadcast(ref, 0)
ok = ok and torch.equal(ref, pb)
# scheme C (the trainer's default): two iterations with the all-gather DEFERRED to the start of the next one
# (all_gather_all + finish before the parameters are read), against two all-reduce iterations
pc = flat_p0.clone(); gc = grads[rank].clone()
pd = flat_p0.clone()
exc = lsdist.ShardedExchange(gc, pc, blocks)
for it in range(2):
    exc.all_gather_all(); exc.finish()                       # previous iteration's shards (a no-op on the first)
    gc.copy_(grads[rank] * (1.0 + it) + 0.01 * pc)           # this iteration's gradient depends on the parameters
    gd = grads[rank] * (1.0 + it) + 0.01 * pd
    dist.all_reduce(gd, op=dist.ReduceOp.SUM)
    pd -= lr * gd
    for i in range(len(blocks)):
        exc.reduce_scatter(i)
    for i in range(len(blocks)):
        exc.wait_grad(i)
        lo, hi = exc.shard_range(i)
        pc[lo:hi] -= lr * exc.shard_grad[i]
exc.all_gather_all(); exc.finish()                           # == FeatureTrainer.sync_params()
ok = ok and torch.allclose(pc, pd, rtol=1e-5, atol=1e-6)
# scheme D (the trainer's exchange since the row form): the logits' gradient is J(x)^T d_coeff with the Jacobian of
# softmax -> top-k -> renormalise (utils/vq_utils.py:9-24), the same on every rank, and d_coeff lives on the top-k
# support only.  The ranks reduce-scatter d_coeff in ROW form ([N, k] instead of [N, 64]) over a second ShardedExchange
# whose blocks are the same row blocks, then apply J^T to their own shard of rows only.  Must equal the all-reduce of the
# dense d logits; the row shards of the two exchanges must coincide.
C, K, N = 64, 4, n_log // 64
def soft_code(x):
    p = torch.softmax(x, dim=1)
    idx = torch.topk(p, K, dim=1).indices.sort(dim=1).values
    m = torch.zeros_like(p).scatter_(1, idx, 1.0)
    y = p * m
    return y / (y.sum(1, keepdim=True) + 1e-10), idx
x0 = flat_p0[:n_log].view(N, C).clone()
gen = torch.Generator().manual_seed(5)
d_rows_all = torch.randn(world, N, K, generator=gen)             # rank r's blend-backward output, row form
xr = x0.clone().requires_grad_(True)
y, idx = soft_code(xr)
dense_g = torch.zeros(N, C).scatter_(1, idx, d_rows_all[rank])   # the same gradient as a dense support-only array
dx = torch.autograd.grad(y, xr, dense_g)[0]
dist.all_reduce(dx, op=dist.ReduceOp.SUM)
p_ref = x0 - lr * dx                                             # all-reduce of d logits, every rank steps everything
pe = flat_p0.clone()
row_blocks = [(a // C, b // C) for a, b in blocks[:-1]]
pex = lsdist.ShardedExchange(torch.zeros_like(pe), pe, blocks)   # parameters (all-gather) -- gradient side unused here
rows_flat = d_rows_all[rank].reshape(-1).clone()
rex = lsdist.ShardedExchange(rows_flat, None, [(a * K, b * K) for a, b in row_blocks])
for i in range(len(row_blocks)):
    rex.reduce_scatter(i)
for i in range(len(row_blocks)):
    rex.wait_grad(i)
    lo, hi = pex.shard_range(i)
    r0, r1 = lo // C, hi // C
    ok = ok and (r0 * K, r1 * K) == rex.shard_range(i)           # the two exchanges shard the same rows
    xs = x0[r0:r1].clone().requires_grad_(True)
    ys, idxs = soft_code(xs)
    gs = torch.zeros(r1 - r0, C).scatter_(1, idxs, rex.shard_grad[i].view(-1, K))
    pe[lo:hi] -= lr * torch.autograd.grad(ys, xs, gs)[0].reshape(-1)
    pex.all_gather_params(i)
pex.finish()
ok = ok and torch.allclose(pe[:n_log].view(N, C), p_ref, rtol=1e-5, atol=1e-6)
try:
    lsdist.ShardedExchange(gb, pb, [(0, 6)])
    ok = False
except ValueError:
    pass
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 3)
"""


def test_sharded_exchange_equals_allreduce_gloo(tmp_path):
    """dist.ShardedExchange (reduce-scatter -> sharded step -> all-gather of the parameters) gives every rank the
    parameters the all-reduce scheme gives, on 2 gloo ranks -- with the dense gradient, with the deferred gather, and
    with the coefficient gradient exchanged in row form ahead of the (linear) top-k backward."""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker_sharded.py"
    script.write_text(_WORKER_SHARDED.format(root=ROOT))
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=180)
        assert p.returncode == 0, out.decode()[-2000:]


# ------------------------------------------------------------------------------------------------
# golden vectors: relevancy query (eval/openclip_encoder.py:114-138), generated by running the reference
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_golden_relevancy_query(tag):
    Q = np.load(os.path.join(ROOT, "tests", "golden", "reference_query.npz"))
    sem = Q[f"{tag}_sem"]                                    # [levels, H, W, 512] as the reference takes it
    embeds = np.concatenate([Q[f"{tag}_pos"], Q[f"{tag}_neg"]], 0)
    got = oracle.relevancy(np.ascontiguousarray(sem.transpose(0, 3, 1, 2)), embeds, Q[f"{tag}_pos"].shape[0])
    assert got.shape == Q[f"{tag}_relev"].shape
    np.testing.assert_allclose(got, Q[f"{tag}_relev"], rtol=0, atol=2e-6)
    assert np.allclose(got[0, :, 0, 0], 0.5)                 # the empty pixel: every similarity is 0


# ---------------------------------------------------------------------------------------------
# distCUDA2: the CPU restatement against outputs of the reference's own compiled SimpleKNN::knn
# (tests/golden/knn_golden.npz, tests/golden/make_knn_golden.py -- generated on a GPU box)
def test_golden_knn_restatement_matches_the_reference():
    from oracle import knn_ref
    z = np.load(os.path.join(ROOT, "tests", "golden", "knn_golden.npz"))
    names = sorted({k.split("/")[0] for k in z.files})
    assert {"uniform_3000", "clusters_5000", "duplicates_1500", "plane_2049", "far_offset_1025", "tiny_1",
            "tiny_4"} <= set(names)
    for n in names:
        pts, want = z[n + "/points"], z[n + "/dist2"]
        for fn in (knn_ref.dist_cuda2_bruteforce, knn_ref.dist_cuda2):
            got = fn(pts)
            assert got.dtype == np.float32 and got.shape == want.shape
            fin = np.isfinite(want)
            assert np.array_equal(np.isinf(got), ~fin), n
            # the fp32 fma is emulated through float64: at most the last bit (the fixture run saw none differ)
            assert np.all(np.abs(got[fin] - want[fin]) <= np.spacing(want[fin])), n
    # duplicates count as neighbours at distance 0; fewer than three other points leave FLT_MAX in the sum:
    # two others -> FLT_MAX / 3, fewer -> +inf
    assert z["duplicates_1500/dist2"][305] == 0.0
    assert np.all(z["tiny_3/dist2"] == np.float32(np.finfo(np.float32).max) / np.float32(3))
    assert np.all(np.isinf(z["tiny_2/dist2"])) and np.all(np.isinf(z["tiny_1/dist2"]))
    # the k-d tree path of the checker against its brute force beyond the brute-force cut-over
    pts = np.random.default_rng(5).normal(0, 1, (6000, 3)).astype(np.float32)
    pts[10:30] = pts[10]                          # more coincident points than the tree returns candidates
    assert np.array_equal(knn_ref.dist_cuda2(pts), knn_ref.dist_cuda2_bruteforce(pts))
