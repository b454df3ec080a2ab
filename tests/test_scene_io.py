"""Scene readers / writers: the PLY and checkpoint layouts of the reference
(/root/reference/scene/gaussian_model.py:67-138, 269-350; scene/dataset_readers.py:120-143), CPU only.  The fixtures
are packed here with ``struct`` following the reference's property order -- independent of the module's own writer."""
import struct

import numpy as np
import pytest
import torch

from langsplatv2_b200 import scene_io, synth


def _trained_ply_bytes(P, M, rng, fmt="binary_little_endian"):
    """A file as GaussianModel.save_ply lays it out: x y z nx ny nz f_dc_0..2 f_rest_0..3(M-1)-1 opacity scale_0..2
    rot_0..3, float32 each; f_dc / f_rest hold [P, 3, *] flattened (features.transpose(1, 2).flatten(1))."""
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + \
            [f"f_rest_{i}" for i in range(3 * (M - 1))] + ["opacity"] + [f"scale_{i}" for i in range(3)] + \
            [f"rot_{i}" for i in range(4)]
    vals = rng.normal(0, 1, (P, len(names))).astype(np.float32)
    vals[:, 3:6] = 0
    head = "ply\nformat %s 1.0\ncomment made by a test\nelement vertex %d\n" % (fmt, P)
    head += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    if fmt == "ascii":
        body = "".join(" ".join(repr(float(x)) for x in row) + "\n" for row in vals).encode()
    else:
        body = b"".join(struct.pack("<%df" % len(names), *row) for row in vals)
    return head.encode("ascii") + body, names, vals


@pytest.mark.parametrize("fmt", ["binary_little_endian", "ascii"])
def test_read_trained_ply_layout(tmp_path, fmt):
    rng = np.random.default_rng(0)
    P, M = 37, 16
    raw, names, vals = _trained_ply_bytes(P, M, rng, fmt)
    path = tmp_path / "point_cloud.ply"
    path.write_bytes(raw)
    s = scene_io.read_ply(str(path))
    c = {n: vals[:, j] for j, n in enumerate(names)}
    assert s["xyz"].shape == (P, 3) and s["features"].shape == (P, M, 3) and s["sh_degree"] == 3
    np.testing.assert_array_equal(s["xyz"].numpy(), np.stack([c["x"], c["y"], c["z"]], 1))
    # features[:, 0, ch] = f_dc_ch ; features[:, 1 + m, ch] = f_rest_{ch * (M - 1) + m}  (gaussian_model.py:316-328, 345)
    for ch in range(3):
        np.testing.assert_array_equal(s["features"][:, 0, ch].numpy(), c[f"f_dc_{ch}"])
        for m in (0, 7, M - 2):
            np.testing.assert_array_equal(s["features"][:, 1 + m, ch].numpy(), c[f"f_rest_{ch * (M - 1) + m}"])
    np.testing.assert_array_equal(s["opacity_logit"][:, 0].numpy(), c["opacity"])
    np.testing.assert_array_equal(s["log_scale"].numpy(), np.stack([c[f"scale_{i}"] for i in range(3)], 1))
    np.testing.assert_array_equal(s["rot_raw"].numpy(), np.stack([c[f"rot_{i}"] for i in range(4)], 1))
    # activations of the op's inputs (gaussian_model.py:28-40)
    torch.testing.assert_close(s["scaling"], torch.exp(s["log_scale"]))
    torch.testing.assert_close(s["opacity"], torch.sigmoid(s["opacity_logit"]))
    torch.testing.assert_close(s["rotation"].norm(dim=1), torch.ones(P))
    if fmt == "binary_little_endian":
        # the module's writer produces the same bytes after the header (and a header the reader takes back)
        out = tmp_path / "sub" / "again.ply"
        scene_io.write_ply(str(out), s)
        again = out.read_bytes()
        assert again.split(b"end_header\n", 1)[1] == raw.split(b"end_header\n", 1)[1]
        s2 = scene_io.read_ply(str(out))
        for k in ("xyz", "features", "log_scale", "rot_raw", "opacity_logit"):
            assert torch.equal(s[k], s2[k])


def test_ply_errors(tmp_path):
    rng = np.random.default_rng(1)
    raw, _, _ = _trained_ply_bytes(5, 16, rng)
    p = tmp_path / "a.ply"
    p.write_bytes(raw[:-7])
    with pytest.raises(ValueError, match="truncated"):
        scene_io.read_ply(str(p))
    p.write_bytes(raw)
    with pytest.raises(ValueError, match="f_rest"):
        scene_io.read_ply(str(p), max_sh_degree=2)                 # the reference asserts the count (:324)
    p.write_bytes(b"plx\n")
    with pytest.raises(ValueError, match="not a PLY"):
        scene_io.read_ply(str(p))
    with pytest.raises(ValueError, match="point-cloud"):
        p.write_bytes(raw)
        scene_io.read_point_cloud(str(p))


def test_point_cloud_ply_round_trip(tmp_path):
    """storePly / fetchPly (dataset_readers.py:120-143): uchar colours come back divided by 255."""
    rng = np.random.default_rng(2)
    xyz = rng.normal(0, 1, (50, 3)).astype(np.float32)
    rgb = rng.integers(0, 256, (50, 3))
    head = "ply\nformat binary_little_endian 1.0\nelement vertex 50\n" + \
           "".join(f"property float {n}\n" for n in ("x", "y", "z", "nx", "ny", "nz")) + \
           "".join(f"property uchar {n}\n" for n in ("red", "green", "blue")) + "end_header\n"
    body = b"".join(struct.pack("<6f3B", *xyz[i], 0, 0, 0, *rgb[i]) for i in range(50))
    p = tmp_path / "points3D.ply"
    p.write_bytes(head.encode() + body)
    pos, col, nrm = scene_io.read_point_cloud(str(p))
    np.testing.assert_array_equal(pos.astype(np.float32), xyz)
    np.testing.assert_allclose(col, rgb / 255.0)
    assert not nrm.any()
    q = tmp_path / "again.ply"
    scene_io.write_point_cloud(str(q), xyz, rgb)
    assert q.read_bytes().split(b"end_header\n", 1)[1] == body


@pytest.mark.parametrize("with_features", [True, False])
def test_checkpoint_tuple_round_trip(tmp_path, with_features):
    """train.py:267 saves (capture(), iteration); restore unpacks 14 fields with language features, 12 without
    (gaussian_model.py:67-138)."""
    src = synth.make_scene(200, seed=3, levels=1)
    P = 200
    par = lambda t: torch.nn.Parameter(t.clone())
    f_dc, f_rest = par(src["features"][:, :1]), par(src["features"][:, 1:])
    head = (2, par(src["xyz"]), f_dc, f_rest, par(src["log_scale"]), par(src["rot_raw"]), par(src["opacity_logit"]))
    tail = (torch.zeros(P), torch.ones(P, 1), torch.full((P, 1), 2.0), {"state": {}, "param_groups": []}, 4.5)
    mid = (par(src["logits"]), par(src["codebooks"])) if with_features else ()
    path = tmp_path / "chkpnt10000.pth"
    torch.save((head + mid + tail, 10000), str(path))
    s = scene_io.read_checkpoint(str(path))
    assert s["iteration"] == 10000 and s["active_sh_degree"] == 2 and s["sh_degree"] == 3
    assert s["spatial_lr_scale"] == 4.5 and float(s["denom"][0]) == 2.0
    for k in ("xyz", "features", "log_scale", "rot_raw", "opacity_logit", "scaling", "rotation", "opacity"):
        assert torch.equal(s[k], src[k]), k
        assert not s[k].requires_grad
    if with_features:
        assert torch.equal(s["logits"], src["logits"]) and torch.equal(s["codebooks"], src["codebooks"])
        assert (s["levels"], s["codebook_size"], s["feat_dim"]) == (1, 64, 512)
    else:
        assert "logits" not in s
    # and back: the pair the reference's restore() unpacks, field for field
    params, it = scene_io.checkpoint_tuple(s, 10000)
    assert it == 10000 and len(params) == (14 if with_features else 12)
    assert params[0] == 2 and torch.equal(params[2], f_dc.detach()) and torch.equal(params[3], f_rest.detach())
    assert torch.equal(params[4], src["log_scale"]) and params[-1] == 4.5
    s2 = scene_io.read_checkpoint((params, it))
    assert torch.equal(s2["features"], s["features"])
    with pytest.raises(ValueError, match="length 3"):
        scene_io.read_checkpoint(((1, 2, 3), 0))


def test_ply_big_endian_and_leading_element(tmp_path):
    """PLY variants plyfile would also read: big-endian payload, and a fixed-size element in front of the vertices."""
    rng = np.random.default_rng(4)
    xyz = rng.normal(0, 1, (9, 3)).astype(np.float32)
    rgb = rng.integers(0, 256, (9, 3))
    head = "ply\nformat binary_big_endian 1.0\nobj_info made by a test\nelement camera 2\nproperty float fx\n" \
           "property int id\nelement vertex 9\n" + \
           "".join(f"property float {n}\n" for n in ("x", "y", "z")) + \
           "".join(f"property uchar {n}\n" for n in ("red", "green", "blue")) + "end_header\n"
    body = struct.pack(">fi", 1.5, 7) + struct.pack(">fi", 2.5, 8) + \
        b"".join(struct.pack(">3f3B", *xyz[i], *rgb[i]) for i in range(9))
    p = tmp_path / "be.ply"
    p.write_bytes(head.encode() + body)
    pos, col, nrm = scene_io.read_point_cloud(str(p))
    np.testing.assert_array_equal(pos.astype(np.float32), xyz)
    np.testing.assert_allclose(col * 255.0, rgb)
    assert nrm.shape == (9, 3) and not nrm.any()                 # no nx/ny/nz in the file
    # a list property in the vertex element is not something the reference writes: refused, not misread
    bad = "ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nproperty list uchar int idx\nend_header\n0 1 5\n"
    p.write_bytes(bad.encode())
    with pytest.raises(ValueError, match="list property"):
        scene_io.read_point_cloud(str(p))


def test_checkpoint_scene_has_what_renderer_and_trainer_read():
    """The dictionary of read_checkpoint carries every key FeatureRenderer / FeatureTrainer take from a scene
    (renderer.py:35-44, trainer.py: xyz, opacity, scaling, rotation, features, logits, codebooks + the sizes)."""
    src = synth.make_scene(64, seed=1, levels=1)
    s = scene_io.read_checkpoint(scene_io.checkpoint_tuple(src, 7))
    for k in ("xyz", "opacity", "scaling", "rotation", "features", "logits", "codebooks", "levels", "codebook_size",
              "sh_degree", "feat_dim"):
        assert k in s, k
    assert s["opacity"].shape == (64, 1) and s["features"].shape == (64, 16, 3) and s["codebooks"].shape == (1, 64, 512)
    assert set(src) - set(s) == set(), set(src) - set(s)         # everything synth.make_scene provides is there
