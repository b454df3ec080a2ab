"""Records how the reference's UNMODIFIED gaussian_renderer/__init__.py::render calls the rasterizer op.

Run in the build container (needs /root/reference; no GPU):  python tests/golden/make_render_calls.py
-> tests/golden/render_calls.json

The reference file is executed as it lies on disk.  Only its environment is stubbed: `diff_gaussian_rasterization`
is a recording module (the real one is the un-vendored CUDA submodule), `plyfile` / `simple_knn` (imported by
scene/gaussian_model.py) are empty shims, and `torch.zeros_like(..., device="cuda")` lands on the CPU because this
container has no GPU.  For every branch of render() (:87-93 quick, :95-98 include_feature, :100-103 RGB, plus the
python-side SH / covariance fallbacks :65-85 and override_color) the recorder stores the settings fields and, per
keyword argument of the op call (:108-119), None-ness, shape, dtype and requires_grad.  tests/test_dropin_gpu.py replays
exactly these calls through langsplatv2_b200 on the GPU box, where /root/reference does not exist.
"""
import importlib
import json
import os
import sys
import types

import torch

REF = os.environ.get("LSV2_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "render_calls.json")
N, H, W, LEVELS, K = 97, 40, 56, 3, 4


def tensor_spec(t):
    if t is None:
        return None
    return {"shape": list(t.shape), "dtype": str(t.dtype).replace("torch.", ""), "requires_grad": bool(t.requires_grad)}


def record_calls(ref_root=REF):
    calls = []

    class GaussianRasterizationSettings:
        def __init__(self, **kw):
            self.kw = kw
            self.__dict__.update(kw)

    class GaussianRasterizer:
        def __init__(self, raster_settings):
            self.raster_settings = raster_settings

        def __call__(self, **kw):
            rs = self.raster_settings.kw
            calls[-1]["settings"] = {k: (tensor_spec(v) if torch.is_tensor(v) else v) for k, v in rs.items()}
            calls[-1]["settings_order"] = list(rs.keys())
            calls[-1]["kwargs"] = {k: tensor_spec(v) for k, v in kw.items()}
            calls[-1]["kwargs_order"] = list(kw.keys())
            n = kw["means3D"].shape[0]
            h, w = rs["image_height"], rs["image_width"]
            return torch.zeros(3, h, w), torch.zeros(192 if rs["quick_render"] else 64, h, w), torch.ones(n, dtype=torch.int32)

    stub = types.ModuleType("diff_gaussian_rasterization")
    stub.GaussianRasterizationSettings = GaussianRasterizationSettings
    stub.GaussianRasterizer = GaussianRasterizer
    saved = {k: sys.modules.get(k) for k in ("diff_gaussian_rasterization", "plyfile", "simple_knn", "simple_knn._C")}
    sys.modules["diff_gaussian_rasterization"] = stub
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    sys.modules["plyfile"] = ply
    knn = types.ModuleType("simple_knn")
    knn_c = types.ModuleType("simple_knn._C")
    knn_c.distCUDA2 = lambda *a, **k: None
    knn._C = knn_c
    sys.modules["simple_knn"] = knn
    sys.modules["simple_knn._C"] = knn_c
    sys.path.insert(0, ref_root)
    real_zeros_like = torch.zeros_like

    def zeros_like(x, *a, **k):
        if str(k.get("device", "")).startswith("cuda") and not torch.cuda.is_available():
            k = dict(k, device=x.device)
        return real_zeros_like(x, *a, **k)

    torch.zeros_like = zeros_like
    try:
        for m in [m for m in sys.modules if m == "gaussian_renderer" or m.startswith("gaussian_renderer.")]:
            del sys.modules[m]
        gr = importlib.import_module("gaussian_renderer")
        src = open(gr.__file__, "rb").read()
        g = torch.Generator().manual_seed(0)

        class PC:                                     # duck-typed GaussianModel: only what render() touches
            active_sh_degree = 3
            max_sh_degree = 3
            get_xyz = torch.randn(N, 3, generator=g)
            get_opacity = torch.rand(N, 1, generator=g)
            get_scaling = torch.rand(N, 3, generator=g) * 0.05
            get_rotation = torch.nn.functional.normalize(torch.randn(N, 4, generator=g))
            get_features = torch.randn(N, 16, 3, generator=g)
            _language_feature_weights = torch.rand(N, LEVELS * K, generator=g)
            _language_feature_indices = torch.randint(0, 64, (N, LEVELS * K), generator=g).float()

            def get_covariance(self, scaling_modifier=1):
                return torch.rand(N, 6, generator=g)

            def get_render_weights(self, topk):
                return torch.rand(N, 64, generator=g).requires_grad_(True)

        class Cam:
            FoVx, FoVy = 1.0471975511965976, 0.7853981633974483
            image_height, image_width = H, W
            world_view_transform = torch.eye(4)
            full_proj_transform = torch.eye(4)
            camera_center = torch.zeros(3)

        class NS:
            def __init__(self, **kw):
                self.__dict__.update(kw)

        bg = torch.tensor([0.1, 0.2, 0.3])
        variants = [
            ("rgb", dict(include_feature=False, quick_render=False), dict(), {}),
            ("include_feature", dict(include_feature=True, quick_render=False), dict(), {}),
            ("quick", dict(include_feature=True, quick_render=True), dict(), {}),
            ("rgb_python_cov3d", dict(include_feature=False, quick_render=False), dict(compute_cov3D_python=True), {}),
            ("rgb_python_sh", dict(include_feature=False, quick_render=False), dict(convert_SHs_python=True), {}),
            ("rgb_override_color", dict(include_feature=False, quick_render=False), dict(),
             dict(override_color=torch.rand(N, 3, generator=g))),
        ]
        for name, optkw, pipekw, extra in variants:
            pipe = NS(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
            pipe.__dict__.update(pipekw)
            opt = NS(topk=K, **optkw)
            calls.append({"branch": name, "opt": optkw, "pipe": dict(pipe.__dict__)})
            out = gr.render(Cam(), PC(), pipe, bg, opt, **extra)
            calls[-1]["returned_keys"] = sorted(out.keys())
            calls[-1]["visibility_is_radii_gt_0"] = bool((out["visibility_filter"] == (out["radii"] > 0)).all())
        import hashlib
        return {"reference_file": "gaussian_renderer/__init__.py", "sha256": hashlib.sha256(src).hexdigest(),
                "N": N, "H": H, "W": W, "calls": calls}
    finally:
        torch.zeros_like = real_zeros_like
        sys.path.remove(ref_root)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        for m in [m for m in sys.modules if m.split(".")[0] in ("gaussian_renderer", "scene", "utils", "arguments")]:
            del sys.modules[m]


if __name__ == "__main__":
    rec = record_calls()
    json.dump(rec, open(OUT, "w"), indent=1, sort_keys=True)
    print("wrote", OUT, "branches:", [c["branch"] for c in rec["calls"]])
