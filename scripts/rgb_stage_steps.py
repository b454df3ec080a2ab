"""A few RGB-stage iterations (forward in RGB mode + L1 loss + the geometry backward) for timing / ncu launch lists.
GPU box only.  Shape of train.py without --include_feature (train.py:169-172), minus ssim and the optimizer."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from langsplatv2_b200 import synth, GaussianRasterizer
from tests import helpers as hp

N = int(os.environ.get("N", 1_000_000)); W = int(os.environ.get("W", 1008)); H = int(os.environ.get("H", 756))
steps = int(os.environ.get("STEPS", 6))
scene = synth.make_scene(N, seed=0, levels=1)
cam = synth.ring_camera(0, 64, W, H)
rs = hp.settings(cam, scene, (0.0, 0.0, 0.0))
leaf = lambda t: t.detach().float().cuda().clone().requires_grad_(True)
p = dict(means3D=leaf(scene["xyz"]), opacities=leaf(scene["opacity"]), shs=leaf(scene["features"]),
         scales=leaf(scene["scaling"]), rotations=leaf(scene["rotation"]))
gt = torch.rand(3, H, W, device="cuda")
ph = torch.zeros((1,), device="cuda")
cap = None


def step():
    global cap
    for v in p.values():
        v.grad = None
    means2D = torch.zeros(N, 3, device="cuda", requires_grad=True)
    rast = GaussianRasterizer(raster_settings=rs, capacity=cap)
    color, _, radii = rast(means2D=means2D, colors_precomp=None, cov3D_precomp=None, language_feature_precomp=ph,
                           language_feature_weights_quick=ph, language_feature_indices=ph, **p)
    loss = (color - gt).abs().mean()
    loss.backward()
    if cap is None:
        cap = int(rast.last_state.num_rendered * 1.2) + 4096
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
evs[0].record()
for i in range(steps):
    loss = step()
    evs[i + 1].record()
torch.cuda.synchronize()
print("per-step ms", [round(evs[i].elapsed_time(evs[i + 1]), 3) for i in range(steps)], "loss", float(loss))
