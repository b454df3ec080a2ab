"""Debug aid: where does the fused decode differ from the checker?  (GPU box only)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oracle
from langsplatv2_b200 import render_decode
from tests import helpers as hp

BG = (0.1, 0.2, 0.3)
N, W, H, seed = 4000, 144, 100, 5
norm = int(os.environ.get("NORM", "1")) != 0
scene, cam = hp.make_case(N, W, H, seed=seed, levels=3, scale_boost=2.0)
_, wq, iq = hp.coeffs(scene, k=4)
pre, binned, fwd = oracle.forward(hp.oracle_scene(scene), hp.oracle_view(cam), mode=2, L=192, weights_quick=wq, indices_quick=iq, bg=BG)
st = hp.run_cuda(scene, cam, BG, hp.MODE_QUICK, wq=wq, iq=iq, render=False)
out = render_decode(st, scene["codebooks"].cuda(), normalize=norm, want_wmap=True)
torch.cuda.synchronize()
cb = scene["codebooks"].numpy()
ref = oracle.decode(fwd["wmap"].reshape(192, -1), cb, normalize=norm).reshape(3, 512, H, W)
got = out.float().cpu().numpy()
err = np.abs(got - ref)
print("max err", err.max(), "ref max", np.abs(ref).max())
print("per level", err.reshape(3, -1).max(1))
print("per 32-chunk (level0)", err[0].reshape(16, 32, -1).max((1, 2)).round(4))
print("per 32-chunk (level2)", err[2].reshape(16, 32, -1).max((1, 2)).round(4))
e = err.max((0, 1))                      # [H, W]
ys, xs = np.nonzero(e > 1e-3)
print("bad pixels", len(ys), "of", H * W)
if len(ys):
    print("rows-in-tile hist", np.bincount(ys % 16, minlength=16))
    print("cols-in-tile hist", np.bincount(xs % 16, minlength=16))
    print("tile hist", np.unique((ys // 16) * ((W + 15) // 16) + xs // 16, return_counts=True))
    y, x = ys[0], xs[0]
    print("example px", y, x, "got", got[:, :4, y, x], "ref", ref[:, :4, y, x])
    ratio = (got * ref).sum(1) / np.maximum((ref * ref).sum(1), 1e-20)
    print("got.ref/ref.ref at bad px", ratio[:, ys[:8], xs[:8]])

gx = (W + 15) // 16; gy = (H + 15) // 16
tb = np.zeros((gy, gx), int)
for y, x in zip(ys, xs): tb[y // 16, x // 16] += 1
print("bad px per tile:\n", tb)
z = (got == 0).all((0, 1)) & (np.abs(ref).sum((0, 1)) > 0)
print("pixels exactly zero but live:", int(z.sum()))
