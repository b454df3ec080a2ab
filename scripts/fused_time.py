"""Quick timing of the fused blend + decode kernel at the metric's size (1M Gaussians, 1920x1080, 3 x top-4 of 64).
usage: python scripts/fused_time.py [frames] [W] [H] [N] [CHW|HWC] [f32|bf16]"""
import sys, time
import torch
sys.path.insert(0, ".")
from langsplatv2_b200 import synth
from langsplatv2_b200.renderer import FeatureRenderer

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 10
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
N = int(sys.argv[4]) if len(sys.argv) > 4 else 1_000_000
scene = synth.make_scene(N, seed=0, levels=3)
layout = sys.argv[5] if len(sys.argv) > 5 else "CHW"
odt = torch.bfloat16 if (len(sys.argv) > 6 and sys.argv[6] == "bf16") else torch.float32
r = FeatureRenderer(scene, topk=4, fused=True, out_layout=layout, out_dtype=odt)
cams = [{k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.ring_camera(v, 8, W, H).items()} for v in range(4)]
r.calibrate(cams)
ev = {}
def hook(tag):
    e = torch.cuda.Event(enable_timing=True); e.record(); ev.setdefault(tag, []).append(e)
for _ in range(3):
    out = r.render(cams[0])
torch.cuda.synchronize()
r.hook = hook
t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
t0.record()
for i in range(frames):
    out = r.render(cams[i % len(cams)])
t1.record()
torch.cuda.synchronize()
dec = [a.elapsed_time(b) for a, b in zip(ev["decode_begin"], ev["decode_end"])]
print("frame ms %.3f  fused-kernel ms: mean %.3f min %.3f max %.3f" % (t0.elapsed_time(t1) / frames, sum(dec) / len(dec), min(dec), max(dec)))
print("out", tuple(out.shape), out.dtype, layout)
r.check_overflow()
