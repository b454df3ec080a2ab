"""Runs a few feature-stage training iterations (for ncu launch lists / timing breakdowns).  GPU box only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from langsplatv2_b200 import synth
from langsplatv2_b200.trainer import FeatureTrainer, synthetic_gt, synthetic_gt_segments

N = int(os.environ.get("N", 1_000_000)); W = int(os.environ.get("W", 1008)); H = int(os.environ.get("H", 756))
steps = int(os.environ.get("STEPS", 4))
scene = synth.make_scene(N, seed=0, levels=3)
cam = synth.ring_camera(0, 64, W, H)
cam = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()}
gt, mask = synthetic_gt(H, W)
if os.environ.get("SEGMENTS") == "1":          # ground truth as (seg, table): loss.decode_cos_loss_segments
    gt, mask = synthetic_gt_segments(H, W), None
tr = FeatureTrainer(scene, topk=4, layers=1)
for _ in range(2):
    tr.step(cam, gt, mask)
torch.cuda.synchronize()
t0 = time.perf_counter()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
evs[0].record()
for i in range(steps):
    loss = tr.step(cam, gt, mask)
    evs[i + 1].record()
torch.cuda.synchronize()
print("per-step ms", [round(evs[i].elapsed_time(evs[i + 1]), 2) for i in range(steps)], "capacity", tr.capacity)
print("ms/iter", (time.perf_counter() - t0) * 1e3 / steps, "loss", float(loss))
