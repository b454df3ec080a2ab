"""Kernel timeline of one captured training iteration at N ranks (torch.profiler / CUPTI): where the exchange sits
relative to the compute.   torchrun --nproc-per-node 2 scripts/ddp_trace.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import profile, ProfilerActivity
from langsplatv2_b200 import synth
from langsplatv2_b200.trainer import FeatureTrainer, synthetic_gt

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N, W, H = 1_000_000, 1008, 756
scene = synth.make_scene(N, seed=0, levels=1)
cam = synth.ring_camera(rank, 64, W, H)
cam = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()}
gt, mask = synthetic_gt(H, W, seed=2 + rank)
tr = FeatureTrainer(scene, topk=4, layers=1, use_graph=True)
tr.calibrate([cam])
for _ in range(8):
    tr.step(cam, gt, mask, view_id=0)
torch.cuda.synchronize(); dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(4):
        tr.step(cam, gt, mask, view_id=0)
    torch.cuda.synchronize()
if rank == 0:
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: e.time_range.start)
    # split into iterations at k_preprocess (the first kernel after the all-gather launch of an iteration)
    starts = [i for i, e in enumerate(ev) if "k_depth_keys" in e.name]
    if len(starts) >= 3:
        a, b = starts[1], starts[2]
        # back up to include whatever precedes k_depth_keys in the iteration (all-gathers, k_preprocess)
        while a > 0 and ev[a - 1].time_range.start > ev[starts[1]].time_range.start - 200:
            a -= 1
        t0 = ev[a].time_range.start
        print(f"iteration: {ev[b].time_range.start - ev[starts[1]].time_range.start:.0f} us between the k_depth_keys of two iterations")
        print(f"{'start us':>9s} {'dur us':>8s} {'stream':>6s}  kernel")
        for e in ev[a:b]:
            nm = e.name.replace("lsv2::", "").replace("void ", "")[:70]
            print(f"{e.time_range.start - t0:9.0f} {e.time_range.end - e.time_range.start:8.0f} {getattr(e, 'device_resource_id', -1):>6}  {nm}")
dist.destroy_process_group()
