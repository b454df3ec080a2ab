"""N-rank check of the sliced gradient all-reduce (FeatureTrainer.overlap_chunks) against the single all-reduce.
   torchrun --nproc-per-node 2 scripts/ddp_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from langsplatv2_b200 import synth
from langsplatv2_b200.trainer import FeatureTrainer, synthetic_gt

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N, W, H = 1_000_000, 1008, 756
scene = synth.make_scene(N, seed=0, levels=3)
cam = synth.ring_camera(rank, 64, W, H)
cam = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in cam.items()}
gt, mask = synthetic_gt(H, W, seed=2 + rank)
res = {}
for chunks in (int(os.environ.get("LSV2_OVERLAP_CHUNKS", "4")), 1):
    os.environ["LSV2_OVERLAP_CHUNKS"] = str(max(chunks, 2))
    tr = FeatureTrainer(scene, topk=4, layers=1)
    tr.overlap_chunks = chunks
    tr.keep_last_grad = True
    tr.step(cam, gt, mask)
    res[chunks] = tr.last_grad
    tr.keep_last_grad = False
    for _ in range(5):
        tr.step(cam, gt, mask)
    torch.cuda.synchronize(); dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(20):
        tr.step(cam, gt, mask)
    t1.record(); torch.cuda.synchronize()
    if rank == 0:
        print("chunks", chunks, "ms/iter %.3f" % (t0.elapsed_time(t1) / 20), flush=True)
    del tr
a, b = [res[k] for k in res if k != 1][0], res[1]
err = float((a - b).abs().max()); ref = float(b.abs().max())
same = torch.tensor([err <= 1e-4 * ref + 1e-12], device="cuda", dtype=torch.int32)
dist.all_reduce(same, op=dist.ReduceOp.MIN)
if rank == 0:
    print("max |sliced - single| = %.3e of max |grad| %.3e ->" % (err, ref), "OK" if int(same) else "MISMATCH", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(same) else 3)
