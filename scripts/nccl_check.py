"""all-reduce bandwidth check under torchrun (debugging aid for the DP training step)."""
import os, torch, torch.distributed as dist
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
x = torch.ones(64 * 1024 * 1024, device="cuda")
for _ in range(3): dist.all_reduce(x)
torch.cuda.synchronize()
b, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
b.record()
for _ in range(10): dist.all_reduce(x)
e.record(); torch.cuda.synchronize()
ms = b.elapsed_time(e) / 10
if dist.get_rank() == 0:
    print(f"allreduce 256 MB: {ms:.3f} ms  algbw {0.268 / ms * 1e3:.1f} GB/s  can_access_peer {torch.cuda.can_device_access_peer(0, 1)}")
dist.destroy_process_group()
