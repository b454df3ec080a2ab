"""GPU box: distCUDA2 of this repo against the reference's compiled SimpleKNN::knn (oracle/_ref) on the bench scene and
on a clustered cloud; CUDA events for ours, wall clock around the (synchronising) reference call."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from langsplatv2_b200 import synth
from langsplatv2_b200.knn import distCUDA2
from oracle import knn_ref

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
g = np.random.default_rng(1)
c = g.normal(0, 2, (64, 3))
clouds = {"uniform": synth.make_scene(P, seed=0, levels=1)["xyz"].float().numpy(),
          "clustered": (c[g.integers(0, 64, P)] + g.normal(0, 0.15, (P, 3)) * g.uniform(0.02, 1, (P, 1))).astype(np.float32)}
for name, pts in clouds.items():
    x = torch.from_numpy(pts).cuda()
    for _ in range(3):
        out = distCUDA2(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = distCUDA2(x)
    e1.record()
    torch.cuda.synchronize()
    ours = e0.elapsed_time(e1) / 10
    line = f"{name} P={P}: ours {ours:.3f} ms"
    if knn_ref.reference_lib() is not None:
        for _ in range(2):
            ref = knn_ref.reference_dist_cuda2(x)
        t0 = time.perf_counter()
        for _ in range(5):
            ref = knn_ref.reference_dist_cuda2(x)
        t = (time.perf_counter() - t0) / 5 * 1e3
        same = torch.equal(ref.view(torch.int32), out.view(torch.int32))
        line += f", reference {t:.3f} ms ({t / ours:.1f}x), bit-identical: {same}"
    print(line)
