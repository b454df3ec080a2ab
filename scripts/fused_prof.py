"""Where the fused kernel's warp roles spend their cycles (instrumented library: `make -C langsplatv2_b200/csrc prof`).
usage: LSV2_LIB=langsplatv2_b200/liblsv2_prof.so python scripts/fused_prof.py [W H N]"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, ".")
assert os.environ.get("LSV2_LIB"), "run with LSV2_LIB=.../liblsv2_prof.so (a -DFD_PROF build)"
from langsplatv2_b200 import synth
from langsplatv2_b200.renderer import FeatureRenderer

W = int(sys.argv[1]) if len(sys.argv) > 1 else 1920
H = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
N = int(sys.argv[3]) if len(sys.argv) > 3 else 1_000_000
scene = synth.make_scene(N, seed=0, levels=3)
r = FeatureRenderer(scene, topk=4, fused=True)
cam = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.ring_camera(0, 8, W, H).items()}
r.calibrate([cam])
for _ in range(3):
    out = r.render(cam)
torch.cuda.synchronize()
scratch = r.last_state.keep[-1]
levels = 3
off = 256 + levels * 512 * 128 + levels * 128 * 128
prof = scratch[off:off + 148 * 6 * 8 * 8].cpu().numpy().view(np.int64).reshape(148, 6, 8).astype(np.float64)
names = {0: ["wait full (producer)", "blend batch compute", "wait w_full (last MMA)", "wait dec_done (prev decode)",
             "pass A (TMEM -> fp16 rows)", "colour out + Gram norm (u_full waits)", "unit setup"],
         1: ["wait stage_free", "scatter (batches 1..)", "unit setup / loop", "stop check + look-ahead links",
             "zero S + issue gathers", "scatter (batch 0 of a unit: first use of its gather)", "masks + fence + arrive + rotate"],
         2: ["wait wraw (blend side)", "wait codebook chunk", "wait F slot free (epilogue)", "issue"],
         3: ["wait inv_full (unit hand-off)", "wait F slot full (MMA)", "tcgen05.ld + stores"],
         4: ["header / mask read", "stage 1: loads + quadratic forms", "exp x4 + alpha", "transmittance chain",
             "fp16 hi/lo split", "A-tile stores", "fence + arrive + votes"]}
roles = ["blend warp 0", "producer", "decode MMA issuer", "epilogue warp 8", "blend warp 0, inside a batch"]
mean = prof.mean(0)
for role in range(5):
    tot = mean[role].sum()
    print(f"{roles[role]}: {tot / 1.965e6:.3f} ms accounted per CTA (mean over 148)")
    for i, nm in enumerate(names[role]):
        print(f"    {nm:45s} {mean[role][i] / 1.965e6:8.3f} ms  {100 * mean[role][i] / max(tot, 1):5.1f}%")
