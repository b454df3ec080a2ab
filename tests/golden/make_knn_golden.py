"""Generates tests/golden/knn_golden.npz: outputs of the REFERENCE's own distCUDA2 core (SimpleKNN::knn, compiled from
/root/reference/submodules/simple-knn/simple_knn.cu into oracle/_ref/libsimple_knn_ref.so by `make -C oracle knn_ref`).
CUDA code, so this runs on a GPU box:  python tests/golden/make_knn_golden.py gpurun_out/knn_golden.npz
(the file is then copied to tests/golden/).  Inputs are stored with the outputs; nothing here reads /root/reference."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from oracle import knn_ref  # noqa: E402


def cases():
    g = np.random.default_rng(7)
    out = {}
    out["uniform_3000"] = g.uniform(-2, 2, (3000, 3)).astype(np.float32)
    c = g.normal(0, 1, (12, 3))
    out["clusters_5000"] = (c[g.integers(0, 12, 5000)] + g.normal(0, 0.05, (5000, 3))).astype(np.float32)
    d = g.uniform(-1, 1, (1500, 3)).astype(np.float32)
    d[100:200] = d[0:100]                       # exact duplicates: distance 0 counts (only j == i is skipped)
    d[300:310] = d[300]                         # ten coincident points
    out["duplicates_1500"] = d
    out["plane_2049"] = np.concatenate([g.uniform(-3, 3, (2049, 2)), np.zeros((2049, 1))], 1).astype(np.float32)
    out["far_offset_1025"] = (g.normal(0, 1e-3, (1025, 3)) + np.array([100.0, -50.0, 25.0])).astype(np.float32)
    for n in (1, 2, 3, 4, 5):
        out[f"tiny_{n}"] = g.uniform(-1, 1, (n, 3)).astype(np.float32)
    return out


def main(path):
    res = {}
    for name, pts in cases().items():
        ref = knn_ref.reference_dist_cuda2(torch.from_numpy(pts).cuda()).cpu().numpy()
        res[name + "/points"] = pts
        res[name + "/dist2"] = ref
        cpu = knn_ref.dist_cuda2(pts)
        with np.errstate(invalid="ignore"):
            same = (cpu == ref) | (np.isinf(cpu) & np.isinf(ref))
        print(f"{name}: P={pts.shape[0]} restatement bit-equal on {int(same.sum())}/{pts.shape[0]}")
    np.savez_compressed(path, **res)
    print("wrote", path)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/knn_golden.npz")
