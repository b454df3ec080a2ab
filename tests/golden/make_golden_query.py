#!/usr/bin/env python
"""Generates tests/golden/reference_query.npz by RUNNING the reference's own relevancy function.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_query.py

eval/openclip_encoder.py is imported unmodified with `open_clip` / `torchvision` stubbed (the file only needs them to
load CLIP weights in __init__, which is not called); `OpenCLIPNetwork.get_max_across_quick` (:114-138, the function
eval_lerf.py:351-352 and backend_renderer.py:204-210 feed with the decoded map) is then evaluated, unbound, on a
seeded unit-norm feature map and seeded unit phrase embeddings.
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
sys.dont_write_bytecode = True
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_query.npz")


def main():
    for name in ("open_clip", "torchvision"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    sys.path.insert(0, os.path.join(REF, "eval"))
    import openclip_encoder
    fn = openclip_encoder.OpenCLIPNetwork.get_max_across_quick
    g = torch.Generator().manual_seed(77)
    out = {}
    for tag, (levels, H, W, P, N) in {"a": (3, 6, 7, 3, 4), "b": (1, 5, 4, 1, 4), "c": (3, 4, 5, 9, 2)}.items():
        sem = torch.nn.functional.normalize(torch.randn(levels, H, W, 512, generator=g), dim=-1)
        sem[0, 0, 0] = 0.0                                             # an empty pixel (feature 0 / (0 + 1e-10))
        pos = torch.nn.functional.normalize(torch.randn(P, 512, generator=g), dim=-1)
        neg = torch.nn.functional.normalize(torch.randn(N, 512, generator=g), dim=-1)
        # make some similarities large so the softmax saturates in places
        pos[0] = torch.nn.functional.normalize(sem[0, H // 2, W // 2] + 0.2 * pos[0], dim=-1)
        me = types.SimpleNamespace(positives=tuple("p%d" % i for i in range(P)), negatives=tuple("n%d" % i for i in range(N)),
                                   pos_embeds=pos, neg_embeds=neg)
        relev = fn(me, sem)
        assert relev.shape == (levels, P, H, W)
        out[f"{tag}_sem"] = sem.numpy()
        out[f"{tag}_pos"] = pos.numpy()
        out[f"{tag}_neg"] = neg.numpy()
        out[f"{tag}_relev"] = relev.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items() if k.endswith("relev")})


if __name__ == "__main__":
    main()
