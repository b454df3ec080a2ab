#!/bin/bash
# N=2 bench under torchrun (GPU box with 2 GPUs)
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 --train-steps 60 --no-reference-style > gpurun_out/r02j_bench_n2.json 2> gpurun_out/r02j_bench_n2.err
echo rc=$?
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r02j_bench_n2.json") if l.startswith("{")][-1])
t = d["train"]
print("N=2 value", d["value"], "e2e", d["e2e"]["value"], "train iter/s/rank", t["value"], "median", t["ms_per_iter_median"], "p95", t["ms_per_iter_p95"], "seg", t["segments_gt"]["ms_per_iter_median"])
PY
