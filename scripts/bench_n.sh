#!/bin/bash
# bench under torchrun on N GPUs of the box: scripts/bench_n.sh N TAG
N=$1; TAG=$2
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --train-steps 60 --no-reference-style > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
echo rc=$?
python - "$N" "$TAG" <<'PY'
import json, sys
n, tag = sys.argv[1], sys.argv[2]
d = json.loads([l for l in open(f"gpurun_out/{tag}_bench_n{n}.json") if l.startswith("{")][-1])
t = d["train"]
print(f"N={n} value", d["value"], "e2e", d["e2e"]["value"], "ms/frame", d["ms_per_step"], "train iter/s/rank", t["value"], "median", t["ms_per_iter_median"], "p95", t["ms_per_iter_p95"], "seg", t["segments_gt"]["ms_per_iter_median"])
PY
