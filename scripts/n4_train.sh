#!/bin/bash
# N-GPU box: train-only bench with the default (row-form) exchange: scripts/n4_train.sh N
N=${1:-4}
python -c "import torch; torch.zeros(1).cuda()"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 --train-steps 60 --no-reference-style --no-query --no-cpu-baseline > gpurun_out/r02v_bench_n${N}_train.json 2> /tmp/n4.err
python - "$N" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/r02v_bench_n{n}_train.json") if l.startswith("{")][-1])
    t = d["train"]
    print(f"N={n} FPS", round(d["value"], 1), "train median", round(t["ms_per_iter_median"], 3), "p95", round(t["ms_per_iter_p95"], 3), "iter/s/rank", round(t["value"], 1), "seg", round(t["segments_gt"]["ms_per_iter_median"], 3), "exchange bytes", t["config"]["grad_exchange_bytes"])
except Exception as e:
    print("FAILED", e); print(open("/tmp/n4.err").read()[-2000:])
PY
