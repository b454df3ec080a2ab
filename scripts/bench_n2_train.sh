#!/bin/bash
# N=2 train-only bench under different NCCL settings (same box): scripts/bench_n2_train.sh
run() {
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --train-steps 60 --no-reference-style --no-query --no-cpu-baseline > /tmp/n2_$tag.json 2> /tmp/n2_$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads([l for l in open(f"/tmp/n2_{tag}.json") if l.startswith("{")][-1])
    t = d["train"]
    print(tag, "train median", round(t["ms_per_iter_median"], 3), "p95", round(t["ms_per_iter_p95"], 3), "seg", round(t["segments_gt"]["ms_per_iter_median"], 3))
except Exception as e:
    print(tag, "FAILED", e)
PY
}
run default X=1
run minch32 NCCL_MIN_NCHANNELS=32
run minch16 NCCL_MIN_NCHANNELS=16
run ctas NCCL_MIN_CTAS=32
run default2 X=1
