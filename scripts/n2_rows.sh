#!/bin/bash
# 2-GPU box: parameter equivalence of the default exchange (row form) against all-reduce, then the train-only bench with
# the row exchange off and on (same box, alternating)
python -c "import torch; torch.zeros(1).cuda()"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 scripts/ddp_equiv.py 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -12 | tee gpurun_out/r02v_ddp_equiv_n2.txt
run() {
  tag=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --train-steps 60 --no-reference-style --no-query --no-cpu-baseline > /tmp/n2_$tag.json 2> /tmp/n2_$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads([l for l in open(f"/tmp/n2_{tag}.json") if l.startswith("{")][-1])
    t = d["train"]
    print(tag, "train median", round(t["ms_per_iter_median"], 3), "p95", round(t["ms_per_iter_p95"], 3), "seg", round(t["segments_gt"]["ms_per_iter_median"], 3), "exchange bytes", t["config"]["grad_exchange_bytes"])
except Exception as e:
    print(tag, "FAILED", e); print(open(f"/tmp/n2_{tag}.err").read()[-1500:])
PY
}
run rows0 LSV2_ROW_GRAD=0
run rows1 LSV2_ROW_GRAD=1
run rows0b LSV2_ROW_GRAD=0
run rows1b LSV2_ROW_GRAD=1
cp /tmp/n2_rows1b.json gpurun_out/r02v_bench_n2_train.json
