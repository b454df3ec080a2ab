#!/bin/bash
# prints the train medians of a short bench run with the row-form gradient off and on (LSV2_ROW_GRAD) (GPU box only)
for d in 0 1 0 1; do
  LSV2_ROW_GRAD=$d python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-reference-style --no-query --train-steps 60 2>/dev/null > /tmp/b_$d.json
  python - "$d" <<'PY'
import json, sys
d = json.loads([l for l in open(f"/tmp/b_{sys.argv[1]}.json") if l.startswith("{")][-1])
t = d["train"]
print("row_grad", sys.argv[1], "median", t["ms_per_iter_median"], "p95", t["ms_per_iter_p95"], "seg", t["segments_gt"]["ms_per_iter_median"])
PY
done
