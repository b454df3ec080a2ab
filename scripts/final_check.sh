#!/bin/bash
# full GPU verification of the tree: gpu tests, smoke, default bench line (+ a launch list of the distCUDA2 kernels)
python -c "import torch; torch.zeros(1).cuda()"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r02w_bench_n1.json 2> gpurun_out/r02w_bench_n1.err; echo bench rc=$?
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r02w_bench_n1.json") if l.startswith("{")][-1])
t = d["train"]
print("FPS", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 3), "train median", round(t["ms_per_iter_median"], 3),
      "seg", round(t["segments_gt"]["ms_per_iter_median"], 3), "scene_init", t.get("scene_init"), "clocks", d["clocks"])
PY
timeout 120 python scripts/knn_time.py > /dev/null 2>&1 && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02w_knn_launches.csv python scripts/knn_time.py 200000 > /dev/null 2>&1; echo ncu rc=$?
