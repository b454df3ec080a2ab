#!/bin/bash
# A/B timing of fused-kernel variants ON THE SAME BOX (boxes differ by +-3 % under the power cap):
# scripts/ab_fused.sh name1 name2 ...   with libraries langsplatv2_b200/ab/liblsv2_<name>.so; two interleaved rounds
for round in 1 2; do
  for v in "$@"; do
    echo -n "$v round $round: "
    LSV2_LIB=langsplatv2_b200/ab/liblsv2_$v.so timeout 120 python scripts/fused_time.py 40 2>&1 | grep "frame ms"
  done
done
