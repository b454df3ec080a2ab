#!/bin/bash
# usage: scripts/gpuretry.sh <outfile> <timeout> <command...>   -- retries gpurun while the pod answers "busy"
out=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $out 2>&1
  if ! grep -q "status=transient" $out; then break; fi
  sleep 120
done
echo DONE >> $out
