#!/bin/bash
# several sweeps in one call: each argument is "n m k|sweep"
mkdir -p gpurun_out
: > gpurun_out/multi_last.log
for spec in "$@"; do
  shape="${spec%%|*}"; sw="${spec#*|}"
  set -- $shape
  timeout 600 python tools/sweep.py --n $1 --m $2 --k $3 "$sw" 2>&1 | grep -E "^timing|^FAILED|^STALLED" >> gpurun_out/multi_last.log
done
cat gpurun_out/multi_last.log
