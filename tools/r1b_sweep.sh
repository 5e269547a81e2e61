#!/bin/bash
# sweep passed as $1 (first_light --sweep syntax), optional extra args in $2
mkdir -p gpurun_out
timeout 900 python tools/sweep.py $2 "$1" > gpurun_out/sweep_last.log 2>&1
cat gpurun_out/sweep_last.log
