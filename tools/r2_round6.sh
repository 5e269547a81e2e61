#!/bin/bash
set -x
mkdir -p gpurun_out
P="python tools/first_light.py --prof f32"
timeout 600 $P > gpurun_out/plain_prof_f32.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 4 -c 1 -f -o gpurun_out/prof_f32_v2 $P > gpurun_out/ncu_prof_f32_v2.log 2>&1
echo "ncu rc=$?"
