#!/bin/bash
mkdir -p gpurun_out
export CNMF_TUNE="DBG=1"
P="python tools/first_light.py --prof f64"
timeout 300 $P > gpurun_out/plain_wonly.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 4 -c 1 -f -o gpurun_out/prof_f64_wonly $P > gpurun_out/ncu_wonly.log 2>&1
echo "rc=$?"
unset CNMF_TUNE
timeout 300 $P > gpurun_out/plain_honly.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 12 -c 1 -f -o gpurun_out/prof_f64_honly $P > gpurun_out/ncu_honly.log 2>&1
echo "rc=$?"
