#!/bin/bash
mkdir -p gpurun_out
export CNMF_MMA32=1
P="python tools/first_light.py --prof f32"
timeout 300 $P > gpurun_out/t32_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma32_kernel -s 4 -c 1 -f -o gpurun_out/t32_prof $P > gpurun_out/t32_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/t32_plain.log
