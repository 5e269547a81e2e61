#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_NW1=4" > gpurun_out/mma4_10M.txt 2>&1
cat gpurun_out/mma4_10M.txt
export CNMF_MMA=3
P="python tools/first_light.py --prof f64"
timeout 300 $P > gpurun_out/mma4_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_kernel -s 4 -c 1 -f -o gpurun_out/mma4_prof $P > gpurun_out/mma4_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/mma4_plain.log
