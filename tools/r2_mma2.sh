#!/bin/bash
# second visit of the tensor-path kernel: row-owner warp count / stage sweeps and one full ncu capture with source
mkdir -p gpurun_out
timeout 900 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=1,CNMF_MMA_NW1=7;CNMF_MMA=1,CNMF_MMA_NW1=6;CNMF_MMA=1,CNMF_MMA_NW1=5;CNMF_MMA=1,CNMF_MMA_NW1=4;CNMF_MMA=1,CNMF_MMA_NW1=3;CNMF_MMA=1,CNMF_MMA_ST=4;CNMF_MMA=1,CNMF_MMA_ST=3;CNMF_MMA=2,CNMF_MMA_NW1=4" > gpurun_out/mma2_sweep.txt 2>&1
cat gpurun_out/mma2_sweep.txt
export CNMF_MMA=1
P="python tools/first_light.py --prof f64"
timeout 300 $P > gpurun_out/mma2_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_kernel -s 4 -c 1 -f -o gpurun_out/mma2_prof $P > gpurun_out/mma2_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/mma2_plain.log
