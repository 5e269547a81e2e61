#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=4;CNMF_MMA=4,CNMF_MMA_NW1=7;CNMF_MMA=4,CNMF_MMA_NW1=6;CNMF_MMA=4,CNMF_MMA_GATE=100;CNMF_MMA=4,CNMF_MMA_PF=2" > gpurun_out/mma7_sweep.txt 2>&1
cat gpurun_out/mma7_sweep.txt
timeout 600 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=4" > gpurun_out/mma7_10M.txt 2>&1
cat gpurun_out/mma7_10M.txt
for cfg in "CNMF_MMA=4"; do
  echo "=== $cfg"
  env $cfg CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so timeout 300 python tools/prof_roles.py f64 2>&1 | tail -18
done > gpurun_out/mma7_roles.txt 2>&1
cat gpurun_out/mma7_roles.txt
