#!/bin/bash
mkdir -p gpurun_out
for cfg in "CNMF_MMA=3" "CNMF_MMA=1 CNMF_MMA_NW1=4" "CNMF_MMA=3 CNMF_MMA_TW=4"; do
  echo "=== $cfg"
  env $cfg CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so timeout 300 python tools/prof_roles.py f64 2>&1 | tail -40
done > gpurun_out/mma6_roles.txt 2>&1
cat gpurun_out/mma6_roles.txt
timeout 600 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=3,CNMF_MMA_TW=1;CNMF_MMA=1,CNMF_MMA_TW=4;CNMF_MMA=0" > gpurun_out/mma6_sweep.txt 2>&1
cat gpurun_out/mma6_sweep.txt
