#!/bin/bash
mkdir -p gpurun_out
for cfg in "CNMF_MMA=3" "CNMF_MMA=3 CNMF_MMA_TW=2"; do
  echo "=== $cfg"
  env $cfg CNMF_TRACE=1 CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so timeout 300 python tools/prof_roles.py f64 2>&1 | grep -v "role\|producer" | tail -45
done > gpurun_out/mma14_trace.txt 2>&1
cat gpurun_out/mma14_trace.txt
