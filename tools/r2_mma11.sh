#!/bin/bash
mkdir -p gpurun_out
for cfg in "CNMF_MMA=3" "CNMF_MMA=3 CNMF_TUNE=PF=2" "CNMF_MMA=1 CNMF_TUNE=NW1=4"; do
  echo "=== $cfg"
  env $cfg CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so timeout 300 python tools/prof_roles.py f64 "$(echo $cfg | grep -o 'CNMF_TUNE=[^ ]*' | cut -d= -f2-)" 2>&1 | tail -17
done > gpurun_out/mma11_roles.txt 2>&1
cat gpurun_out/mma11_roles.txt
