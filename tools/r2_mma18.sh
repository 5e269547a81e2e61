#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_mma_kernel.py -x -q -m gpu > gpurun_out/mma18_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mma18_pytest.log; tail -4 gpurun_out/mma18_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for cfg in "CNMF_MMA=3" "CNMF_MMA=3 CNMF_TUNE=HG=55464546" "CNMF_MMA=3 CNMF_TUNE=HG=54464556"; do
  echo "=== $cfg"
  env $cfg CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so timeout 300 python tools/prof_roles.py f64 "$(echo $cfg | grep -o 'CNMF_TUNE=[^ ]*' | cut -d= -f2-)" 2>&1 | tail -17 | grep "role H\|producer"
done > gpurun_out/mma18_roles.txt 2>&1
cat gpurun_out/mma18_roles.txt
