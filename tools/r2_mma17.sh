#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 600 python tests/mma_check.py --n 1000000 --iters 20 "CNMF_MMA=3" > gpurun_out/mma17_check.txt 2>&1
grep -E "BAD|===|timing|Error|error|TIMEOUT" gpurun_out/mma17_check.txt
timeout 900 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=3,CNMF_TUNE=HGEVEN=1;CNMF_MMA=3;CNMF_MMA=3,CNMF_TUNE=HG=44564456;CNMF_MMA=3,CNMF_TUNE=HG=44554466;CNMF_MMA=3,CNMF_TUNE=HG=55454554;CNMF_MMA=3,CNMF_TUNE=HG=45464546;CNMF_MMA=3,CNMF_TUNE=HGEVEN=1" > gpurun_out/mma17_sweep.txt 2>&1
cat gpurun_out/mma17_sweep.txt
timeout 900 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=3,CNMF_TUNE=HGEVEN=1;CNMF_MMA=3;CNMF_MMA=3,CNMF_TUNE=HGEVEN=1;CNMF_MMA=3" > gpurun_out/mma17_10M.txt 2>&1
cat gpurun_out/mma17_10M.txt
