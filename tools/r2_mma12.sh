#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tests/mma_check.py --n 1000000 --iters 20 "CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=1,CNMF_MMA_TW=4;CNMF_MMA=3" > gpurun_out/mma12_check.txt 2>&1
grep -E "BAD|===|timing|Error|error|TIMEOUT" gpurun_out/mma12_check.txt
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 900 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=3,CNMF_MMA_TW=4,CNMF_TUNE=NW1=6;CNMF_MMA=3,CNMF_MMA_TW=4,CNMF_TUNE=NW1=5;CNMF_MMA=1,CNMF_MMA_TW=4;CNMF_MMA=1,CNMF_MMA_TW=4,CNMF_TUNE=NW1=4;CNMF_MMA=1,CNMF_MMA_TW=4,CNMF_TUNE=NW1=5;CNMF_MMA=1,CNMF_TUNE=NW1=4" > gpurun_out/mma12_sweep.txt 2>&1
cat gpurun_out/mma12_sweep.txt
timeout 600 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4" > gpurun_out/mma12_10M.txt 2>&1
cat gpurun_out/mma12_10M.txt
