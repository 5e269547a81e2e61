#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 900 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=3;CNMF_MMA=3,CNMF_TUNE=HINT=1000;CNMF_MMA=3,CNMF_TUNE=HINT=20000;CNMF_MMA=3,CNMF_TUNE=HINT=20000,GATENS=256;CNMF_MMA=3,CNMF_TUNE=HINT=200,GATENS=256;CNMF_MMA=3" > gpurun_out/mma16_sweep.txt 2>&1
cat gpurun_out/mma16_sweep.txt
timeout 900 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=3;CNMF_MMA=3,CNMF_TUNE=HINT=20000;CNMF_MMA=3,CNMF_TUNE=HINT=20000,GATENS=256;CNMF_MMA=3,CNMF_TUNE=HINT=1000,GATENS=256;CNMF_MMA=3" > gpurun_out/mma16_10M.txt 2>&1
cat gpurun_out/mma16_10M.txt
