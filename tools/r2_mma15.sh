#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 900 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=3;CNMF_MMA=4;CNMF_MMA=4,CNMF_MMA_TW=2;CNMF_MMA=4,CNMF_TUNE=NW1=7;CNMF_MMA=3;CNMF_MMA=4" > gpurun_out/mma15_sweep.txt 2>&1
cat gpurun_out/mma15_sweep.txt
timeout 600 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=3;CNMF_MMA=4;CNMF_MMA=3;CNMF_MMA=4" > gpurun_out/mma15_10M.txt 2>&1
cat gpurun_out/mma15_10M.txt
