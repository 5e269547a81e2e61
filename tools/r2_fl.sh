#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
{
timeout 600 python tools/mma_sweep.py --f32 --n 1000000 --iters 20 --k 6 "CNMF_TUNE=FL=0;CNMF_TUNE=FL=8;CNMF_TUNE=FL=128;CNMF_TUNE=FL=1024"
timeout 600 python tools/mma_sweep.py --f32 --n 1700000 --iters 20 --k 8 --m 184 "CNMF_TUNE=FL=0;CNMF_TUNE=FL=8;CNMF_TUNE=FL=128;CNMF_TUNE=FL=1024"
} > gpurun_out/fl_k1.txt 2>&1
cat gpurun_out/fl_k1.txt
