#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
{
timeout 600 python tools/mma_sweep.py --f32 --n 1000000 --iters 20 --k 6 "CNMF_MMA32=0;CNMF_MMA32=1;CNMF_MMA32=1,CNMF_TUNE=NW1=6;CNMF_MMA32=1,CNMF_TUNE=NW1=5;CNMF_MMA32=1,CNMF_TUNE=NW1=4;CNMF_MMA32=1,CNMF_TUNE=ST=3;CNMF_MMA32=1,CNMF_TUNE=FL=8;CNMF_MMA32=1,CNMF_TUNE=FL=256"
timeout 600 python tools/mma_sweep.py --f32 --n 1700000 --iters 20 --k 8 --m 184 "CNMF_MMA32=0;CNMF_MMA32=1;CNMF_MMA32=1,CNMF_TUNE=NW1=5;CNMF_MMA32=1,CNMF_TUNE=ST=4;CNMF_MMA32=1,CNMF_MMA_TW=4;CNMF_MMA32=1,CNMF_MMA_TW=1"
} > gpurun_out/t32_3.txt 2>&1
cat gpurun_out/t32_3.txt
