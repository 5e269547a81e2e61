#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tests/mma_check.py --f32 --n 1000000 --iters 20 "CNMF_MMA32=1" 2>&1 | grep -E "BAD|===|timing|rror|TIMEOUT" > gpurun_out/t32_4.txt
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
{
timeout 600 python tools/mma_sweep.py --f32 --n 1000000 --iters 20 --k 6 "CNMF_MMA32=0;CNMF_MMA32=1;CNMF_MMA32=1,CNMF_TUNE=FL=8;CNMF_MMA32=1,CNMF_TUNE=FL=256;CNMF_MMA32=1,CNMF_MMA_TW=2"
timeout 600 python tools/mma_sweep.py --f32 --n 1700000 --iters 20 --k 8 --m 184 "CNMF_MMA32=0;CNMF_MMA32=1;CNMF_MMA32=1,CNMF_MMA_TW=4"
timeout 600 python tools/mma_sweep.py --f32 --n 10000000 --iters 30 --k 6 "CNMF_MMA32=0;CNMF_MMA32=1"
} >> gpurun_out/t32_4.txt 2>&1
cat gpurun_out/t32_4.txt
