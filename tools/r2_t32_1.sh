#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tests/mma_check.py --f32 --n 1000000 --iters 20 "CNMF_MMA32=0;CNMF_MMA32=1;CNMF_MMA32=1,CNMF_MMA_TW=2;CNMF_MMA32=1,CNMF_MMA_TW=1" > gpurun_out/t32_1.txt 2>&1
cat gpurun_out/t32_1.txt
