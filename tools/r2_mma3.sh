#!/bin/bash
mkdir -p gpurun_out
timeout 120 tools/micro/_bin/r2_dmma > gpurun_out/micro_r2_dmma.txt 2>&1; cat gpurun_out/micro_r2_dmma.txt
timeout 1200 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=0;CNMF_MMA=1;CNMF_MMA=1,CNMF_MMA_NW1=4;CNMF_MMA=1,CNMF_MMA_GATE=200;CNMF_MMA=1,CNMF_MMA_NW1=4,CNMF_MMA_GATE=200;CNMF_MMA=2;CNMF_MMA=2,CNMF_MMA_NW1=4;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_NW1=4;CNMF_MMA=3,CNMF_MMA_NW1=5;CNMF_MMA=3,CNMF_MMA_NW1=3;CNMF_MMA=3,CNMF_MMA_TW=4,CNMF_MMA_NW1=4" > gpurun_out/mma3_sweep.txt 2>&1
cat gpurun_out/mma3_sweep.txt
