#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tests/mma_check.py --n 1000000 --iters 20 "CNMF_MMA=3;CNMF_MMA=1" > gpurun_out/mma5_check.txt 2>&1
grep -E "BAD|===|timing|Error|error" gpurun_out/mma5_check.txt
timeout 1200 python tools/mma_sweep.py --n 1000000 --iters 20 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_PF=1;CNMF_MMA=3,CNMF_MMA_PF=2;CNMF_MMA=3,CNMF_MMA_PF=4;CNMF_MMA=3,CNMF_MMA_NW1=6;CNMF_MMA=3,CNMF_MMA_NW1=4;CNMF_MMA=3,CNMF_MMA_GATE=100;CNMF_MMA=1,CNMF_MMA_NW1=4" > gpurun_out/mma5_sweep.txt 2>&1
cat gpurun_out/mma5_sweep.txt
timeout 600 python tools/mma_sweep.py --n 10000000 --iters 40 "CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_PF=2" > gpurun_out/mma5_10M.txt 2>&1
cat gpurun_out/mma5_10M.txt
