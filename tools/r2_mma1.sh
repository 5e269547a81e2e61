#!/bin/bash
# first GPU visit of the fp64 tensor-path kernel: parity on ragged shapes + timing at 1M rows, per variant
mkdir -p gpurun_out
timeout 1500 python tests/mma_check.py --n 1000000 --iters 20 "CNMF_MMA=0;CNMF_MMA=1;CNMF_MMA=1,CNMF_MMA_TW=1;CNMF_MMA=1,CNMF_MMA_TW=4;CNMF_MMA=2" > gpurun_out/mma1.txt 2>&1
cat gpurun_out/mma1.txt
