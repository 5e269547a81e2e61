#!/bin/bash
# policy boundaries of the fp64 tensor-path kernel on the final code: k = 4, 5 at 301 columns, k = 6 at 81 .. 137 columns
mkdir -p gpurun_out
{
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 5 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 4 "CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 6 --m 81 "CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 6 --m 112 "CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 6 --m 136 "CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 8 --m 81 "CNMF_MMA=0;CNMF_MMA=3"
timeout 300 python tools/mma_sweep.py --n 800000 --iters 20 --k 6 --m 383 "CNMF_MMA=0;CNMF_MMA=3"
} > gpurun_out/policy.txt 2>&1
cat gpurun_out/policy.txt
