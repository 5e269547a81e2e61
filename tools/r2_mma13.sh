#!/bin/bash
mkdir -p gpurun_out
{
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 8 "CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4"
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 7 "CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4"
timeout 300 python tools/mma_sweep.py --n 1700000 --iters 20 --k 8 --m 184 "CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 8 --m 81 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=3,CNMF_MMA_TW=1;CNMF_MMA=1"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 6 --m 150 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=1"
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 6 --m 200 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4"
timeout 300 python tools/mma_sweep.py --n 10000000 --iters 40 --k 6 "CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=3;CNMF_MMA=3,CNMF_MMA_TW=4;CNMF_MMA=3"
} > gpurun_out/mma13_shapes.txt 2>&1
cat gpurun_out/mma13_shapes.txt
