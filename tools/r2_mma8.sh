#!/bin/bash
mkdir -p gpurun_out
{
for k in 5 6 7 8; do
  timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k $k "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=1"
done
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 8 --m 184 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=1"
timeout 300 python tools/mma_sweep.py --n 2000000 --iters 20 --k 6 --m 81 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=1"
timeout 300 python tools/mma_sweep.py --n 500000 --iters 20 --k 6 --m 640 "CNMF_MMA=0;CNMF_MMA=3;CNMF_MMA=1"
timeout 300 python tools/mma_sweep.py --n 100000 --iters 50 --k 6 --m 301 "CNMF_MMA=0;CNMF_MMA=3"
} > gpurun_out/mma8_shapes.txt 2>&1
cat gpurun_out/mma8_shapes.txt
