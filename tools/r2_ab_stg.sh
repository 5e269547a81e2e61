#!/bin/bash
# same-box A/B: product library (W' stored before the hand-over) against the tuning library built from the new source
mkdir -p gpurun_out
T=$PWD/cnmf_b200/libcnmf_b200_tune.so
{
timeout 600 python tools/mma_sweep.py --f32 --n 1000000 --iters 20 --k 6 "A=old;CNMF_LIB=$T;A=old;CNMF_LIB=$T"
timeout 600 python tools/mma_sweep.py --f32 --n 1700000 --iters 20 --k 8 --m 184 "A=old;CNMF_LIB=$T;A=old;CNMF_LIB=$T"
timeout 600 python tools/mma_sweep.py --n 100000 --iters 100 --k 4 "A=old;CNMF_LIB=$T;A=old;CNMF_LIB=$T"
timeout 600 python tools/mma_sweep.py --n 1000000 --iters 20 --k 5 "A=old;CNMF_LIB=$T;A=old;CNMF_LIB=$T"
} > gpurun_out/ab_stg.txt 2>&1
cat gpurun_out/ab_stg.txt
