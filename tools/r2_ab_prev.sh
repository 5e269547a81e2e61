#!/bin/bash
# same-box A/B of two PRODUCT builds: cnmf_b200/libcnmf_b200_prev.so (the build before a change) against libcnmf_b200.so
mkdir -p gpurun_out
P=$PWD/cnmf_b200/libcnmf_b200_prev.so
{
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 6 "CNMF_LIB=$P;NEW=1;CNMF_LIB=$P;NEW=1"
timeout 300 python tools/mma_sweep.py --n 1000000 --iters 20 --k 8 "CNMF_LIB=$P;NEW=1"
timeout 400 python tools/mma_sweep.py --n 10000000 --iters 30 --k 6 "CNMF_LIB=$P;NEW=1;CNMF_LIB=$P;NEW=1"
} > gpurun_out/ab_prev.txt 2>&1
cat gpurun_out/ab_prev.txt
timeout 600 python -m pytest tests/test_mma_kernel.py -x -q -m gpu 2>&1 | tail -2
