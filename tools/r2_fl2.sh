#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/fl2_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/fl2_pytest.log; tail -3 gpurun_out/fl2_pytest.log
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
{
timeout 600 python tools/mma_sweep.py --f32 --n 1000000 --iters 20 --k 6 "CNMF_TUNE=FL=0;CNMF_TUNE=FL=8;CNMF_TUNE=FL=1024"
timeout 600 python tools/mma_sweep.py --f32 --n 1700000 --iters 20 --k 8 --m 184 "CNMF_TUNE=FL=0;CNMF_TUNE=FL=8;CNMF_TUNE=FL=1024"
timeout 600 python tools/mma_sweep.py --f32 --n 10000000 --iters 30 --k 6 "CNMF_TUNE=FL=0"
} > gpurun_out/fl2_k1.txt 2>&1
cat gpurun_out/fl2_k1.txt
