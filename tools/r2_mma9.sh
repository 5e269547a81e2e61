#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_mma_kernel.py -x -q -m gpu > gpurun_out/mma9_pytest_mma.log 2>&1; echo "rc=$?" >> gpurun_out/mma9_pytest_mma.log
tail -15 gpurun_out/mma9_pytest_mma.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/mma9_pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/mma9_pytest_gpu.log
tail -15 gpurun_out/mma9_pytest_gpu.log
