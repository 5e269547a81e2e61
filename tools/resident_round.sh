#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ragged or all_k or routes or batched or restarts" > gpurun_out/resident_tests.log 2>&1
tail -n 15 gpurun_out/resident_tests.log
timeout 600 python tools/resident_check.py > gpurun_out/resident_check.log 2>&1
cat gpurun_out/resident_check.log | grep -v "^NVIDIA"
