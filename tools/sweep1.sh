#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size" > gpurun_out/full.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/full.log
