#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 500 python tools/bench_configs.py > gpurun_out/configs.txt 2>&1; cut -c1-200 gpurun_out/configs.txt | head -9
timeout 300 python tools/sweep.py --n 100000 --m 301 --k 4 "f64:;f64:RP1=2;f64:C2=1;f64:NW1=5" > gpurun_out/sweep13.txt 2>&1; cut -c1-220 gpurun_out/sweep13.txt
