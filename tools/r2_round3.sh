#!/bin/bash
# Round 2, GPU visit: role H software-pipelined row loop (HMODE=1, tuning build): parity through it, then A/B timings.
set -x
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
CNMF_TUNE=HMODE=1,WMODE=1 timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_h1.log 2>&1; echo "pytest HMODE=1 rc=$?" >> gpurun_out/pytest_h1.log
tail -5 gpurun_out/pytest_h1.log
timeout 900 python tools/first_light.py --sweep "f64:;f64:HMODE=1;f64:WMODE=1;f64:HMODE=1,WMODE=1;f64:HMODE=1,WMODE=1,MAXW=16;f64:MAXW=16;f64:HMODE=1,WMODE=1,NT=4,NOPROD=2;f64:HMODE=1,WMODE=1,G=16,RP1=2;f64:HMODE=1,WMODE=1,NW1=5;f32:;f32:HMODE=1;f32:WMODE=1;f32:HMODE=1,WMODE=1;f32:HMODE=1,WMODE=1,RP1=1" > gpurun_out/sweep3.txt 2>&1; echo "sweep rc=$?"
cat gpurun_out/sweep3.txt
timeout 600 python tools/first_light.py --sweep "f32:;f32:HMODE=1;f32:HMODE=1,WMODE=1;f32:HMODE=1,WMODE=1,RP1=1" --n 1700000 --m 184 --k 8 > gpurun_out/sweep3_c3.txt 2>&1; cat gpurun_out/sweep3_c3.txt
timeout 600 python tools/first_light.py --sweep "f64:;f64:HMODE=1,WMODE=1" --n 100000 --m 301 --k 4 > gpurun_out/sweep3_c2.txt 2>&1; cat gpurun_out/sweep3_c2.txt
