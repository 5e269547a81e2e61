#!/bin/bash
# Round 2, GPU visit 2: the column-split role W (G=4, tuning build): parity suite through it, then timing sweeps.
set -x
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
CNMF_TUNE=G=4 timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_local_group.py -x -q > gpurun_out/pytest_g4.log 2>&1; echo "pytest G=4 rc=$?" >> gpurun_out/pytest_g4.log
tail -15 gpurun_out/pytest_g4.log
CNMF_TUNE=G=4,NW1=4 timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -k "golden or every_k or all_k or ragged" > gpurun_out/pytest_g4_nw4.log 2>&1; echo "pytest G=4,NW1=4 rc=$?" >> gpurun_out/pytest_g4_nw4.log
tail -5 gpurun_out/pytest_g4_nw4.log
timeout 900 python tools/first_light.py --sweep "f64:;f64:G=4;f64:G=4,WBAL=0;f64:G=4,NW1=5;f64:G=4,NW1=4;f64:G=4,NW1=4,MAXW=10;f64:G=4,NW1=5,MAXW=11;f64:G=4,NW1=4,WBAL=0;f64:G=4,NT=2;f64:G=4,NW1=3,MAXW=9;f32:;f32:G=4;f32:G=4,RP1=1;f32:G=4,NW1=4;f32:G=4,WBAL=0" > gpurun_out/sweep2.txt 2>&1; echo "sweep rc=$?"
cat gpurun_out/sweep2.txt
timeout 600 python tools/first_light.py --sweep "f32:;f32:G=4;f32:G=4,RP1=1" --n 1700000 --m 184 --k 8 > gpurun_out/sweep2_c3.txt 2>&1; cat gpurun_out/sweep2_c3.txt
timeout 600 python tools/first_light.py --sweep "f64:;f64:G=4" --n 100000 --m 301 --k 4 > gpurun_out/sweep2_c2.txt 2>&1; cat gpurun_out/sweep2_c2.txt
