#!/bin/bash
mkdir -p gpurun_out
CNMF_TUNE=NOPROD=2,NT=2 timeout 120 python tools/first_light.py > gpurun_out/parity.txt 2>&1; echo "parity rc=$?"; tail -5 gpurun_out/parity.txt | cut -c1-150
timeout 900 python tools/sweep.py "f64:;f64:NOPROD=2;f64:NOPROD=2,NT=2;f64:NOPROD=2,NT=2,DBG=1;f64:NT=2;f32:;f32:NOPROD=2;f32:NOPROD=2,NT=2" > gpurun_out/sweep28.txt 2>&1
timeout 600 python tools/sweep.py --n 1700000 --m 184 --k 8 "f32:;f32:NOPROD=2" >> gpurun_out/sweep28.txt 2>&1
timeout 600 python tools/sweep.py --n 100000 --m 301 --k 4 "f64:;f64:NOPROD=2;f64:NOPROD=2,NT=2" >> gpurun_out/sweep28.txt 2>&1
cat gpurun_out/sweep28.txt | grep "timing\|FAILED\|STALL" | cut -c1-200
