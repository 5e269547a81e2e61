#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sweep.py --n 100000 --m 301 --k 4 "f64:;f64:DBG=2;f64:;f64:DBG=2;f64:;f64:DBG=2" > gpurun_out/sweep27.txt 2>&1
timeout 600 python tools/sweep.py --n 1700000 --m 184 --k 8 "f32:;f32:DBG=2;f32:;f32:DBG=2" >> gpurun_out/sweep27.txt 2>&1
timeout 600 python tools/sweep.py --n 10000000 "f32:;f32:DBG=2;f64:;f64:DBG=2" >> gpurun_out/sweep27.txt 2>&1
grep "timing\|FAILED\|STALL" gpurun_out/sweep27.txt | cut -c1-130
