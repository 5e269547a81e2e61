#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/sweep.py "f64:;f64:DBG=2;f64:;f64:DBG=2;f32:;f32:DBG=2;f32:;f32:DBG=2" > gpurun_out/sweep26.txt 2>&1
timeout 900 python tools/sweep.py --n 10000000 "f64:;f64:DBG=2;f64:;f64:DBG=2" >> gpurun_out/sweep26.txt 2>&1
cat gpurun_out/sweep26.txt | grep "timing\|FAILED\|STALL" | cut -c1-160
