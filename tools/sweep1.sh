#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/first_light.py > gpurun_out/parity.txt 2>&1; tail -6 gpurun_out/parity.txt | cut -c1-150
timeout 900 python tools/sweep.py "f64:;f64:DBG=1;f64:DEAL=0;f32:;f32:DBG=1" > gpurun_out/sweep6.txt 2>&1
cat gpurun_out/sweep6.txt | cut -c1-250
