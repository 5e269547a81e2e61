#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sweep.py "f64:MAXW=16,NT=2;f64:MAXW=16,NT=2,NW1=7;f64:MAXW=16,NT=2,NW1=8;f64:MAXW=16,NT=2,DBG=1;f64:MAXW=16,NT=1;f32:MAXW=16,NT=2,RP1=1;f32:MAXW=16,NT=2,RP1=1,NW1=7" > gpurun_out/sweep16.txt 2>&1
cat gpurun_out/sweep16.txt | cut -c1-250
