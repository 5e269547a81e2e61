#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sweep.py "f64:;f64:RP1=2,NW1=4;f64:RP1=2,NW1=4,NT=2;f64:RP1=2,NW1=5,NT=2;f64:RP1=2,NW1=3,NT=2;f64:RP1=2,NW1=4,NT=2,DEAL=0;f64:RP1=2,NW1=4,ST=5,NT=2,PF=4" > gpurun_out/sweep18.txt 2>&1
cat gpurun_out/sweep18.txt | cut -c1-250
