#!/bin/bash
set -x
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 900 python tools/first_light.py --sweep "f64:;f64:GATENS=100;f64:GATENS=300;f64:LOOKAHEAD=0;f64:GATENS=200,LOOKAHEAD=0;f64:NW1=5;f64:MAXW=16;f64:NT=4,NOPROD=2;f64:G=16,RP1=2;f64:ST=4;f32:;f32:GATENS=200;f32:LOOKAHEAD=0;f32:G=16;f32:MAXW=16" > gpurun_out/sweep5.txt 2>&1
cat gpurun_out/sweep5.txt
unset CNMF_LIB
P="python tools/first_light.py --prof f64"
timeout 600 $P > gpurun_out/plain_prof_f64.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 4 -c 1 -f -o gpurun_out/prof_f64_v2 $P > gpurun_out/ncu_prof_f64_v2.log 2>&1
echo "ncu rc=$?"
