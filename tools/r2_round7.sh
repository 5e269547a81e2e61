#!/bin/bash
# Variant ranking at the BENCH size under sustained load (the 1M-row sweeps never reach the 1 kW power cap)
set -x
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_tune.so
timeout 1500 python tools/first_light.py --n 10000000 --iters 40 --sweep "f64:;f64:G=16,RP1=2;f64:WREG=1;f64:WREG=1,CR=4;f64:NT=4,NOPROD=2;f64:NW1=5;f64:NW1=4;f64:C2=1;f64:MAXW=16;f64:ST=4;f64:" > gpurun_out/sweep7.txt 2>&1
cat gpurun_out/sweep7.txt
unset CNMF_LIB
timeout 600 python tools/first_light.py --n 10000000 --iters 40 --sweep "f64:;f32:" > gpurun_out/sweep7_product.txt 2>&1
cat gpurun_out/sweep7_product.txt
