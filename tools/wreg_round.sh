#!/bin/bash
# one GPU call: parity of the WREG variants, then timing sweeps next to the default kernel
mkdir -p gpurun_out
timeout 420 python tests/wreg_check.py > gpurun_out/wreg_parity.log 2>&1; echo "parity exit $?" >> gpurun_out/wreg_parity.log
S64="f64:;f64:WREG=1;f64:WREG=1,CR=4;f64:WREG=1,CR=1;f64:WREG=1,NOPROD=2;f64:WREG=1,NW1=5;f64:DBG=1;f64:WREG=1,DBG=1;f64:WREG=1,ST=4"
S32="f32:;f32:WREG=1;f32:WREG=1,CR=4;f32:WREG=1,CR=1;f32:WREG=1,NOPROD=2;f32:DBG=1;f32:WREG=1,DBG=1"
timeout 600 python tools/sweep.py "$S64;$S32" > gpurun_out/wreg_sweep.log 2>&1
timeout 300 python tools/sweep.py --n 1700000 --m 184 --k 8 "f32:;f32:WREG=1;f32:WREG=1,CR=4;f32:WREG=1,C2=2;f32:WREG=1,CR=4,C2=2" >> gpurun_out/wreg_sweep.log 2>&1
timeout 200 python tools/sweep.py --n 100000 --m 301 --k 4 "f64:;f64:WREG=1;f64:WREG=1,CR=4;f64:WREG=1,CR=8" >> gpurun_out/wreg_sweep.log 2>&1
tail -n 60 gpurun_out/wreg_parity.log gpurun_out/wreg_sweep.log
