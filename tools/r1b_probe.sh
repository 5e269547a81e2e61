#!/bin/bash
# resource probes of the diagnostics build: sweep passed as $1
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so
timeout 900 python tools/sweep.py $2 "$1" 2>&1 | grep -E "^timing|^FAILED|^STALLED|^NVIDIA" > gpurun_out/probe_last.log
cat gpurun_out/probe_last.log
