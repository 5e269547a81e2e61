#!/bin/bash
mkdir -p gpurun_out
export CNMF_LIB=$PWD/cnmf_b200/libcnmf_b200_prof.so
export CNMF_RESIDENT=1
timeout 300 python - > gpurun_out/resident_prof.log 2>&1 <<'PY'
import sys, os
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
import resident_check as rc
import bench_configs as bc
rc.one(1000, 81, 2, np.float64, 500)
rc.one(1000, 81, 2, np.float32, 500)
bc.batched("4: 74 x (512 x 81) k=3 f32", 74, 512, 81, 3, np.float32, 40)
bc.batched("4: 37 x (512 x 81) k=3 f64", 37, 512, 81, 3, np.float64, 40)
PY
grep -v "^NVIDIA" gpurun_out/resident_prof.log | tail -20
