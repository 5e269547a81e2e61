#!/bin/bash
mkdir -p gpurun_out
export CNMF_RESIDENT=1
for bg in 8 4; do
export CNMF_RES_BG=$bg
echo "CNMF_RES_BG=$bg"
timeout 300 python - 2>&1 <<'PY' | grep -v "^NVIDIA"
import sys, os
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
import resident_check as rc
import bench_configs as bc
bc.batched("4: 4096 x (512 x 81) k=3 f32 batched", 4096, 512, 81, 3, np.float32, 40)
rc.one(1000, 81, 2, np.float32, 500)
PY
done > gpurun_out/resident_ab.log
cat gpurun_out/resident_ab.log
