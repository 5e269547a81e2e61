#!/bin/bash
# one full ncu capture of the resident cluster kernel: config 1 (one problem, 8 CTAs) and 74 problems of config 4
mkdir -p gpurun_out
cat > /tmp/res_run.py <<'PY'
import sys
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
import resident_check as rc
import bench_configs as bc
if sys.argv[1] == "one":
    rc.one(1000, 81, 2, np.float64, 200)
else:
    bc.batched("74 x (512 x 81) k=3 f32", 74, 512, 81, 3, np.float32, 40)
PY
export CNMF_RESIDENT=1
for w in one batch; do
timeout 300 python /tmp/res_run.py $w > gpurun_out/res_plain_$w.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resident_kernel -s 1 -c 1 -f -o gpurun_out/prof_resident_$w python /tmp/res_run.py $w > gpurun_out/ncu_resident_$w.log 2>&1
echo "ncu resident $w rc=$?"
done
