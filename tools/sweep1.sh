#!/bin/bash
for w in 1 0 1 0; do
echo "== wpre $w"
CNMF_BATCHED_WPRE=$w timeout 300 python - <<PY 2>&1 | grep "^4:" | cut -c1-110
import sys, numpy as np
sys.path.insert(0, "tools")
import bench_configs as bc
bc.batched("4: f32", 4096, 512, 81, 3, np.float32, 40)
bc.batched("4: f64", 4096, 512, 81, 3, np.float64, 40)
PY
done
