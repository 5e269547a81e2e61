#!/bin/bash
# one full ncu capture (with source) of the fp32 stream kernel at config 3's shape (1.7M x 184, k = 8)
mkdir -p gpurun_out
cat > /tmp/c3.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from cnmf_b200 import Plan, _lib, synth
dev = torch.device("cuda", 0)
rows, m, k = 1_700_000, 184, 8
ld = _lib.padded_ld(m, _lib.F32)
X = torch.empty((rows, ld), dtype=torch.float32, device=dev); V = torch.empty_like(X)
hs = torch.from_numpy(synth.basis(k, m, 0)).to(dev)
_lib.check(_lib.load().cnmf_synth(_lib.F32, X.data_ptr(), V.data_ptr(), rows, 0, m, ld, k, hs.data_ptr(), 0, 2, 0.1, None))
rng = np.random.default_rng(1)
W = torch.from_numpy(rng.random((rows, k)).astype(np.float32)).to(dev); H = torch.from_numpy(rng.random((k, m)).astype(np.float32)).to(dev)
hist = torch.zeros(8, dtype=torch.float64, device=dev)
plan = Plan(rows, m, k, np.float32, ld=ld)
plan.solve(X, V, W, H, hist, 4); torch.cuda.synchronize(); print(plan.info, hist.cpu().numpy()[:5])
PY
timeout 300 python /tmp/c3.py > gpurun_out/c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 2 -c 1 -f -o gpurun_out/c3_prof python /tmp/c3.py > gpurun_out/c3_ncu.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/c3_plain.log
