#!/bin/bash
mkdir -p gpurun_out
python - <<PY
import time, numpy as np, os
import cnmf
from cnmf_b200 import synth
from cnmf_b200._api import nmf as _n
from oracle import cnmf_oracle as orc
X, V = synth.spectra(1000, 81, 2, seed=1, vmode="ones")
W0, H0 = orc.init_factors(X, V, 2, seed=2)
for it in (500, 500, 500, 50, 50):
    t0 = time.perf_counter()
    cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=it, tol=0)
    dt = time.perf_counter() - t0
    print(it, "call %.2f ms" % (dt * 1e3), {k: round(v * 1e3, 3) for k, v in _n.last_timing.items()})
PY
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu4.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu4.log
