#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu3.log
python - <<PY
import time, numpy as np, os
import cnmf
from cnmf_b200 import synth
from oracle import cnmf_oracle as orc
X, V = synth.spectra(1000, 81, 2, seed=1, vmode="ones")
W0, H0 = orc.init_factors(X, V, 2, seed=2)
for route in ("0", "1", "0", "1"):
    os.environ["CNMF_ONE_CTA"] = route
    cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=500, tol=0)
    t0 = time.perf_counter()
    for _ in range(5):
        cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=500, tol=0)
    dt = (time.perf_counter() - t0) / 5
    print("route", route, "1000x81 k=2, 500 iterations, whole cnmf.nmf() call: %.2f ms (%.0f it/s end to end)" % (dt * 1e3, 500 / dt))
PY
