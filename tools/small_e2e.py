"""End-to-end time of one cnmf.nmf() call on the upstream package's example size (1 000 x 81, k = 2,
V = 1, NumPy in / NumPy out, warm context): default route (resident cluster kernel) and CNMF_ONE_CTA=0
(streaming plan)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cnmf  # noqa: E402
from cnmf_b200 import init_factors, synth  # noqa: E402

X, V = synth.spectra(1000, 81, 2, seed=1, vmode="ones", dtype=np.float64)
W0, H0 = init_factors(1000, 81, 2, float(X.mean()), seed=2, dtype=np.float64)
for route in (None, "0"):
    if route is None:
        os.environ.pop("CNMF_ONE_CTA", None)
    else:
        os.environ["CNMF_ONE_CTA"] = route
    for iters in (50, 500):
        for _ in range(3):
            cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
        t = []
        for _ in range(20):
            t0 = time.perf_counter()
            W, H, hist = cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
            t.append(time.perf_counter() - t0)
        print("route=%s iterations=%d: median %.3f ms per call (min %.3f)  phases %s" % (
            "default(resident)" if route is None else "streaming", iters, 1e3 * np.median(t), 1e3 * min(t),
            {k: round(1e3 * v, 3) for k, v in cnmf.nmf.last_timing.items()}), flush=True)
