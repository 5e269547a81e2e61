"""GPU bring-up parity: small solves against the oracle with diagnostics, one line per check
(run under `timeout` on the GPU box: python tests/bringup_parity.py)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cnmf  # noqa: E402
from cnmf_b200 import synth  # noqa: E402
from oracle import cnmf_oracle as orc  # noqa: E402


def rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b)) / np.max(np.abs(b)))


def parity(n, m, k, dtype, iters, vmode="ivar_mask"):
    X, V = synth.spectra(n, m, k, seed=3, vmode=vmode, p_missing=0.05, dtype=dtype)
    W0, H0 = orc.init_factors(X, V, k, seed=4, dtype=dtype)
    t0 = time.time()
    W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
    t1 = time.time()
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    print("parity n=%d m=%d k=%d %s it=%d: W %.2e H %.2e chi %.2e  chi0 %.6g/%.6g  (%.2fs)" % (
        n, m, k, np.dtype(dtype).name, iters, rel(W, Wr), rel(H, Hr), float(np.max(np.abs(hist - hr) / hr)),
        hist[0], hr[0], t1 - t0), flush=True)



if __name__ == "__main__":
    parity(64, 20, 2, np.float64, 5)
    parity(1000, 81, 2, np.float64, 50)
    parity(1000, 81, 2, np.float32, 50)
    parity(5000, 301, 6, np.float64, 30)
    parity(5000, 301, 6, np.float32, 30)
    parity(40000, 184, 8, np.float32, 10, "mask")
