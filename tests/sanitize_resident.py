"""Small resident-kernel and streaming solves for compute-sanitizer (racecheck / memcheck)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cnmf  # noqa: E402
from cnmf_b200 import synth  # noqa: E402
from oracle import cnmf_oracle as orc  # noqa: E402

for (n, m, k, dt) in ((1000, 81, 2, np.float64), (700, 81, 3, np.float32), (60, 20, 4, np.float64)):
    X, V = synth.spectra(n, m, k, seed=1, vmode="ivar_mask", p_missing=0.05, dtype=dt)
    W0, H0 = orc.init_factors(X, V, k, seed=2, dtype=dt)
    W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=4, tol=0)
    Wr, Hr, hr = orc.solve(X, V, W0, H0, 4)
    print(n, m, k, np.dtype(dt).name, "max rel W", float(np.max(np.abs(W - Wr)) / np.max(np.abs(Wr))), flush=True)
B = 3
cs = []
for b in range(B):
    X, V = synth.spectra(512, 81, 3, seed=10 + b, vmode="ivar", dtype=np.float32)
    W0, H0 = orc.init_factors(X, V, 3, seed=20 + b, dtype=np.float32)
    cs.append((X, V, W0, H0))
X, V, W0, H0 = (np.stack([c[i] for c in cs]) for i in range(4))
W, H, hist = cnmf.nmf(X, 3, weights=V, W0=W0, H0=H0, max_iter=3, tol=0)
print("batched", W.shape, hist[:, -1], flush=True)
