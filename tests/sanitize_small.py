"""Small solves for compute-sanitizer (one tool per GPU visit): single, ragged, batched, W-only."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cnmf
from cnmf_b200 import synth
from oracle import cnmf_oracle as orc

for (n, m, k, dt) in ((700, 81, 2, np.float64), (513, 301, 6, np.float32), (300, 184, 8, np.float32), (65, 33, 11, np.float64)):
    X, V = synth.spectra(n, m, k, seed=1, vmode="ivar_mask", p_missing=0.1, dtype=dt)
    W0, H0 = orc.init_factors(X, V, k, seed=2, dtype=dt)
    W, H, h = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=6, tol=0)
    Wr, Hr, hr = orc.solve(X, V, W0, H0, 6)
    print(n, m, k, np.dtype(dt).name, float(np.max(np.abs(W - Wr)) / np.max(Wr)), float(np.max(np.abs(h - hr) / hr)), flush=True)
Xb = np.stack([synth.spectra(130, 33, 3, seed=s, dtype=np.float32)[0] for s in range(3)])
W, H, h = cnmf.nmf(Xb, 3, max_iter=5, tol=0, seed=3)
print("batched", W.shape, np.isfinite(h).all(), flush=True)
