"""Small solves of the fp64 tensor-path kernel for compute-sanitizer (run under `compute-sanitizer --tool memcheck|racecheck|synccheck`,
one tool per call): both forms, 1 / 2 / 4 column parts, ragged rows and widths, an odd rows x k product, few SMs so that the ring wraps."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cnmf  # noqa: E402
from cnmf_b200 import synth  # noqa: E402
from oracle import cnmf_oracle as orc  # noqa: E402

os.environ["CNMF_ONE_CTA"] = "0"
for mode, tw in (("3", "4"), ("3", "2"), ("1", "4"), ("3", "1")):
    os.environ["CNMF_MMA"], os.environ["CNMF_MMA_TW"] = mode, tw
    for (n, m, k) in ((2003, 301, 6), (77, 81, 7), (9, 20, 3)):
        X, V = synth.spectra(n, m, k, seed=1, vmode="ivar_mask", p_missing=0.1, dtype=np.float64)
        W0, H0 = orc.init_factors(X, V, k, seed=2, dtype=np.float64)
        W, H, h = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=3, tol=0)
        Wr, Hr, hr = orc.solve(X, V, W0, H0, 3)
        print(mode, tw, n, m, k, float(np.max(np.abs(W - Wr)) / np.max(Wr)), float(np.max(np.abs(h - hr) / hr)), flush=True)
