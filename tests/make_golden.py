"""Writes tests/golden/*.npz: oracle outputs for small seeded cases.

Run here (CPU) with `python tests/make_golden.py`.  Inputs are regenerated
from the seed by cnmf_b200.synth (host NumPy), so only outputs are stored.
The upstream snapshot holds no fixtures; these pin the
oracle against itself over time and give the GPU tests fixed targets.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cnmf_b200 import synth  # noqa: E402
from oracle import cnmf_oracle as orc  # noqa: E402

CASES = {
    # name: n, m, k, vmode, p_missing, dtype, iters, seed
    "c1_absorption_1000x81_k2_f64": (1000, 81, 2, "ones", 0.0, "float64", 500, 11),
    "c2_iop_2000x301_k4_f64": (2000, 301, 4, "ivar", 0.0, "float64", 100, 12),
    "c3_scene_3000x184_k8_f32": (3000, 184, 8, "mask", 0.1, "float32", 100, 13),
    "c4_station_512x81_k3_f32": (512, 81, 3, "ivar", 0.0, "float32", 100, 14),
    "c5_shard_3001x301_k6_f64": (3001, 301, 6, "ivar_mask", 0.05, "float64", 100, 15),
    "c5_shard_3001x301_k6_f32": (3001, 301, 6, "ivar_mask", 0.05, "float32", 100, 15),
}


def inputs(name):
    n, m, k, vmode, p, dtype, iters, seed = CASES[name]
    X, V = synth.spectra(n, m, k, seed=seed, vmode=vmode, p_missing=p, dtype=np.dtype(dtype))
    W0, H0 = orc.init_factors(X, V, k, seed=seed + 1, dtype=np.dtype(dtype))
    return X, V, W0, H0, k, iters


def main():
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    for name in CASES:
        X, V, W0, H0, k, iters = inputs(name)
        W, H, hist = orc.solve(X, V, W0, H0, iters)
        np.savez_compressed(os.path.join(out, name + ".npz"), W=W, H=H, hist=hist,
                            x_sum=np.float64(X.sum(dtype=np.float64)), v_sum=np.float64(V.sum(dtype=np.float64)))
        print(name, hist[0], hist[-1])


if __name__ == "__main__":
    main()
