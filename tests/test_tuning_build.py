"""The tuning build (cnmf_b200/build.py --tune -> libcnmf_b200_tune.so) carries kernel variants that only the
CNMF_TUNE sweep knobs can select (16 lanes per row, register-resident role W, producer-less forms, the 512-thread
cap).  They never ship, but they must stay correct while they exist: each variant runs a handful of the parity cases
in a child process (the library path and the knobs are read once per process)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TUNE_LIB = os.path.join(ROOT, "cnmf_b200", "libcnmf_b200_tune.so")

CHILD = r"""
import sys
import numpy as np
sys.path.insert(0, %r)
import cnmf
from cnmf_b200 import synth, Plan
from oracle import cnmf_oracle as orc
for (n, m, k, dt, tol) in ((3001, 301, 6, np.float64, 1e-10), (2050, 301, 6, np.float32, 1e-4), (4000, 184, 8, np.float32, 1e-4),
                           (1777, 301, 4, np.float64, 1e-10)):
    X, V = synth.spectra(n, m, k, seed=n, vmode="ivar_mask", p_missing=0.07, dtype=dt)
    W0, H0 = orc.init_factors(X, V, k, seed=n + 1, dtype=dt)
    try:
        W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=12, tol=0)
    except Exception as exc:
        if "no compiled" in str(exc) or "error -3" in str(exc):
            print("SKIP", n, m, k, dt.__name__, str(exc)[:80])
            continue
        raise
    Wr, Hr, hr = orc.solve(X, V, W0, H0, 12)
    rel = lambda a, b: float(np.max(np.abs(a.astype(np.float64) - b)) / np.max(np.abs(b)))
    assert rel(W, Wr) <= tol and rel(H, Hr) <= tol and float(np.max(np.abs(hist - hr) / hr)) <= tol, (n, m, k, rel(W, Wr), rel(H, Hr))
    print("OK", n, m, k, dt.__name__)
""" % ROOT

VARIANTS = ["G=16", "G=16,RP1=2", "WREG=1", "NOPROD=1", "NOPROD=2,NT=4", "MAXW=16", "NT=2", "NW1=4", "ST=3"]
# environment settings that exist in the tuning build only: the fp32 tensor-path kernel (TF32 x 3, mma32_kernel.cuh) with
# 4 / 2 / 1 column parts, the 17-warp form of the fp64 tensor-path kernel
ENV_VARIANTS = [{"CNMF_MMA32": "1"}, {"CNMF_MMA32": "1", "CNMF_MMA_TW": "2"}, {"CNMF_MMA32": "1", "CNMF_MMA_TW": "1"},
                {"CNMF_MMA": "4"}]


@pytest.mark.gpu
@pytest.mark.parametrize("tune", VARIANTS)
def test_tuning_variant_matches_oracle(tune):
    if not os.path.exists(TUNE_LIB):
        pytest.skip("tuning build absent (python cnmf_b200/build.py --tune)")
    env = dict(os.environ, CNMF_LIB=TUNE_LIB, CNMF_TUNE=tune, CNMF_ONE_CTA="0")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("extra", ENV_VARIANTS, ids=lambda e: ",".join("%s=%s" % kv for kv in e.items()))
def test_tuning_only_kernel_matches_oracle(extra):
    if not os.path.exists(TUNE_LIB):
        pytest.skip("tuning build absent (python cnmf_b200/build.py --tune)")
    env = dict(os.environ, CNMF_LIB=TUNE_LIB, CNMF_ONE_CTA="0", **extra)
    env.pop("CNMF_TUNE", None)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("OK") == 4, r.stdout
