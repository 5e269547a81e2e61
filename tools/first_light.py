"""GPU bring-up timings and the sweep / profiling entry points (`--sweep`, `--prof`).  The parity
checks that went with them live in tests/bringup_parity.py (the oracle is test infrastructure)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cnmf_b200 import Plan, _lib, synth  # noqa: E402


def rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b)) / np.max(np.abs(b)))


def timing(n, m, k, dtype, iters=20, tune=None):
    if tune is not None:
        os.environ["CNMF_TUNE"] = tune
    code = _lib.F32 if dtype == np.float32 else _lib.F64
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    ld = _lib.padded_ld(m, code)
    dev = torch.device("cuda", 0)
    X = torch.empty((n, ld), dtype=tdt, device=dev)
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, 0)).to(dev)
    _lib.check(_lib.load().cnmf_synth(code, X.data_ptr(), V.data_ptr(), n, 0, m, ld, k, hs.data_ptr(), 0, 3, 0.05, None))
    rng = np.random.default_rng(1)
    W = torch.from_numpy(rng.random((n, k)).astype(dtype)).to(dev)
    H = torch.from_numpy(rng.random((k, m)).astype(dtype)).to(dev)
    hist = torch.zeros(iters + 1, dtype=torch.float64, device=dev)
    plan = Plan(n, m, k, dtype, ld=ld)
    info = plan.info
    plan.solve(X, V, W, H, hist, 5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    plan.solve(X, V, W, H, hist, iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (iters + 1)   # + the prologue pass
    gbs = info["bytes_per_iter"] / ms / 1e6
    h = hist.cpu().numpy()
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    plan.chi2(X, V, W, H, out)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        plan.chi2(X, V, W, H, out)
    e1.record()
    torch.cuda.synchronize()
    ms_h = e0.elapsed_time(e1) / 10
    print("timing n=%d m=%d k=%d %s tune=%s: %.3f ms/iter  %.0f GB/s (%.1f%% of 6447)  grid=%d thr=%d st=%d R=%d S2=%d  chi %.4g->%.4g mono=%s  Honly=%.3f ms" % (
        n, m, k, np.dtype(dtype).name, tune, ms, gbs, gbs / 64.475, info["grid"], info["threads"], info["stages"],
        info["tile_rows"], info["h_streams"], h[0], h[-1], bool(np.all(np.diff(h) <= 1e-6 * h[:-1])), ms_h), flush=True)
    plan.close()
    os.environ.pop("CNMF_TUNE", None)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), flush=True)
    if "--sweep" in sys.argv:
        opt = lambda name, d: int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else d
        n, m, k, its = opt("--n", 1_000_000), opt("--m", 301), opt("--k", 6), opt("--iters", 20)
        for tune in sys.argv[sys.argv.index("--sweep") + 1].split(";"):
            dt = np.float64 if tune.startswith("f64:") else np.float32
            try:
                timing(n, m, k, dt, iters=its, tune=tune.split(":", 1)[1])
            except Exception as exc:
                print("FAILED %s: %r" % (tune, exc), flush=True)
        sys.exit(0)
    if "--prof" in sys.argv:
        dt = np.float32 if sys.argv[sys.argv.index("--prof") + 1] == "f32" else np.float64
        timing(1_000_000, 301, 6, dt, iters=3, tune=os.environ.get("CNMF_TUNE"))
        sys.exit(0)
    if "--time" in sys.argv:
        for dt in (np.float32, np.float64):
            timing(1_000_000, 301, 6, dt)
        timing(1_000_000, 301, 6, np.float32, tune="GRAPH=0")
        timing(1_000_000, 301, 6, np.float32, tune="C2=2,S2=4")
        timing(1_000_000, 301, 6, np.float32, tune="G=16")
        timing(1_000_000, 301, 6, np.float32, tune="NW1=4")
        timing(1_000_000, 301, 6, np.float64, tune="C2=2,S2=2")
        timing(1_000_000, 301, 6, np.float64, tune="RP1=2")
        timing(100_000, 301, 4, np.float64, iters=100)
        timing(1_700_000, 184, 8, np.float32)
