"""Every BASELINE.json config on one B200: iterations/s, achieved GB/s and the fraction of the
measured HBM roofline (algorithmic bytes = rows * M * 2 * sizeof(T) per iteration), device-resident
operands, CUDA events.  Parity of these shapes is the business of tests/; this only times them.

    python tools/bench_configs.py [--iters N]  ->  one line per config + a JSON summary line
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cnmf_b200 import Plan, _lib, synth  # noqa: E402

try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0
DEV = torch.device("cuda", 0)
LIB = _lib.load()


def operands(n, m, k, dtype, vmode, p_missing, batch=1, seed=7):
    code = _lib.F32 if dtype == np.float32 else _lib.F64
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    ld = _lib.padded_ld(m, code)
    rows = n * batch
    X = torch.empty((rows, ld), dtype=tdt, device=DEV)
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, seed)).to(DEV)
    _lib.check(LIB.cnmf_synth(code, X.data_ptr(), V.data_ptr(), rows, 0, m, ld, k, hs.data_ptr(), seed,
                              synth.VMODES[vmode], p_missing, None))
    rng = np.random.default_rng(seed + 1)
    W = torch.from_numpy(rng.random((rows, k)).astype(dtype)).to(DEV)
    H = torch.from_numpy(rng.random((batch * k, m)).astype(dtype)).to(DEV)
    return X, V, W, H, ld


def single(name, n, m, k, dtype, vmode, p, iters):
    X, V, W, H, ld = operands(n, m, k, dtype, vmode, p)
    hist = torch.zeros(2 * iters + 8, dtype=torch.float64, device=DEV)
    plan = Plan(n, m, k, dtype, ld=ld)
    info = plan.info
    plan.solve(X, V, W, H, hist, 5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    plan.solve(X, V, W, H, hist, iters, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    h = hist.cpu().numpy()[:iters + 6]
    ok = bool(np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]))
    plan.close()
    return report(name, info["bytes_per_iter"], ms, ok, "grid=%d thr=%d st=%d R=%d" % (
        info["grid"], info["threads"], info["stages"], info["tile_rows"]))


def batched(name, b, n, m, k, dtype, iters):
    X, V, W, H, ld = operands(n, m, k, dtype, "ivar", 0.0, batch=b)
    hist = torch.zeros((b, iters + 1), dtype=torch.float64, device=DEV)
    niter = torch.zeros(b, dtype=torch.int32, device=DEV)
    plan = Plan(n, m, k, dtype, ld=ld, batch=b)
    info = plan.info
    W1, H1 = W.clone(), H.clone()
    plan.solve_batched(X, V, W1, H1, hist, niter, 3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    plan.solve_batched(X, V, W, H, hist, niter, iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    h = hist.cpu().numpy()
    ok = bool(np.all(np.isfinite(h)) and np.all(np.diff(h, axis=1) <= 1e-6 * h[:, :-1]))
    plan.close()
    return report(name, info["bytes_per_iter"], ms, ok, "%s, thr=%d smem=%d (effective GB/s)" % (
        "resident clusters of %d CTAs" % (info["variant"] - 700000) if info["variant"] >= 700000 else "one CTA per problem, tiles re-read from L2",
        info["threads"], info["smem_bytes"]))


def resident_single(name, n, m, k, dtype, vmode, p, iters):
    """one problem on the resident cluster kernel (the entry point's default route when it fits)"""
    X, V, W, H, ld = operands(n, m, k, dtype, vmode, p)
    hist = torch.zeros((1, iters + 1), dtype=torch.float64, device=DEV)
    niter = torch.zeros(1, dtype=torch.int32, device=DEV)
    plan = Plan(n, m, k, dtype, ld=ld, variant=_lib.VARIANT_RESIDENT)
    info = plan.info
    plan.solve_batched(X, V, W.clone(), H.clone(), hist, niter, 3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    plan.solve_batched(X, V, W, H, hist, niter, iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    h = hist.cpu().numpy()[0]
    ok = bool(np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]))
    plan.close()
    return report(name, info["bytes_per_iter"], ms, ok, "resident cluster kernel: %d CTAs x %d rows, thr=%d smem=%d, one launch for all iterations" % (
        info["grid"], info["tile_rows"], info["threads"], info["smem_bytes"]))


def report(name, bytes_iter, ms, ok, note):
    gbs = bytes_iter / ms / 1e6
    r = {"config": name, "ms_per_iter": ms, "it_per_s": 1e3 / ms, "gb_per_s": gbs, "frac_of_measured_peak": gbs / PEAK,
         "frac_of_8000": gbs / 8000.0, "chi2_monotone": ok, "note": note}
    print("%-46s %9.3f ms/it %10.1f it/s %8.0f GB/s  %5.1f%% of %.0f  (%4.1f%% of 8000)  mono=%s  %s" % (
        name, ms, 1e3 / ms, gbs, 100 * gbs / PEAK, PEAK, 100 * gbs / 8000.0, ok, note), flush=True)
    return r


if __name__ == "__main__":
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 40
    print(torch.cuda.get_device_name(0), "peak", PEAK, "GB/s", flush=True)
    out = []
    out.append(single("1: 1,000 x 81 k=2 f64 V=1 (launch-bound)", 1000, 81, 2, np.float64, "ones", 0.0, 200))
    out.append(resident_single("1: 1,000 x 81 k=2 f64 V=1 (default route: resident)", 1000, 81, 2, np.float64, "ones", 0.0, 500))
    out.append(single("2: 100,000 x 301 k=4 f64 ivar", 100_000, 301, 4, np.float64, "ivar", 0.0, 200))
    out.append(single("3: 1,700,000 x 184 k=8 f32 mask 10%", 1_700_000, 184, 8, np.float32, "mask", 0.10, iters))
    out.append(batched("4: 4096 x (512 x 81) k=3 f32 batched", 4096, 512, 81, 3, np.float32, iters))
    out.append(batched("4: 4096 x (512 x 81) k=3 f64 batched", 4096, 512, 81, 3, np.float64, iters))
    out.append(single("5: 10,000,000 x 301 k=6 f32 ivar_mask 5%", 10_000_000, 301, 6, np.float32, "ivar_mask", 0.05, iters))
    out.append(single("5: 10,000,000 x 301 k=6 f64 ivar_mask 5%", 10_000_000, 301, 6, np.float64, "ivar_mask", 0.05, iters))
    print(json.dumps({"when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "peak_gbs": PEAK, "configs": out}))
