"""Timing of the fp64 stream pass per environment setting (CNMF_MMA, CNMF_MMA_TW, CNMF_MMA_NW1, CNMF_MMA_ST), one child
process per setting (a hang costs one timeout).  Parity of the settings is tests/mma_check.py's business.
    python tools/mma_sweep.py [--n ROWS] [--iters N] [--k K] [--m M] "CNMF_MMA=0;CNMF_MMA=1,CNMF_MMA_TW=4"
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(rows, iters, m, k, f32=False):
    import torch
    from cnmf_b200 import Plan, _lib, synth
    dev = torch.device("cuda", 0)
    code, tdt, ndt = (_lib.F32, torch.float32, np.float32) if f32 else (_lib.F64, torch.float64, np.float64)
    ld = _lib.padded_ld(m, code)
    X = torch.empty((rows, ld), dtype=tdt, device=dev)
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, 0)).to(dev)
    _lib.check(_lib.load().cnmf_synth(code, X.data_ptr(), V.data_ptr(), rows, 0, m, ld, k, hs.data_ptr(), 0, 3, 0.05, None))
    rng = np.random.default_rng(1)
    W = torch.from_numpy(rng.random((rows, k)).astype(ndt)).to(dev)
    H = torch.from_numpy(rng.random((k, m)).astype(ndt)).to(dev)
    hist = torch.zeros(iters + 1, dtype=torch.float64, device=dev)
    plan = Plan(rows, m, k, ndt, ld=ld)
    info = plan.info
    plan.solve(X, V, W, H, hist, 5)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.solve(X, V, W, H, hist, iters)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (iters + 1))
    h = hist.cpu().numpy()
    print("timing %d x %d k=%d: %.3f ms/iter %.0f GB/s variant=%d thr=%d st=%d chi -> %.8g mono=%s" % (
        rows, m, k, best, info["bytes_per_iter"] / best / 1e6, info["variant"], info["threads"], info["stages"],
        h[-1], bool(np.all(np.diff(h) <= (1e-6 if f32 else 1e-9) * h[:-1]))), flush=True)
    plan.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    opt = lambda name, d: int(args[args.index(name) + 1]) if name in args else d
    n, iters, m, k = opt("--n", 1_000_000), opt("--iters", 20), opt("--m", 301), opt("--k", 6)
    if "--child" in args:
        child(n, iters, m, k, "--f32" in args)
        sys.exit(0)
    for cfg in args[-1].split(";"):
        env = dict(os.environ)
        for kv in cfg.split(","):
            if "=" in kv:
                a, b = kv.split("=", 1)
                env[a] = b
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--n", str(n), "--iters", str(iters),
                                "--m", str(m), "--k", str(k)] + (["--f32"] if "--f32" in args else []), env=env, capture_output=True, text=True, timeout=180)
            print("%-44s %s" % (cfg, (r.stdout.strip() or r.stderr[-800:])), flush=True)
        except subprocess.TimeoutExpired:
            print("%-44s TIMEOUT" % cfg, flush=True)
