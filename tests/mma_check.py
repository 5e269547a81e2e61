"""Bring-up of the fp64 tensor-path kernel (mma_kernel.cuh): parity against the oracle on small ragged shapes,
then timing, one child process per environment setting so that a hang costs one timeout, not the visit.
    python tests/mma_check.py [--n ROWS] [--iters N] "CNMF_MMA=0;CNMF_MMA=1;CNMF_MMA=1,CNMF_MMA_TW=4"
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b)) / np.max(np.abs(b)))


def child(n, iters, f32=False):
    import torch
    import cnmf
    from cnmf_b200 import Plan, _lib, synth
    from oracle import cnmf_oracle as orc
    os.environ["CNMF_ONE_CTA"] = "0"
    for (rows, m, k, its, upd) in ((64, 20, 2, 5, "both"), (5003, 301, 6, 12, "both"), (2001, 81, 4, 10, "both"), (1237, 184, 8, 8, "both"),
                                   (999, 301, 5, 6, "W"), (1000, 301, 6, 6, "H"), (777, 77, 7, 6, "both"), (40, 301, 3, 4, "both"), (3, 17, 1, 4, "both")):
        dt = np.float32 if f32 else np.float64
        X, V = synth.spectra(rows, m, k, seed=3, vmode="ivar_mask", p_missing=0.05, dtype=dt)
        W0, H0 = orc.init_factors(X, V, k, seed=4, dtype=dt)
        W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=its, tol=0, update=upd)
        Wr, Hr, hr = orc.solve(X, V, W0, H0, its, update_W=upd in ("both", "W"), update_H=upd in ("both", "H"))
        e = (rel(W, Wr), rel(H, Hr), float(np.max(np.abs(hist - hr) / hr)))
        print("parity %d x %d k=%d %s: W %.1e H %.1e chi %.1e %s" % (rows, m, k, upd, e[0], e[1], e[2], "ok" if max(e) < (1e-4 if f32 else 1e-10) else "BAD"), flush=True)
    dev = torch.device("cuda", 0)
    code, tdt, ndt = (_lib.F32, torch.float32, np.float32) if f32 else (_lib.F64, torch.float64, np.float64)
    for (rows, m, k) in ((n, 301, 6), (100000, 301, 4)) + (((1700000, 184, 8),) if f32 else ()):
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
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.solve(X, V, W, H, hist, iters)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / (iters + 1)
        h = hist.cpu().numpy()
        print("timing %d x %d k=%d: %.3f ms/iter %.0f GB/s variant=%d thr=%d st=%d smem=%d chi %.6g -> %.6g mono=%s" % (
            rows, m, k, ms, info["bytes_per_iter"] / ms / 1e6, info["variant"], info["threads"], info["stages"], info["smem_bytes"],
            h[0], h[-1], bool(np.all(np.diff(h) <= (1e-6 if f32 else 1e-9) * h[:-1]))), flush=True)
        plan.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    opt = lambda name, d: int(args[args.index(name) + 1]) if name in args else d
    n, iters = opt("--n", 1_000_000), opt("--iters", 20)
    if "--child" in args:
        child(n, iters, "--f32" in args)
        sys.exit(0)
    for cfg in args[-1].split(";"):
        env = dict(os.environ)
        for kv in cfg.split(","):
            if "=" in kv:
                a, b = kv.split("=", 1)
                env[a] = b
        print("=== " + cfg, flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--n", str(n), "--iters", str(iters)] + (["--f32"] if "--f32" in args else []),
                               env=env, capture_output=True, text=True, timeout=240)
            print(r.stdout + r.stderr[-2000:], flush=True)
        except subprocess.TimeoutExpired as exc:
            print("TIMEOUT\n" + (exc.stdout.decode() if exc.stdout else ""), flush=True)
