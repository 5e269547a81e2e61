"""GPU A/B of the resident cluster kernel (resident.cu) against the streaming batched kernel:
config 4 (4096 x 512 x 81, k = 3) in fp32 / fp64 and config 1 (one 1 000 x 81 problem, k = 2) through
the one-problem-per-cluster route.  CNMF_RESIDENT=0 selects the streaming kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402

import bench_configs as bc  # noqa: E402
from cnmf_b200 import Plan, _lib  # noqa: E402


def one(n, m, k, dtype, iters):
    """single problem through a batch-of-one plan (what CNMF_ONE_CTA=1 does)"""
    X, V, W, H, ld = bc.operands(n, m, k, dtype, "ones", 0.0)
    hist = torch.zeros((1, iters + 1), dtype=torch.float64, device=bc.DEV)
    niter = torch.zeros(1, dtype=torch.int32, device=bc.DEV)
    d = _lib.Desc()
    plan = Plan.__new__(Plan)
    import ctypes as C
    d.n_rows, d.ld, d.m, d.k, d.dtype, d.batch = n, ld, m, k, (_lib.F32 if dtype == np.float32 else _lib.F64), 1
    d.eps, d.comm, d.sm_limit, d.variant = 0.0, None, 0, 1
    h = C.c_void_p()
    _lib.check(bc.LIB.cnmf_plan_create(C.byref(h), C.byref(d)))
    info = _lib.Info()
    bc.LIB.cnmf_plan_info(h, C.byref(info))

    def run(W_, H_, it):
        _lib.check(bc.LIB.cnmf_solve_batched(h, X.data_ptr(), V.data_ptr(), W_.data_ptr(), H_.data_ptr(), hist.data_ptr(),
                                             niter.data_ptr(), it, 0.0, 3, None))
    run(W.clone(), H.clone(), 3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(W, H, iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    hh = hist.cpu().numpy()[0]
    print("one problem %d x %d k=%d %s RESIDENT=%s: %d iterations %.3f ms = %.2f us/iteration  variant=%d grid=%d smem=%d  chi %.5g -> %.5g mono=%s" % (
        n, m, k, np.dtype(dtype).name, os.environ.get("CNMF_RESIDENT", "1"), iters, ms, 1e3 * ms / iters, info.variant,
        info.grid, info.smem_bytes, hh[0], hh[-1], bool(np.all(np.diff(hh) <= 1e-9 * hh[:-1]))), flush=True)
    bc.LIB.cnmf_plan_destroy(h)


if __name__ == "__main__":
    for res in ("1", "0"):
        os.environ["CNMF_RESIDENT"] = res
        print("CNMF_RESIDENT=" + res, flush=True)
        bc.batched("4: 4096 x (512 x 81) k=3 f32 batched", 4096, 512, 81, 3, np.float32, 40)
        bc.batched("4: 4096 x (512 x 81) k=3 f64 batched", 4096, 512, 81, 3, np.float64, 40)
        bc.batched("4096 x (128 x 81) k=3 f32 batched", 4096, 128, 81, 3, np.float32, 40)
        one(1000, 81, 2, np.float64, 500)
        one(1000, 81, 2, np.float32, 500)
        one(5000, 81, 2, np.float32, 200)
