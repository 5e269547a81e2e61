"""cnmf_chi2_static: the entry point that allocates nothing (its scratch is a file-scope device array).

Collected last on purpose.  The GPU test calls the new entry point through
the C ABI and compares it with the existing one (cnmf_chi2 on a plan) on two calls in a row -- the second on a
smaller problem, so that stale partial sums of the first call sit in the shared scratch behind the ones it
reads."""
import ctypes as C

import numpy as np
import pytest

from cnmf_b200 import _lib

TOL = {np.float64: 1e-10, np.float32: 1e-4}    # the parity tolerances of tests/test_gpu_parity.py


def _case(n, m, k, seed, dtype):
    rng = np.random.default_rng(seed)
    X = (rng.random((n, m)) * 3).astype(dtype)
    V = ((rng.random((n, m)) * 4 + 0.05) * (rng.random((n, m)) > 0.1)).astype(dtype)
    W = (rng.random((n, k)) + 0.01).astype(dtype)
    H = (rng.random((k, m)) + 0.01).astype(dtype)
    return X, V, W, H


def test_static_entry_point_rejects_sizes_beyond_its_scratch():
    """Size checks come before any device work, so this runs without a GPU (pointers are never followed)."""
    lib = _lib.load()
    p = C.c_void_p(256)
    too_many = _lib.CHI2_STATIC_MAX_ROWS + 1
    assert lib.cnmf_chi2_static(_lib.F64, p, p, p, p, too_many, 301, 301, 6, p, None) == _lib.ERR_UNSUPPORTED
    assert b"CNMF_CHI2_STATIC_MAX_ROWS" in lib.cnmf_last_error()
    wide = _lib.CHI2_STATIC_MAX_H_BYTES // (8 * 16) + 1          # k = 16 fp64 rows of H just over the limit
    assert lib.cnmf_chi2_static(_lib.F64, p, p, p, p, 1000, wide, wide, 16, p, None) == _lib.ERR_UNSUPPORTED
    assert b"CNMF_CHI2_STATIC_MAX_H_BYTES" in lib.cnmf_last_error()
    assert lib.cnmf_chi2_static(_lib.F64, p, p, p, p, 1000, 300, 301, 6, p, None) == _lib.ERR_ARG      # ld < m
    assert lib.cnmf_chi2_static(_lib.F64, None, p, p, p, 1000, 301, 301, 6, p, None) == _lib.ERR_ARG
    assert lib.cnmf_chi2_static(7, p, p, p, p, 1000, 301, 301, 6, p, None) == _lib.ERR_ARG


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_static_entry_point_matches_plan_chi2_on_two_calls_in_a_row(dtype):
    import torch
    from cnmf_b200 import Plan, prepare_device
    from cnmf_b200._api import _stream_ptr
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    code = _lib.F64 if dtype == np.float64 else _lib.F32
    # first call: 9 row blocks with a ragged last one; second call: 2 blocks (7 stale partials stay behind them)
    for n, m, k, seed in ((8300, 301, 6, 5), (1500, 81, 3, 6)):
        X, V, W, H = _case(n, m, k, seed, dtype)
        Xp, Vp = prepare_device(torch.from_numpy(X).to(dev), torch.from_numpy(V).to(dev))
        Wd, Hd = torch.from_numpy(W).to(dev), torch.from_numpy(H).to(dev)
        ref = torch.zeros(1, dtype=torch.float64, device=dev)
        plan = Plan(n, m, k, dtype, ld=Xp.shape[1])
        plan.chi2(Xp, Vp, Wd, Hd, ref)
        torch.cuda.synchronize()
        plan.close()
        out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
        _lib.check(lib.cnmf_chi2_static(code, Xp.data_ptr(), Vp.data_ptr(), Wd.data_ptr(), Hd.data_ptr(), n,
                                        Xp.shape[1], m, k, out.data_ptr(), _stream_ptr(None)))
        torch.cuda.synchronize()
        a, b = float(out.item()), float(ref.item())
        assert b > 0 and abs(a - b) <= TOL[dtype] * b, (n, a, b)
        # the same operands without the pitch padding (ld = m): no layout requirement on this entry point
        Xd, Vd = Xp[:, :m].contiguous(), Vp[:, :m].contiguous()
        out.fill_(-1.0)
        _lib.check(lib.cnmf_chi2_static(code, Xd.data_ptr(), Vd.data_ptr(), Wd.data_ptr(), Hd.data_ptr(), n, m, m, k,
                                        out.data_ptr(), _stream_ptr(None)))
        torch.cuda.synchronize()
        assert abs(float(out.item()) - b) <= TOL[dtype] * b
