"""The row-sharded schedule and the fused exchange + H-update kernel inside ONE process
(cnmf_comm_create_local): ranks are host threads, the peers' windows are reached through peer access.
On a one-GPU box every rank lives on device 0 ("virtual shards"), so the driver's
single-GPU test run exercises the exchange arithmetic; with more GPUs the ranks spread over them."""
import numpy as np
import pytest

from cnmf_b200 import synth
from oracle import cnmf_oracle as orc

pytestmark = pytest.mark.gpu
TOL = {np.dtype(np.float64): 1e-10, np.dtype(np.float32): 1e-4}


def _case(n=20011, m=301, k=6, dtype=np.float64):
    X, V = synth.spectra(n, m, k, seed=77, vmode="ivar_mask", p_missing=0.1, dtype=dtype)
    W0, H0 = orc.init_factors(X, V, k, seed=78, dtype=dtype)
    return X, V, W0, H0


def _devices(world):
    import torch
    nd = torch.cuda.device_count()
    return [r % nd for r in range(world)]


def _rel(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))) / np.max(np.abs(b)))


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_sharded_solve_through_peer_windows_matches_oracle(dtype, world):
    """`world` ranks (8 = the BASELINE box) against the single-process oracle and against each other."""
    import cnmf
    X, V, W0, H0 = _case(dtype=dtype)
    iters = 25
    W, H, hist = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0, devices=_devices(world))
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    tol = TOL[np.dtype(dtype)]
    assert W.shape == Wr.shape and len(hist) == iters + 1
    assert _rel(W, Wr) <= tol and _rel(H, Hr) <= tol
    assert float(np.max(np.abs(hist - hr) / hr)) <= tol
    # the sharded result against the one-GPU result of the same library (summation order only)
    W1, H1, h1 = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
    assert _rel(W, W1) <= tol and _rel(H, H1) <= tol and float(np.max(np.abs(hist - h1) / h1)) <= tol


def test_every_rank_holds_the_same_H_and_stops_at_the_same_iteration():
    import ctypes as C
    from cnmf_b200 import LocalGroup, _lib, shard_rows
    lib = _lib.load()
    X, V, W0, H0 = _case(n=6007, m=81, k=3)
    world = 4
    grp = LocalGroup(_devices(world))
    vp = lambda a: a.ctypes.data_as(C.c_void_p)

    def body(c):
        lo, hi = shard_rows(X.shape[0], world, c.rank)
        Xr, Vr = np.ascontiguousarray(X[lo:hi]), np.ascontiguousarray(V[lo:hi])
        W, H, hist = W0[lo:hi].copy(), H0.copy(), np.full(501, np.nan)
        n_iter = C.c_int32(0)
        rc = lib.cnmf_nmf_host(_lib.F64, hi - lo, 81, 3, vp(Xr), vp(Vr), None, vp(W), vp(H), vp(hist), 500, 1e-3, 0.0,
                               _lib.UPDATE_BOTH, c.handle, C.byref(n_iter), None, None)
        return rc, H, hist[:n_iter.value + 1]
    try:
        res = grp.run(body)
    finally:
        grp.close()
    Wr, Hr, hr = orc.solve(X, V, W0, H0, 500, tol=1e-3)
    for rc, H, hist in res:
        assert rc == 0
        assert np.array_equal(H, res[0][1]) and np.array_equal(hist, res[0][2])   # bit-identical: no broadcast needed
        assert len(hist) == len(hr) and _rel(H, Hr) <= 1e-10


def test_a_missing_peer_ends_a_fixed_count_solve_with_an_error():
    """Rank 1 never runs: rank 0's exchange kernel gives up after the (shortened) timeout and the
    tol = 0 solve returns an error instead of rc = 0 with an unwritten history -- through cnmf_solve
    (plan level) and through the host entry point (whose first exchange is the input verdict)."""
    import ctypes as C
    import torch
    from cnmf_b200 import LocalGroup, Plan, _lib, prepare_device
    lib = _lib.load()
    X, V, W0, H0 = _case(n=1000, m=81, k=2)
    grp = LocalGroup([0, 0], timeout=0.5)

    def plan_level(c):
        dev = torch.device("cuda", 0)
        Xp, Vp = prepare_device(torch.from_numpy(X).to(dev), torch.from_numpy(V).to(dev))
        W, H = torch.from_numpy(W0).to(dev), torch.from_numpy(H0).to(dev)
        hist = torch.full((11,), float("nan"), dtype=torch.float64, device=dev)
        plan = Plan(1000, 81, 2, np.float64, ld=Xp.shape[-1], comm=c)
        try:
            plan.solve(Xp, Vp, W, H, hist, 10, tol=0.0)
        except _lib.CnmfError as exc:
            return exc.code, str(exc), hist.cpu().numpy()
        finally:
            plan.close()
        return 0, "", hist.cpu().numpy()
    try:
        rc, text, hist = grp.run(plan_level, only=[0])[0]
    finally:
        grp.close()
    assert rc == _lib.ERR_NCCL and "timed out" in text
    assert np.all(np.isnan(hist))

    grp = LocalGroup([0, 0], timeout=0.5)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)

    def host_level(c):
        W, H, hist = W0.copy(), H0.copy(), np.full(11, np.nan)
        n_iter = C.c_int32(-1)
        rc = lib.cnmf_nmf_host(_lib.F64, 1000, 81, 2, vp(X), vp(V), None, vp(W), vp(H), vp(hist), 10, 0.0, 0.0,
                               _lib.UPDATE_BOTH, c.handle, C.byref(n_iter), None, None)
        return rc, lib.cnmf_last_error().decode(), hist
    try:
        rc, text, hist = grp.run(host_level, only=[0])[0]
    finally:
        grp.close()
    assert rc == _lib.ERR_NCCL and "timed out" in text and np.all(np.isnan(hist))


def test_one_bad_shard_fails_every_rank_together():
    import cnmf
    X, V, W0, H0 = _case(n=4000, m=81, k=2)
    X = X.copy()
    X[3999, 5] = -1.0          # lands in the last rank's block
    V[3999, 5] = 1.0
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=5, tol=0, devices=_devices(3))
