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


def _rel(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))) / np.max(np.abs(b)))


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_virtual_shards_on_one_gpu_match_oracle(dtype, world):
    """`world` ranks (8 = the BASELINE box) on ONE GPU: per-rank plans and fused kernels, all ranks' exchange +
    H update in one cooperative launch (cnmf_solve_group) -- against the oracle, the one-GPU solve and each other."""
    import cnmf
    from cnmf_b200 import solve_virtual_shards
    X, V, W0, H0 = _case(dtype=dtype)
    iters = 25
    W, Hs, hists = solve_virtual_shards(X, 6, world, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    tol = TOL[np.dtype(dtype)]
    assert W.shape == Wr.shape and all(len(h) == iters + 1 for h in hists)
    for H, hist in zip(Hs, hists):
        assert np.array_equal(H, Hs[0]) and np.array_equal(hist, hists[0])      # bit-identical on every rank
    assert _rel(W, Wr) <= tol and _rel(Hs[0], Hr) <= tol
    assert float(np.max(np.abs(hists[0] - hr) / hr)) <= tol
    W1, H1, h1 = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
    assert _rel(W, W1) <= tol and _rel(Hs[0], H1) <= tol and float(np.max(np.abs(hists[0] - h1) / h1)) <= tol


def test_virtual_shards_stop_at_the_same_iteration():
    from cnmf_b200 import solve_virtual_shards
    X, V, W0, H0 = _case(n=6007, m=81, k=3)
    W, Hs, hists = solve_virtual_shards(X, 3, 4, weights=V, W0=W0, H0=H0, max_iter=500, tol=1e-3)
    Wr, Hr, hr = orc.solve(X, V, W0, H0, 500, tol=1e-3)
    for H, hist in zip(Hs, hists):
        assert len(hist) == len(hr) and np.array_equal(hist, hists[0]) and np.array_equal(H, Hs[0])
    assert _rel(Hs[0], Hr) <= 1e-10 and _rel(W, Wr) <= 1e-10


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_devices_keyword_shards_over_the_gpus_of_the_box(dtype):
    """cnmf.nmf(X, k, devices=range(G)) from plain python: one host thread and one plan per GPU, exchange through
    peer windows (no NCCL, no torchrun), W returned as one array."""
    import cnmf
    g = _gpus()
    if g < 2:
        pytest.skip("needs >= 2 GPUs (virtual shards cover the one-GPU box)")
    X, V, W0, H0 = _case(dtype=dtype)
    iters = 25
    W, H, hist = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0, devices=range(g))
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    tol = TOL[np.dtype(dtype)]
    assert _rel(W, Wr) <= tol and _rel(H, Hr) <= tol and float(np.max(np.abs(hist - hr) / hr)) <= tol
    with pytest.raises(ValueError):
        cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=2, tol=0, devices=[0, 0])


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
    if _gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    X, V, W0, H0 = _case(n=4000, m=81, k=2)
    X = X.copy()
    X[3999, 5] = -1.0          # lands in the last rank's block
    V[3999, 5] = 1.0
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=5, tol=0, devices=range(min(_gpus(), 3)))
