"""GPU parity: the CUDA path (through the C ABI / cnmf.nmf) against the NumPy
oracle on the same seeded inputs.  Tolerances are BASELINE.json's: 1e-10
relative in fp64, 1e-4 in fp32; arrays are compared as max|d| / max|ref|, the
chi2 trace per entry."""
import os

import numpy as np
import pytest

from cnmf_b200 import synth
from oracle import cnmf_oracle as orc
from tests import make_golden as mg

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = {np.dtype(np.float64): 1e-10, np.dtype(np.float32): 1e-4}


def rel(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))) / np.max(np.abs(b)))


def assert_parity(got, ref, dtype, scale=1.0):
    tol = TOL[np.dtype(dtype)] * scale
    W, H, hist = got
    Wr, Hr, hr = ref
    assert len(hist) == len(hr)
    assert rel(W, Wr) <= tol, ("W", rel(W, Wr))
    assert rel(H, Hr) <= tol, ("H", rel(H, Hr))
    # chi2 entries that have fallen to rounding noise (exact fits) are compared on the scale of chi2[0]
    floor = np.maximum(hr, 1e-13 * hr[0])
    assert float(np.max(np.abs(hist - hr) / floor)) <= tol, ("chi2", float(np.max(np.abs(hist - hr) / floor)))


def case(n, m, k, seed, vmode="ivar", p=0.0, dtype=np.float64):
    X, V = synth.spectra(n, m, k, seed=seed, vmode=vmode, p_missing=p, dtype=dtype)
    W0, H0 = orc.init_factors(X, V, k, seed=seed + 1, dtype=dtype)
    return X, V, W0, H0


@pytest.mark.parametrize("route", ["0", None], ids=["streaming", "default-route"])
@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_golden_vectors(name, route, monkeypatch):
    """default-route: a problem that fits a cluster's shared memory runs on the resident cluster kernel."""
    import cnmf
    if route is not None:
        monkeypatch.setenv("CNMF_ONE_CTA", route)
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    X, V, W0, H0, k, iters = mg.inputs(name)
    got = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=iters, tol=0)
    assert got[0].dtype == X.dtype and got[2].dtype == np.float64
    assert_parity(got, (g["W"], g["H"], g["hist"]), X.dtype)


# The host entry point can send a single problem through the batched solver (CNMF_ONE_CTA=1,
# desc.variant = CNMF_VARIANT_ONE_CTA); small shapes are checked through both routes.
# "one-cta" runs the resident cluster kernel where the problem fits a cluster's shared memory
# (resident.cu; k <= 8) and the streaming batched kernel otherwise; "one-cta-streamed" forces the latter.
ROUTES = [pytest.param("0", id="streaming"), pytest.param("1", id="one-cta"), pytest.param("1s", id="one-cta-streamed")]


def set_route(monkeypatch, route):
    monkeypatch.setenv("CNMF_ONE_CTA", route[0])
    monkeypatch.setenv("CNMF_RESIDENT", "0" if route.endswith("s") else "1")


@pytest.mark.parametrize("route", ROUTES)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n,m,k", [(1, 5, 1), (3, 7, 2), (17, 4, 3), (1003, 83, 5), (257, 301, 7), (999, 50, 8),
                                   (400, 33, 11), (130, 64, 16), (2050, 184, 6)])
def test_ragged_shapes_and_every_k_family(n, m, k, dtype, route, monkeypatch):
    import cnmf
    set_route(monkeypatch, route)
    X, V, W0, H0 = case(n, m, k, seed=n + m + k, vmode="ivar_mask", p=0.1, dtype=dtype)
    got = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=30, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 30), dtype)


@pytest.mark.parametrize("route", ROUTES)
@pytest.mark.parametrize("k", range(1, 17))
def test_all_k_fp64(k, route, monkeypatch):
    import cnmf
    set_route(monkeypatch, route)
    X, V, W0, H0 = case(300, 41, k, seed=k)
    got = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=10, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 10), np.float64)


@pytest.mark.parametrize("route", ROUTES)
def test_small_problem_routes_agree_on_the_stopping_iteration(route, monkeypatch):
    """tol > 0, W-only and H-only fits through both routes of the host entry point."""
    import cnmf
    set_route(monkeypatch, route)
    X, V, W0, H0 = case(1000, 81, 2, seed=3, vmode="ones")
    got = cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=500, tol=1e-4)
    ref = orc.solve(X, V, W0, H0, 500, tol=1e-4)
    assert len(got[2]) == len(ref[2]) and len(ref[2]) < 501
    assert_parity(got, ref, np.float64)
    for upd in ("W", "H"):
        got = cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=20, tol=0, update=upd)
        assert_parity(got, orc.solve(X, V, W0, H0, 20, update_W=upd == "W", update_H=upd == "H"), np.float64)


def test_config2_reduced_iop_fp64():
    import cnmf
    X, V, W0, H0 = case(8000, 301, 4, seed=21)
    got = cnmf.nmf(X, 4, weights=V, W0=W0, H0=H0, max_iter=60, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 60), np.float64)


def test_config3_reduced_scene_fp32_mask_keyword():
    import cnmf
    f = np.float32
    X, V, W0, H0 = case(20000, 184, 8, seed=22, vmode="mask", p=0.1, dtype=f)
    Xn = X.copy()
    Xn[V == 0] = np.nan      # missing pixels may hold anything
    got = cnmf.nmf(Xn, 8, mask=V > 0, W0=W0, H0=H0, max_iter=60, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 60), f)
    # the fp32 run also tracks the fp64 oracle
    ref64 = orc.solve(X.astype(np.float64), V.astype(np.float64), W0.astype(np.float64), H0.astype(np.float64), 60)
    assert_parity(got, ref64, f)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_config5_reduced_many_tiles_per_cta(dtype):
    """Enough rows that every persistent CTA walks many tiles and (fp32) flushes."""
    import cnmf
    X, V, W0, H0 = case(60000, 301, 6, seed=23, vmode="ivar_mask", p=0.05, dtype=dtype)
    got = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=20, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 20), dtype)


def test_tolerance_stop_matches_oracle():
    import cnmf
    X, V, W0, H0 = case(1000, 81, 2, seed=31)
    ref = orc.solve(X, V, W0, H0, 500, tol=1e-4)
    got = cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=500, tol=1e-4)
    assert 1 <= len(ref[2]) - 1 < 500
    assert_parity(got, ref, np.float64)


@pytest.mark.parametrize("update", ["W", "H"])
def test_single_factor_updates(update):
    import cnmf
    X, V, W0, H0 = case(700, 81, 3, seed=41)
    got = cnmf.nmf(X, 3, weights=V, W0=W0, H0=H0, max_iter=25, tol=0, update=update)
    ref = orc.solve(X, V, W0, H0, 25, update_W=update == "W", update_H=update == "H")
    assert_parity(got, ref, np.float64)
    fixed, start = (got[1], H0) if update == "W" else (got[0], W0)
    assert np.array_equal(fixed, start)


def test_masked_row_and_column_stay_finite_and_zero():
    import cnmf
    X, V, W0, H0 = case(300, 81, 2, seed=51)
    V[7, :] = 0
    V[:, 11] = 0
    got = cnmf.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=20, tol=0)
    assert np.all(got[0][7] == 0) and np.all(got[1][:, 11] == 0)
    assert_parity(got, orc.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=20, tol=0), np.float64)


def test_exact_factorisation_is_a_fixed_point():
    import cnmf
    rng = np.random.default_rng(5)
    Ws, Hs = rng.random((300, 3)) + 0.1, rng.random((3, 40)) + 0.1
    X = Ws @ Hs
    V = rng.random(X.shape) + 0.5
    W, H, hist = cnmf.nmf(X, 3, weights=V, W0=Ws, H0=Hs, max_iter=3, tol=0)
    assert np.max(np.abs(W - Ws)) < 1e-12 and np.max(np.abs(H - Hs)) < 1e-12
    assert hist.max() < 1e-22


def test_torch_cuda_input_stays_on_device_and_is_deterministic():
    import torch
    import cnmf
    X, V, W0, H0 = case(5000, 184, 4, seed=61, vmode="ivar_mask", p=0.1)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(a).to(dev)
    out1 = cnmf.nmf(t(X), 4, weights=t(V), W0=t(W0), H0=t(H0), max_iter=40, tol=0)
    out2 = cnmf.nmf(t(X), 4, weights=t(V), W0=t(W0), H0=t(H0), max_iter=40, tol=0)
    assert all(o.is_cuda for o in out1)
    for a, b in zip(out1, out2):
        assert torch.equal(a, b)              # fixed-order reductions: bit-reproducible
    host = cnmf.nmf(X, 4, weights=V, W0=W0, H0=H0, max_iter=40, tol=0)
    for a, b in zip(out1, host):
        assert np.array_equal(a.cpu().numpy(), b)
    assert_parity(tuple(o.cpu().numpy() for o in out1), orc.solve(X, V, W0, H0, 40), np.float64)


def test_grid_size_does_not_change_the_answer_beyond_rounding():
    """Virtual shards: the same rows reduced through 1, 7 and all SMs' partial blocks."""
    import torch
    from cnmf_b200 import Plan, prepare_device
    X, V, W0, H0 = case(9000, 81, 3, seed=71)
    dev = torch.device("cuda", 0)
    Xp, Vp = prepare_device(torch.from_numpy(X).to(dev), torch.from_numpy(V).to(dev))
    ref = orc.solve(X, V, W0, H0, 15)
    for lim in (1, 7, 0):
        plan = Plan(9000, 81, 3, np.float64, ld=Xp.shape[1], sm_limit=lim)
        W, H = torch.from_numpy(W0).to(dev), torch.from_numpy(H0).to(dev)
        hist = torch.zeros(16, dtype=torch.float64, device=dev)
        assert plan.solve(Xp, Vp, W, H, hist, 15) == 15
        torch.cuda.synchronize()
        assert_parity((W.cpu().numpy(), H.cpu().numpy(), hist.cpu().numpy()), ref, np.float64)
        plan.close()


def test_chi2_entry_point():
    import torch
    from cnmf_b200 import Plan, prepare_device
    X, V, W0, H0 = case(2500, 301, 6, seed=81, dtype=np.float32)
    dev = torch.device("cuda", 0)
    Xp, Vp = prepare_device(torch.from_numpy(X).to(dev), torch.from_numpy(V).to(dev))
    plan = Plan(2500, 301, 6, np.float32, ld=Xp.shape[1])
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    plan.chi2(Xp, Vp, torch.from_numpy(W0).to(dev), torch.from_numpy(H0).to(dev), out)
    ref = orc.chi2(X.astype(np.float64), V.astype(np.float64), W0.astype(np.float64), H0.astype(np.float64))
    assert abs(float(out.item()) - ref) <= 1e-5 * ref


def test_data_errors_raise_value_error():
    import cnmf
    X, V, W0, H0 = case(100, 20, 2, seed=91)
    with pytest.raises(ValueError):
        cnmf.nmf(-X - 1.0, 2)
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, weights=-V)
    Xn = X.copy()
    Xn[3, 3] = np.inf
    with pytest.raises(ValueError):
        cnmf.nmf(Xn, 2, weights=np.ones_like(V))
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, W0=-W0, H0=H0)


def test_default_start_is_the_oracles():
    import cnmf
    X, V, _, _ = case(500, 81, 2, seed=95)
    got = cnmf.nmf(X, 2, weights=V, max_iter=30, tol=0, seed=7)
    ref = orc.nmf(X, 2, weights=V, max_iter=30, tol=0, seed=7)
    assert_parity(got, ref, np.float64)


def test_device_generator_matches_host_generator():
    import ctypes as C
    import torch
    from cnmf_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    n, m, k, seed, row0 = 777, 301, 6, 5, 123456789
    Hs = synth.basis(k, m, seed)
    for dtype, code, tdt, tol in ((np.float64, _lib.F64, torch.float64, 1e-9), (np.float32, _lib.F32, torch.float32, 1e-5)):
        ld = _lib.padded_ld(m, code)
        Xd = torch.empty((n, ld), dtype=tdt, device=dev)
        Vd = torch.empty_like(Xd)
        hs = torch.from_numpy(Hs).to(dev)
        _lib.check(lib.cnmf_synth(code, Xd.data_ptr(), Vd.data_ptr(), n, row0, m, ld, k, hs.data_ptr(), seed,
                                  synth.VMODES["ivar_mask"], 0.05, None))
        torch.cuda.synchronize()
        X, V = synth.spectra(n, m, k, seed=seed, vmode="ivar_mask", p_missing=0.05, dtype=dtype, row0=row0)
        Xg, Vg = Xd.cpu().numpy(), Vd.cpu().numpy()
        assert np.all(Xg[:, m:] == 0) and np.all(Vg[:, m:] == 0)
        assert np.array_equal(Vg[:, :m] == 0, V == 0)
        assert np.max(np.abs(Xg[:, :m] - X)) <= tol * np.max(X)
        assert np.max(np.abs(Vg[:, :m] - V) / np.maximum(V, 1)) <= tol


@pytest.mark.parametrize("resident", ["1", "0"], ids=["resident", "streamed"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_batched_equals_looped(dtype, resident, monkeypatch):
    """T7: the one-problem-per-cluster / per-CTA kernels against per-problem oracle solves (and the single-problem path)."""
    import cnmf
    monkeypatch.setenv("CNMF_RESIDENT", resident)
    B, n, m, k = 6, 512, 81, 3
    cs = [case(n, m, k, seed=100 + b, dtype=dtype) for b in range(B)]
    X, V, W0, H0 = (np.stack([c[i] for c in cs]) for i in range(4))
    W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=40, tol=0)
    assert W.shape == (B, n, k) and H.shape == (B, k, m) and hist.shape == (B, 41)
    for b in range(B):
        ref = orc.solve(X[b], V[b], W0[b], H0[b], 40)
        assert_parity((W[b], H[b], hist[b]), ref, dtype)
    single = cnmf.nmf(X[2], k, weights=V[2], W0=W0[2], H0=H0[2], max_iter=40, tol=0)
    assert_parity((W[2], H[2], hist[2]), single, dtype)


@pytest.mark.parametrize("resident", ["1", "0"], ids=["resident", "streamed"])
def test_batched_ragged_and_tolerance(resident, monkeypatch):
    import cnmf
    monkeypatch.setenv("CNMF_RESIDENT", resident)
    B, n, m, k = 5, 130, 33, 4
    cs = [case(n, m, k, seed=200 + b, vmode="ivar_mask", p=0.1) for b in range(B)]
    X, V, W0, H0 = (np.stack([c[i] for c in cs]) for i in range(4))
    W, H, hist = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=300, tol=1e-4)
    for b in range(B):
        Wr, Hr, hr = orc.solve(X[b], V[b], W0[b], H0[b], 300, tol=1e-4)
        nb = len(hr)
        assert np.all(np.isnan(hist[b, nb:])) and not np.any(np.isnan(hist[b, :nb]))
        assert_parity((W[b], H[b], hist[b, :nb]), (Wr, Hr, hr), np.float64)


@pytest.mark.parametrize("n,m,k,dtype", [(1003, 83, 3, np.float64), (999, 50, 8, np.float32), (1000, 81, 2, np.float64),
                                        (257, 301, 4, np.float64), (2050, 40, 3, np.float32)])
def test_resident_cluster_kernel_repeats_bit_for_bit(n, m, k, dtype, monkeypatch):
    """The resident cluster kernel sums in a fixed order (streams, then ranks): two runs give the same bits,
    every CTA of the cluster applies the same H update, and a ragged last rank changes nothing."""
    import cnmf
    from cnmf_b200 import Plan, _lib
    monkeypatch.setenv("CNMF_RESIDENT", "1")
    p = Plan(n, m, k, dtype, variant=_lib.VARIANT_RESIDENT)
    info = p.info
    p.close()
    assert info["variant"] >= 700000 and info["grid"] == info["variant"] - 700000  # one cluster, CS CTAs
    X, V, W0, H0 = case(n, m, k, seed=n + k, vmode="ivar_mask", p=0.07, dtype=dtype)
    a = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=25, tol=0)
    b = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=25, tol=0)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
    assert_parity(a, orc.solve(X, V, W0, H0, 25), dtype)


def test_restarts_pick_the_lowest_chi2_and_match_single_runs():
    import cnmf
    from cnmf_b200 import init_factors
    X, V, _, _ = case(300, 81, 3, seed=5, vmode="ivar_mask", p=0.05)
    W, H, hist, finals = cnmf.nmf_restarts(X, 3, n_restarts=4, weights=V, max_iter=60, tol=0, seed=9)
    assert finals.shape == (4,) and hist[-1] == finals.min() and len(hist) == 61
    b = int(np.argmin(finals))
    mean = float(X[V > 0].sum()) / float((V > 0).sum())
    W0, H0 = init_factors(300, 81, 3, mean, seed=[9, b])
    assert_parity((W, H, hist), orc.solve(X, V, W0, H0, 60), np.float64)


def test_class_shim_matches_the_function():
    import cnmf
    X, V, W0, H0 = case(500, 50, 4, seed=21, vmode="ivar")
    m = cnmf.NMF(X, W=W0.copy(), H=H0.copy(), V=V)
    chi, sec = m.SolveNMF(maxiters=25, tol=0)
    ref = orc.solve(X, V, W0, H0, 25)
    assert_parity((m.W, m.H, m.chi2_history), ref, np.float64)
    assert chi == m.chi2_history[-1] == m.cost and sec > 0
    m2 = cnmf.NMF(X, W=W0.copy(), H=ref[1].copy(), V=V)
    m2.SolveNMF(maxiters=5, tol=0, W_only=True)
    assert np.array_equal(m2.H, ref[1])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_config5_full_size_properties(dtype):
    """BASELINE.json configs[4] at its full 10,000,000 x 301, k = 6 (device-generated operands):
    the oracle cannot run there, so the size-independent properties are checked instead --
    chi2 is a sum over rows (whole = sum of three uneven row blocks, on the same W and H: the
    'checksum of checksums'), chi2 falls monotonically, W and H stay non-negative and finite,
    and a 5,000-row slice of the result is one oracle step away from itself."""
    import torch
    from cnmf_b200 import Plan, _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    n, m, k, seed = 10_000_000, 301, 6, 7
    code = _lib.F32 if dtype == np.float32 else _lib.F64
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    ld = _lib.padded_ld(m, code)
    need = 2 * n * ld * np.dtype(dtype).itemsize + (4 << 30)
    if torch.cuda.mem_get_info(0)[0] < need:
        pytest.skip("needs %.0f GB of free device memory" % (need / 2**30))
    X = torch.empty((n, ld), dtype=tdt, device=dev)
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, seed)).to(dev)
    _lib.check(lib.cnmf_synth(code, X.data_ptr(), V.data_ptr(), n, 0, m, ld, k, hs.data_ptr(), seed,
                              synth.VMODES["ivar_mask"], 0.05, None))
    g = torch.Generator(device=dev).manual_seed(11)
    W = torch.rand((n, k), dtype=tdt, device=dev, generator=g)
    H = torch.rand((k, m), dtype=tdt, device=dev, generator=g)
    iters = 6
    hist = torch.full((iters + 1,), float("nan"), dtype=torch.float64, device=dev)
    plan = Plan(n, m, k, dtype, ld=ld)
    plan.solve(X, V, W, H, hist, iters - 1)
    # one more iteration, keeping the state before it for the slice check below
    sl = slice(4_999_000, 5_004_000)
    Wb = W[sl].clone()
    plan.solve(X, V, W, H, hist, 1, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME)
    h = hist.cpu().numpy()
    assert np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]), h
    assert float(W.min()) >= 0 and float(H.min()) >= 0
    assert bool(torch.isfinite(W).all()) and bool(torch.isfinite(H).all())
    # chi2 over the whole set = sum over row blocks
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    plan.chi2(X, V, W, H, out)
    whole = float(out.item())
    plan.close()
    parts = 0.0
    for lo, hi in ((0, 3_333_332), (3_333_332, 3_333_342), (3_333_342, n)):   # even starts keep fp32 W rows 16-byte aligned
        p = Plan(hi - lo, m, k, dtype, ld=ld)
        p.chi2(X[lo:hi], V[lo:hi], W[lo:hi], H, out)
        parts += float(out.item())
        p.close()
    assert abs(whole - parts) <= (1e-12 if dtype == np.float64 else 1e-6) * whole
    assert abs(whole - h[-1]) <= (1e-10 if dtype == np.float64 else 1e-4) * whole   # history's last entry is chi2(W, H)
    # the W update of the last iteration (it uses the H that iteration produced) on a slice of rows,
    # against the oracle's formula
    Xs, Vs = X[sl, :m].cpu().numpy().astype(np.float64), V[sl, :m].cpu().numpy().astype(np.float64)
    Wo, Ho = Wb.cpu().numpy().astype(np.float64), H.cpu().numpy().astype(np.float64)
    eps = 1e-12 if dtype == np.float32 else 1e-16
    Wn = Wo * ((Vs * Xs) @ Ho.T) / np.maximum((Vs * (Wo @ Ho)) @ Ho.T, eps)
    assert rel(W[sl].cpu().numpy(), Wn) <= TOL[np.dtype(dtype)]


# ---------------------------------------------------------------------------
# Config 4 at its full size: 4 096 problems of 512 x 81, k = 3 in one launch
# (28 waves of CTAs, tiles re-read through L2, the K4 / K4r routing rule).
# The generator's global row index makes problem b the rows [512 b, 512 b + 512).
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("resident", ["0", "1"], ids=["K4-streamed", "K4r-resident"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_config4_full_batch_against_looped_solves(dtype, resident, monkeypatch):
    import torch
    import cnmf
    from cnmf_b200 import _lib
    monkeypatch.setenv("CNMF_RESIDENT", resident)
    B, n, m, k, iters = 4096, 512, 81, 3, 20
    dev = torch.device("cuda", 0)
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    code = _lib.F32 if dtype == np.float32 else _lib.F64
    lib = _lib.load()
    X = torch.empty((B * n, m), dtype=tdt, device=dev)   # dense pitch: cnmf.nmf pads on the device
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, 44)).to(dev)
    _lib.check(lib.cnmf_synth(code, X.data_ptr(), V.data_ptr(), B * n, 0, m, m, k, hs.data_ptr(), 44,
                              synth.VMODES["ivar"], 0.0, None))
    rng = np.random.default_rng(45)
    W0 = torch.from_numpy(rng.random((B, n, k)).astype(dtype)).to(dev)
    H0 = torch.from_numpy(rng.random((B, k, m)).astype(dtype)).to(dev)
    W, H, hist = cnmf.nmf(X.view(B, n, m), k, weights=V.view(B, n, m), W0=W0, H0=H0, max_iter=iters, tol=0)
    assert W.shape == (B, n, k) and hist.shape == (B, iters + 1)
    h = hist.cpu().numpy()
    assert np.all(np.isfinite(h)) and np.all(np.diff(h, axis=1) <= 1e-5 * h[:, :-1])
    pick = np.unique(np.concatenate([[0, 1, 147, 148, 2047, 4094, 4095], np.random.default_rng(3).integers(0, B, 12)]))
    Xh, Vh = X.view(B, n, m)[pick].cpu().numpy(), V.view(B, n, m)[pick].cpu().numpy()
    for i, b in enumerate(pick):
        ref = orc.solve(Xh[i], Vh[i], W0[b].cpu().numpy(), H0[b].cpu().numpy(), iters)
        assert_parity((W[b].cpu().numpy(), H[b].cpu().numpy(), h[b]), ref, dtype)
        # ... and the batch member equals the same problem solved alone through the entry point
        single = cnmf.nmf(Xh[i], k, weights=Vh[i], W0=W0[b].cpu().numpy(), H0=H0[b].cpu().numpy(), max_iter=iters, tol=0)
        assert_parity((W[b].cpu().numpy(), H[b].cpu().numpy(), h[b]), single, dtype)


def test_large_pageable_input_goes_through_the_staging_ring():
    """Pageable NumPy arrays above 32 MB are uploaded through the pinned staging ring (host.cu, Stager): X, weights and
    mask each take that path here; a pinned copy of the same arrays (the direct path) must give identical bits."""
    import torch
    import cnmf
    n, m, k = 70_000, 301, 4
    X, V, W0, H0 = case(n, m, k, seed=31, vmode="ivar", p=0.0)
    mask = np.random.default_rng(5).random((n, m)) > 0.07
    a = cnmf.nmf(X, k, weights=V, mask=mask, W0=W0, H0=H0, max_iter=3, tol=0)
    ref = orc.nmf(X, k, weights=V, mask=mask, W0=W0, H0=H0, max_iter=3, tol=0)
    assert_parity(a, ref, np.float64)
    pin = lambda arr: torch.from_numpy(arr).pin_memory().numpy()
    b = cnmf.nmf(pin(X), k, weights=pin(V), mask=pin(mask.astype(np.uint8)), W0=W0, H0=H0, max_iter=3, tol=0)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
