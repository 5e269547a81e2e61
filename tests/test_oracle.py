"""Pins for the CPU oracle (the upstream snapshot has no tests: SURVEY.md section 4)."""
import numpy as np
import pytest

from oracle import cnmf_oracle as orc
from cnmf_b200 import synth


def _case(n=1000, m=81, k=2, seed=0, vmode="ivar", p=0.0, dtype=np.float64):
    X, V = synth.spectra(n, m, k, seed=seed, vmode=vmode, p_missing=p, dtype=dtype)
    W0, H0 = orc.init_factors(X, V, k, seed=seed + 1, dtype=dtype)
    return X, V, W0, H0


def test_T1_chi2_monotone_and_nonnegative():
    X, V, W0, H0 = _case()
    W, H, hist = orc.solve(X, V, W0, H0, 500)
    assert len(hist) == 501
    assert np.all(np.diff(hist) <= 1e-12 * hist[:-1])
    assert W.min() >= 0 and H.min() >= 0
    assert hist[-1] < 0.05 * hist[0]


def test_T2_sklearn_unweighted_cross_check():
    from sklearn.decomposition._nmf import _multiplicative_update_h, _multiplicative_update_w
    X, _, W0, H0 = _case(vmode="ones")
    V = np.ones_like(X)
    W, H = W0.copy(), H0.copy()
    # sklearn's W-then-H on X^T is our H-then-W on X (roles swap under transpose)
    Wt, Ht = H0.T.copy(), W0.T.copy()
    for _ in range(200):
        Wt, *_ = _multiplicative_update_w(X.T, Wt, Ht, beta_loss=2, l1_reg_W=0, l2_reg_W=0, gamma=1.0)
        Ht = _multiplicative_update_h(X.T, Wt, Ht, beta_loss=2, l1_reg_H=0, l2_reg_H=0, gamma=1.0)
    W, H, _ = orc.solve(X, V, W0, H0, 200)
    assert np.max(np.abs(H - Wt.T)) <= 1e-10 * np.max(np.abs(H))
    assert np.max(np.abs(W - Ht.T)) <= 1e-10 * np.max(np.abs(W))


def test_T3_exact_factorisation_is_a_fixed_point():
    rng = np.random.default_rng(5)
    Ws, Hs = rng.random((300, 3)) + 0.1, rng.random((3, 40)) + 0.1
    X = Ws @ Hs
    V = rng.random(X.shape) + 0.5
    W, H, hist = orc.solve(X, V, Ws, Hs, 3)
    assert np.max(np.abs(W - Ws)) < 1e-13 and np.max(np.abs(H - Hs)) < 1e-13
    assert hist.max() < 1e-24


def test_T4_masked_entries_do_not_matter():
    X, V, W0, H0 = _case(n=400, vmode="ivar_mask", p=0.2)
    rng = np.random.default_rng(9)
    X2 = X.copy()
    X2[V == 0] = rng.random(int((V == 0).sum())) * 100
    a = orc.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=50, tol=0)
    b = orc.nmf(X2, 2, weights=V, W0=W0, H0=H0, max_iter=50, tol=0)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
    X3 = X.copy()
    X3[V == 0] = np.nan
    c = orc.nmf(X3, 2, weights=V, W0=W0, H0=H0, max_iter=50, tol=0)
    assert np.array_equal(a[0], c[0])
    # mask= keyword is the same as zero weights
    d = orc.nmf(X3, 2, weights=np.where(V > 0, V, 1.0), mask=V > 0, W0=W0, H0=H0, max_iter=50, tol=0)
    assert np.array_equal(a[1], d[1])


def test_fully_masked_row_and_column_stay_finite():
    X, V, W0, H0 = _case(n=200)
    V[7, :] = 0
    V[:, 11] = 0
    W, H, hist = orc.nmf(X, 2, weights=V, W0=W0, H0=H0, max_iter=20, tol=0)
    assert np.all(np.isfinite(W)) and np.all(np.isfinite(H)) and np.all(np.isfinite(hist))
    assert np.all(W[7] == 0) and np.all(H[:, 11] == 0)


@pytest.mark.parametrize("g", [2, 4, 8])
def test_T6_row_sharded_schedule_equals_plain_loop(g):
    X, V, W0, H0 = _case(n=1003, m=81, k=3, vmode="ivar_mask", p=0.1)
    W, H, hist = orc.solve(X, V, W0, H0, 100)
    Ws, Hs, hs = orc.solve_sharded(X, V, W0, H0, 100, g)
    assert np.max(np.abs(W - Ws)) <= 1e-11 * np.max(np.abs(W))
    assert np.max(np.abs(H - Hs)) <= 1e-11 * np.max(np.abs(H))
    assert np.max(np.abs(hist - hs) / hist) <= 1e-11


def test_tolerance_stops_early_and_history_length():
    X, V, W0, H0 = _case()
    W, H, hist = orc.solve(X, V, W0, H0, 500, tol=1e-4)
    n = len(hist) - 1
    assert 1 <= n < 500
    assert (hist[-2] - hist[-1]) <= 1e-4 * hist[-2]
    assert np.all((hist[:-2] - hist[1:-1]) > 1e-4 * hist[:-2])


def test_float32_tracks_float64():
    X, V, W0, H0 = _case(n=2000, m=184, k=4, vmode="mask", p=0.1)
    W, H, hist = orc.solve(X, V, W0, H0, 100)
    f = np.float32
    W32, H32, h32 = orc.solve(X.astype(f), V.astype(f), W0.astype(f), H0.astype(f), 100)
    assert W32.dtype == f and h32.dtype == np.float64
    assert np.max(np.abs(W32 - W)) <= 1e-4 * np.max(np.abs(W))
    assert np.max(np.abs(H32 - H)) <= 1e-4 * np.max(np.abs(H))
    assert np.max(np.abs(h32 - hist) / hist) <= 1e-4


def test_argument_errors():
    X, V, W0, H0 = _case(n=50)
    with pytest.raises(ValueError):
        orc.nmf(-X - 1, 2)
    with pytest.raises(ValueError):
        orc.nmf(X, 2, weights=-V)
    with pytest.raises(ValueError):
        orc.nmf(X, 2, weights=V[:10])
    with pytest.raises(ValueError):
        orc.nmf(X, 2, W0=W0[:10], H0=H0)
    Xn = X.copy()
    Xn[0, 0] = np.nan
    with pytest.raises(ValueError):
        orc.nmf(Xn, 2)


def test_batched_equals_looped():
    Xs, Vs, Ws, Hs = zip(*[_case(n=64, m=81, k=3, seed=s) for s in range(4)])
    X, V, W0, H0 = map(np.stack, (Xs, Vs, Ws, Hs))
    W, H, hist = orc.nmf(X, 3, weights=V, W0=W0, H0=H0, max_iter=30, tol=0)
    for b in range(4):
        w, h, c = orc.solve(X[b], V[b], W0[b], H0[b], 30)
        assert np.array_equal(W[b], w) and np.array_equal(H[b], h) and np.array_equal(hist[b], c)
