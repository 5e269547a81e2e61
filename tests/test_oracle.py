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


# ---------------------------------------------------------------------------
# Independent pins of the WEIGHTED update (V != 1).  The reference offers none
# (/root/reference/cnmf/__init__.py is empty), so the numerators / denominators
# the oracle forms are checked against three things that do not reuse its algebra:
# reverse-mode differentiation of chi2, a scalar triple loop, and scikit-learn's
# unweighted update through the rank-one-V identity.
# ---------------------------------------------------------------------------
def _weighted_case(n=57, m=23, k=3, seed=3, masked=True):
    rng = np.random.default_rng(seed)
    X = rng.random((n, m)) * 3
    V = rng.random((n, m)) * 4 + 0.05          # non-uniform weights
    if masked:
        V = V * (rng.random((n, m)) > 0.2)     # and a 0/1 mask folded in
    W = rng.random((n, k)) + 0.01
    H = rng.random((k, m)) + 0.01
    return X, V, W, H


@pytest.mark.parametrize("masked", [False, True])
def test_P1_numerators_and_denominators_are_the_gradient_split_of_chi2(masked):
    """d chi2 / dW = 2 (den_W - num_W) and d chi2 / dH = 2 (den_H - num_H): autograd of
    sum V (X - WH)^2 pins both pairs without touching the oracle's formulas."""
    import torch
    X, V, W, H = _weighted_case(masked=masked)
    Wt = torch.tensor(W, dtype=torch.float64, requires_grad=True)
    Ht = torch.tensor(H, dtype=torch.float64, requires_grad=True)
    chi = (torch.tensor(V) * (torch.tensor(X) - Wt @ Ht) ** 2).sum()
    chi.backward()
    assert abs(float(chi.detach()) - orc.chi2(X, V, W, H)) <= 1e-12 * float(chi.detach())
    num_h, den_h = orc.h_partials(X, V, W, H)
    gh = 2.0 * (den_h - num_h)
    assert np.max(np.abs(gh - Ht.grad.numpy())) <= 1e-12 * np.max(np.abs(gh))
    # the W pair is recovered from the update itself: W' = W num / den with num, den > 0
    VX, VWH = V * X, V * (W @ H)
    num_w, den_w = VX @ H.T, VWH @ H.T
    gw = 2.0 * (den_w - num_w)
    assert np.max(np.abs(gw - Wt.grad.numpy())) <= 1e-12 * np.max(np.abs(gw))
    Wn = orc.w_update(X, V, W, H, 1e-16)
    assert np.max(np.abs(Wn * den_w - W * num_w)) <= 1e-13 * np.max(np.abs(W * num_w))
    # autograd alone cannot tell num from den (only their difference): the split is pinned by
    # positivity -- num collects the X terms, den the WH terms -- checked by zeroing X
    n0, d0 = orc.h_partials(np.zeros_like(X), V, W, H)
    assert np.all(n0 == 0) and np.max(np.abs(d0 - den_h)) <= 1e-13 * np.max(den_h)


def test_P2_brute_force_triple_loop():
    """One H-then-W iteration on a 7 x 5, k = 2 case with plain Python loops."""
    X, V, W, H = _weighted_case(n=7, m=5, k=2, seed=11)
    n, m, k = 7, 5, 2
    eps = 1e-16

    def wh(Wc, Hc, i, j):
        s = 0.0
        for a in range(k):
            s += Wc[i][a] * Hc[a][j]
        return s
    Wl, Hl = W.tolist(), H.tolist()
    Hn = [[0.0] * m for _ in range(k)]
    for a in range(k):
        for j in range(m):
            num = den = 0.0
            for i in range(n):
                num += Wl[i][a] * V[i, j] * X[i, j]
                den += Wl[i][a] * V[i, j] * wh(Wl, Hl, i, j)
            Hn[a][j] = Hl[a][j] * num / max(den, eps)
    Wn = [[0.0] * k for _ in range(n)]
    for i in range(n):
        for a in range(k):
            num = den = 0.0
            for j in range(m):
                num += V[i, j] * X[i, j] * Hn[a][j]
                den += V[i, j] * wh(Wl, Hn, i, j) * Hn[a][j]
            Wn[i][a] = Wl[i][a] * num / max(den, eps)
    chi = 0.0
    for i in range(n):
        for j in range(m):
            chi += V[i, j] * (X[i, j] - wh(Wn, Hn, i, j)) ** 2
    Wo, Ho, co = orc.iteration(X, V, W, H, eps)
    assert np.max(np.abs(Wo - np.array(Wn))) <= 1e-13 * np.max(Wo)
    assert np.max(np.abs(Ho - np.array(Hn))) <= 1e-13 * np.max(Ho)
    assert abs(co - chi) <= 1e-13 * chi


def test_P3_rank_one_weights_reduce_to_sklearn_unweighted():
    """V = r c^T: the weighted problem on (X, W, H) is the unweighted one on
    (diag(sqrt r) X diag(sqrt c), diag(sqrt r) W, H diag(sqrt c)), so scikit-learn's MU solver
    (sklearn/decomposition/_nmf.py:521-723) is a library-backed pin of the V != 1 arithmetic."""
    from sklearn.decomposition._nmf import _multiplicative_update_h, _multiplicative_update_w
    rng = np.random.default_rng(21)
    n, m, k = 180, 41, 3
    X = rng.random((n, m)) * 2
    r, c = rng.random(n) * 3 + 0.2, rng.random(m) * 5 + 0.1
    V = np.outer(r, c)
    W0, H0 = rng.random((n, k)) + 0.05, rng.random((k, m)) + 0.05
    sr, sc = np.sqrt(r), np.sqrt(c)
    Xs = sr[:, None] * X * sc[None, :]
    # sklearn's W-then-H on Xs^T is our H-then-W on Xs (roles swap under transpose)
    Wt, Ht = (H0 * sc[None, :]).T.copy(), (sr[:, None] * W0).T.copy()
    iters = 60
    for _ in range(iters):
        Wt, *_ = _multiplicative_update_w(Xs.T, Wt, Ht, beta_loss=2, l1_reg_W=0, l2_reg_W=0, gamma=1.0)
        Ht = _multiplicative_update_h(Xs.T, Wt, Ht, beta_loss=2, l1_reg_H=0, l2_reg_H=0, gamma=1.0)
    W, H, hist = orc.solve(X, V, W0, H0, iters)
    Hs, Ws = Wt.T / sc[None, :], Ht.T / sr[:, None]
    assert np.max(np.abs(H - Hs)) <= 1e-10 * np.max(np.abs(H))
    assert np.max(np.abs(W - Ws)) <= 1e-10 * np.max(np.abs(W))
    chi_s = float(np.sum((Xs - Ht.T @ Wt.T) ** 2))
    assert abs(hist[-1] - chi_s) <= 1e-10 * chi_s
