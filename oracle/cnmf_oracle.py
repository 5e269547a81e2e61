"""CPU oracle for the weighted / masked multiplicative-update NMF iteration.

THIS IS TEST INFRASTRUCTURE, NOT THE PRODUCT.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  The product path (``cnmf_b200``) never does.

PARITY UNPINNED.  The upstream snapshot at /root/reference holds no NMF code
(`/root/reference/cnmf/__init__.py` is 0 bytes, `/root/reference/setup.py:1-41`
is packaging only), so no reference function, test or
golden vector pins this arithmetic.  The oracle follows the update equations
stated in BASELINE.json `north_star` (restated in SURVEY.md section 7.1):

    chi2 = sum_ij V_ij (X_ij - (WH)_ij)^2
    H <- H * [W^T (V*X)] / max(W^T (V*(WH)), eps)
    W <- W * [(V*X) H^T] / max((V*(WH)) H^T, eps)

with every free choice fixed here, once, and mirrored by the CUDA path:

* order inside one iteration: H first, then W (with the new H), then chi2;
* guard: ``new = old * num / max(den, eps)``; eps = 1e-16 (fp64), 1e-12 (fp32);
  W and H themselves are not floored;
* chi2 history is the raw sum (no normalisation), always float64, entry 0 is
  chi2 at (W0, H0), so its length is ``n_iter + 1``;
* stop when ``it == max_iter`` or, if ``tol > 0``,
  ``(hist[it-1] - hist[it]) <= tol * hist[it-1]``;
* ``weights`` and ``mask`` (True/1 = good) multiply into one V; X is zero-filled
  where V == 0 so that non-finite values under the mask cannot poison sums.

Independent pins used by tests/test_oracle.py: scikit-learn's MU solver at
V == 1 (sklearn/decomposition/_nmf.py:726-880), the exact fixed point
X = W*H*, the monotone chi2 property and mask invariance.
"""
from __future__ import annotations

import numpy as np

EPS = {np.dtype(np.float64): 1e-16, np.dtype(np.float32): 1e-12}


def default_eps(dtype) -> float:
    return EPS[np.dtype(dtype)]


def resolve_dtype(X, dtype=None):
    if dtype is not None:
        dt = np.dtype(dtype)
    else:
        dt = np.asarray(X).dtype
        if dt not in (np.dtype(np.float32), np.dtype(np.float64)):
            dt = np.dtype(np.float64)
    if dt not in EPS:
        raise ValueError("dtype must be float32 or float64")
    return dt


def prepare(X, weights=None, mask=None, dtype=None):
    """Fold weights and mask into one V and zero-fill X under V == 0.

    Returns (X, V) in the compute dtype.  Raises ValueError on the argument
    errors the entry point documents.
    """
    dt = resolve_dtype(X, dtype)
    X = np.array(X, dtype=dt, copy=True)
    if X.ndim not in (2, 3):
        raise ValueError("X must be (N, M) or (B, N, M)")
    if weights is None:
        V = np.ones_like(X)
    else:
        V = np.array(weights, dtype=dt, copy=True)
        if V.shape != X.shape:
            raise ValueError("weights must have the shape of X")
        if not np.all(np.isfinite(V)):
            raise ValueError("weights must be finite")
        if np.any(V < 0):
            raise ValueError("weights must be non-negative")
    if mask is not None:
        mk = np.asarray(mask)
        if mk.shape != X.shape:
            raise ValueError("mask must have the shape of X")
        V = V * (mk != 0).astype(dt)
    good = V > 0
    X[~good] = 0
    if not np.all(np.isfinite(X)):
        raise ValueError("X must be finite where the weight is positive")
    if np.any(X < 0):
        raise ValueError("X must be non-negative where the weight is positive")
    return X, V


def init_factors(X, V, k, seed=None, dtype=None):
    """Seeded U(0,1) start scaled so that W0 H0 has the mean of X."""
    dt = resolve_dtype(X, dtype)
    rng = np.random.default_rng(seed)
    n, m = X.shape[-2], X.shape[-1]
    lead = X.shape[:-2]
    good = V > 0
    mean = float(X[good].mean()) if good.any() else 1.0
    scale = np.sqrt(max(mean, 1e-300) / k)
    W0 = (rng.random(lead + (n, k)) * scale).astype(dt)
    H0 = (rng.random(lead + (k, m)) * scale).astype(dt)
    return W0, H0


def chi2(X, V, W, H) -> float:
    """sum V (X - WH)^2, accumulated in float64 whatever the compute dtype."""
    R = X - W @ H
    return float(np.sum(V * R * R, dtype=np.float64))


def h_partials(X, V, W, H):
    """k x M numerator and denominator of the H update for one row block."""
    VX = V * X
    VWH = V * (W @ H)
    return W.T @ VX, W.T @ VWH


def w_update(X, V, W, H, eps):
    """Row-local W update (needs only the rows it is given and H)."""
    VX = V * X
    VWH = V * (W @ H)
    num = VX @ H.T
    den = VWH @ H.T
    return W * num / np.maximum(den, W.dtype.type(eps))


def iteration(X, V, W, H, eps, update_W=True, update_H=True):
    """One H-then-W iteration; returns (W, H, chi2 after both updates)."""
    dt = W.dtype
    if update_H:
        num, den = h_partials(X, V, W, H)
        H = (H * num / np.maximum(den, dt.type(eps))).astype(dt)
    if update_W:
        W = w_update(X, V, W, H, eps).astype(dt)
    return W, H, chi2(X, V, W, H)


def converged(prev, cur, tol) -> bool:
    return tol is not None and tol > 0 and (prev - cur) <= tol * prev


def solve(X, V, W0, H0, max_iter, tol=0.0, eps=None, update_W=True, update_H=True):
    """The loop on prepared operands.  Returns (W, H, hist float64[n_iter+1])."""
    dt = X.dtype
    eps = default_eps(dt) if eps is None else eps
    W = np.array(W0, dtype=dt, copy=True)
    H = np.array(H0, dtype=dt, copy=True)
    hist = [chi2(X, V, W, H)]
    for _ in range(int(max_iter)):
        W, H, c = iteration(X, V, W, H, eps, update_W, update_H)
        hist.append(c)
        if not np.isfinite(c):
            raise RuntimeError("chi2 became non-finite")
        if converged(hist[-2], hist[-1], tol):
            break
    return W, H, np.asarray(hist, dtype=np.float64)


def solve_sharded(X, V, W0, H0, max_iter, n_shards, tol=0.0, eps=None):
    """Same loop with rows split into n_shards contiguous blocks; only the
    k x M partials and the chi2 scalar cross shard boundaries.
    This is the schedule the CUDA path runs: partials of (W, H) ->
    [sum, H update, per shard: W update, chi2, next partials]."""
    dt = X.dtype
    eps = default_eps(dt) if eps is None else eps
    n = X.shape[0]
    bounds = shard_bounds(n, n_shards)
    W = np.array(W0, dtype=dt, copy=True)
    H = np.array(H0, dtype=dt, copy=True)

    def partials():
        num = np.zeros(H.shape, np.float64)
        den = np.zeros(H.shape, np.float64)
        c = 0.0
        for lo, hi in bounds:
            a, b = h_partials(X[lo:hi], V[lo:hi], W[lo:hi], H)
            num += a
            den += b
            c += chi2(X[lo:hi], V[lo:hi], W[lo:hi], H)
        return num, den, c

    num, den, c = partials()
    hist = [c]
    for _ in range(int(max_iter)):
        H = (H * num.astype(dt) / np.maximum(den.astype(dt), dt.type(eps))).astype(dt)
        for lo, hi in bounds:
            W[lo:hi] = w_update(X[lo:hi], V[lo:hi], W[lo:hi], H, eps)
        num, den, c = partials()
        hist.append(c)
        if converged(hist[-2], hist[-1], tol):
            break
    return W, H, np.asarray(hist, dtype=np.float64)


def shard_bounds(n, n_shards):
    """Contiguous, as-even-as-possible row blocks (first n % G get one more)."""
    base, rem = divmod(int(n), int(n_shards))
    out, lo = [], 0
    for g in range(int(n_shards)):
        hi = lo + base + (1 if g < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def nmf(X, k, *, weights=None, mask=None, W0=None, H0=None, max_iter=1000,
        tol=1e-5, dtype=None, seed=None, eps=None, update_W=True, update_H=True):
    """Oracle restatement of the entry point (single or batched X)."""
    X, V = prepare(X, weights, mask, dtype)
    dt = X.dtype
    if k < 1:
        raise ValueError("k must be >= 1")
    if W0 is None or H0 is None:
        Wi, Hi = init_factors(X, V, k, seed, dt)
        W0 = Wi if W0 is None else W0
        H0 = Hi if H0 is None else H0
    W0 = np.asarray(W0, dtype=dt)
    H0 = np.asarray(H0, dtype=dt)
    if W0.shape != X.shape[:-1] + (k,) or H0.shape != X.shape[:-2] + (k, X.shape[-1]):
        raise ValueError("W0/H0 shape mismatch")
    if np.any(W0 < 0) or np.any(H0 < 0):
        raise ValueError("W0 and H0 must be non-negative")
    if X.ndim == 2:
        return solve(X, V, W0, H0, max_iter, tol, eps, update_W, update_H)
    Ws, Hs, hs = [], [], []
    for b in range(X.shape[0]):
        w, h, c = solve(X[b], V[b], W0[b], H0[b], max_iter, tol, eps, update_W, update_H)
        Ws.append(w)
        Hs.append(h)
        hs.append(c)
    n = max(len(c) for c in hs)
    hist = np.full((X.shape[0], n), np.nan)
    for b, c in enumerate(hs):
        hist[b, :len(c)] = c
    return np.stack(Ws), np.stack(Hs), hist


# ---------------------------------------------------------------------------
# Timed CPU baseline: the same iteration, row blocks fanned over host threads
# (NumPy releases the GIL inside ufuncs and BLAS), float64 partial sums.
# ---------------------------------------------------------------------------
def iteration_threaded(X, V, W, H, eps, pool, n_blocks):
    dt = W.dtype
    bounds = [b for b in shard_bounds(X.shape[0], n_blocks) if b[1] > b[0]]
    parts = list(pool.map(lambda b: h_partials(X[b[0]:b[1]], V[b[0]:b[1]], W[b[0]:b[1]], H), bounds))
    num = sum(p[0].astype(np.float64) for p in parts)
    den = sum(p[1].astype(np.float64) for p in parts)
    H = (H * num.astype(dt) / np.maximum(den.astype(dt), dt.type(eps))).astype(dt)

    def wblock(b):
        lo, hi = b
        W[lo:hi] = w_update(X[lo:hi], V[lo:hi], W[lo:hi], H, eps)
        return chi2(X[lo:hi], V[lo:hi], W[lo:hi], H)

    c = float(sum(pool.map(wblock, bounds)))
    return W, H, c
