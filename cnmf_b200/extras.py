"""Callers either side of the hot path: best-of-R restarts,
a class-style shim and result files.  Everything numerical still happens in libcnmf_b200.so.

The upstream snapshot defines none of these (its package is empty), so the names follow the
conventions SURVEY.md section 8b records: the `NMF(...).SolveNMF(...)` shape is the NonnegMFPy-style
convention IOP codes commonly use -- recalled, not verifiable from /root/reference.
"""
from __future__ import annotations

import time

import numpy as np

from ._api import init_factors, nmf


def nmf_restarts(X, k, *, n_restarts=8, weights=None, mask=None, max_iter=1000, tol=1e-5, dtype=None, seed=0,
                 eps=None):
    """Best of `n_restarts` seeded random starts (MU is start-sensitive).

    Small problems (the per-station regime) run as ONE batched launch -- one CTA per restart;
    larger ones run the streaming kernel once per start.  Returns (W, H, chi2_history, finals)
    of the best start, `finals` being the last chi2 of every start."""
    X = np.asarray(X)
    if X.ndim != 2:
        raise ValueError("X must be (N, M)")
    if n_restarts < 1:
        raise ValueError("n_restarts must be >= 1")
    n, m = X.shape
    dt = np.dtype(dtype) if dtype is not None else (X.dtype if X.dtype in (np.float32, np.float64) else np.dtype(np.float64))
    good = np.ones(X.shape, bool) if weights is None else (np.asarray(weights) > 0)
    if mask is not None:
        good &= np.asarray(mask) != 0
    mean = float(X[good].sum(dtype=np.float64)) / max(float(good.sum()), 1.0)
    starts = [init_factors(n, m, k, mean, seed=[int(seed), r], dtype=dt) for r in range(n_restarts)]
    if n * m * n_restarts <= (1 << 26) and n_restarts > 1:
        rep = lambda a: None if a is None else np.broadcast_to(np.asarray(a), (n_restarts,) + X.shape)
        W, H, hist = nmf(rep(X), k, weights=rep(weights), mask=rep(mask), W0=np.stack([s[0] for s in starts]),
                         H0=np.stack([s[1] for s in starts]), max_iter=max_iter, tol=tol, dtype=dt, eps=eps)
        finals = np.array([h[~np.isnan(h)][-1] for h in hist])
        b = int(np.argmin(finals))
        return W[b], H[b], hist[b][~np.isnan(hist[b])], finals
    best, finals = None, []
    for W0, H0 in starts:
        W, H, hist = nmf(X, k, weights=weights, mask=mask, W0=W0, H0=H0, max_iter=max_iter, tol=tol, dtype=dt, eps=eps)
        finals.append(hist[-1])
        if best is None or hist[-1] < best[2][-1]:
            best = (W, H, hist)
    return best[0], best[1], best[2], np.array(finals)


class NMF:
    """Class-style shim: `NMF(X, W=, H=, V=, M=, n_components=).SolveNMF(maxiters=, tol=)`.

    X : (N, M) data; V : weights (inverse variance); M : mask, True = good.  After SolveNMF the
    factors are in `.W`, `.H`, the trace in `.chi2_history`; it returns (chi2, seconds)."""

    def __init__(self, X, W=None, H=None, V=None, M=None, n_components=5, seed=None):
        self.X = np.asarray(X)
        self.V, self.M = V, M
        self.n_components = int(W.shape[1]) if W is not None else (int(H.shape[0]) if H is not None else int(n_components))
        self.W, self.H = W, H
        self.seed = seed
        self.chi2_history = None

    def SolveNMF(self, maxiters=1000, tol=1e-5, W_only=False, H_only=False):
        if W_only and H_only:
            raise ValueError("W_only and H_only exclude each other")
        t0 = time.perf_counter()
        self.W, self.H, self.chi2_history = nmf(self.X, self.n_components, weights=self.V, mask=self.M, W0=self.W,
                                                H0=self.H, max_iter=maxiters, tol=tol, seed=self.seed,
                                                update="W" if W_only else ("H" if H_only else "both"))
        return float(self.chi2_history[-1]), time.perf_counter() - t0

    @property
    def cost(self):
        return None if self.chi2_history is None else float(self.chi2_history[-1])


def save_result(path, W, H, chi2_history, wavelengths=None, **meta):
    """W, H, chi2 trace (+ wavelength grid and free-form metadata) -> one .npz file."""
    W, H = np.asarray(W), np.asarray(H)
    if W.shape[-1] != H.shape[-2]:
        raise ValueError("W and H disagree on k")
    if wavelengths is not None and len(wavelengths) != H.shape[-1]:
        raise ValueError("wavelengths must have one entry per column of H")
    extra = {"meta_" + k: np.asarray(v) for k, v in meta.items()}
    np.savez_compressed(path, W=W, H=H, chi2_history=np.asarray(chi2_history, np.float64),
                        wavelengths=np.asarray([] if wavelengths is None else wavelengths, np.float64), **extra)


def load_result(path):
    """-> dict with W, H, chi2_history, wavelengths (None if absent) and the metadata."""
    with np.load(path) as z:
        out = {"W": z["W"], "H": z["H"], "chi2_history": z["chi2_history"],
               "wavelengths": z["wavelengths"] if z["wavelengths"].size else None}
        out.update({k[5:]: z[k] for k in z.files if k.startswith("meta_")})
    return out
