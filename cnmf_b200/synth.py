"""Synthetic absorption / backscatter-like spectra.

Host (NumPy) restatement of the counter-based generator that the device kernel
``cnmf_synth_*`` (csrc/synth.cu) implements: both hash the 64-bit element
index with the same splitmix64 finaliser, so a slice generated on the host
equals the same slice generated on the GPU to rounding of exp/log/cos.

    H*  : k basis spectra on the wavelength grid, max-normalised (host only)
    W*  : LogNormal(0, 0.5) abundances, one normal per (row, component)
    S   = W* H*,  sigma = 0.02 S + 1e-4,  X = max(S + sigma * eps, 0)
    V   : "ones" | "ivar" (1/sigma^2) | "mask" (0/1) | "ivar_mask"
"""
from __future__ import annotations

import numpy as np

GOLD = np.uint64(0x9E3779B97F4A7C15)
STRM = np.uint64(0xD1B54A32D192ED03)
M1 = np.uint64(0xBF58476D1CE4E5B9)
M2 = np.uint64(0x94D049BB133111EB)

STREAM_ABUND = 1
STREAM_NOISE = 2
STREAM_MASK = 3

VMODES = {"ones": 0, "ivar": 1, "mask": 2, "ivar_mask": 3}


def wavelengths(m: int) -> np.ndarray:
    if m == 81:
        return 350.0 + 5.0 * np.arange(81)
    if m == 301:
        return 400.0 + 1.0 * np.arange(301)
    if m == 184:
        return 400.0 + 3.0 * np.arange(184)
    return np.linspace(400.0, 700.0, m)


def basis(k: int, m: int, seed: int = 0) -> np.ndarray:
    """k IOP-like basis spectra (float64, k x m), each scaled to max 1."""
    lam = wavelengths(m)
    rng = np.random.default_rng(1000003 * (seed + 1))
    H = np.empty((k, m))
    for a in range(k):
        u = rng.random(3)
        fam = a % 4
        if fam == 0:      # CDOM / detritus exponential
            s = 0.010 + 0.010 * u[0]
            h = np.exp(-s * (lam - 440.0))
        elif fam == 1:    # phytoplankton: blue + red Gaussian
            c1 = 430.0 + 20.0 * u[0]
            h = np.exp(-0.5 * ((lam - c1) / 40.0) ** 2) \
                + (0.3 + 0.4 * u[1]) * np.exp(-0.5 * ((lam - 675.0) / 12.0) ** 2)
        elif fam == 2:    # particulate backscatter power law
            eta = 2.0 * u[0]
            h = (lam / 550.0) ** (-eta)
        else:             # water-like term rising into the red
            h = np.exp((lam - lam.max()) / (40.0 + 40.0 * u[0])) + 0.02
        H[a] = h / h.max()
    return H


def _hash(seed: int, stream: int, idx: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + GOLD * (idx.astype(np.uint64) + np.uint64(1)) \
            + STRM * np.uint64(stream)
        z ^= z >> np.uint64(30)
        z *= M1
        z ^= z >> np.uint64(27)
        z *= M2
        z ^= z >> np.uint64(31)
    return z


def _uniform_pair(h: np.ndarray):
    u1 = ((h >> np.uint64(32)).astype(np.float64) + 0.5) * (1.0 / 4294967296.0)
    u2 = ((h & np.uint64(0xFFFFFFFF)).astype(np.float64) + 0.5) * (1.0 / 4294967296.0)
    return u1, u2


def _normal(seed: int, stream: int, idx: np.ndarray) -> np.ndarray:
    u1, u2 = _uniform_pair(_hash(seed, stream, idx))
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


def spectra(n: int, m: int, k: int, seed: int = 0, vmode: str = "ivar",
            p_missing: float = 0.0, dtype=np.float64, row0: int = 0,
            Hstar: np.ndarray | None = None):
    """Rows row0 .. row0+n of the synthetic set.  Returns X, V (n x m, dtype)."""
    if vmode not in VMODES:
        raise ValueError("vmode must be one of %s" % sorted(VMODES))
    Hs = basis(k, m, seed) if Hstar is None else np.asarray(Hstar, np.float64)
    rows = (np.arange(n, dtype=np.uint64) + np.uint64(row0))
    ia = rows[:, None] * np.uint64(k) + np.arange(k, dtype=np.uint64)[None, :]
    Wstar = np.exp(0.5 * _normal(seed, STREAM_ABUND, ia))
    # the device sums a = 0..k-1 in order; do the same to stay bit-close
    S = np.zeros((n, m))
    for a in range(k):
        S += Wstar[:, a:a + 1] * Hs[a][None, :]
    ie = rows[:, None] * np.uint64(m) + np.arange(m, dtype=np.uint64)[None, :]
    sig = 0.02 * S + 1e-4
    X = np.maximum(S + sig * _normal(seed, STREAM_NOISE, ie), 0.0)
    mode = VMODES[vmode]
    V = 1.0 / (sig * sig) if (mode & 1) else np.ones_like(S)
    if mode & 2:
        u1, _ = _uniform_pair(_hash(seed, STREAM_MASK, ie))
        V = np.where(u1 < p_missing, 0.0, V)
        X = np.where(u1 < p_missing, 0.0, X)
    return X.astype(dtype), V.astype(dtype)
