"""The fragment algebra of the fp64 tensor-path kernel (cnmf_b200/csrc/mma_kernel.cuh), replayed on the CPU: a lane-level
NumPy model of mma.sync.m8n8k4.f64 (A[lane/4][lane%4], B[lane%4][lane/4], C/D[lane/4][2(lane%4)+j]) runs role W and role H
of one 8-row tile exactly as the kernel indexes them -- sigma / rho permutations, the column pair (role W) resp. row pair
(role H) of an accumulator becoming the contracted index of the next product -- and must give the oracle's numerators and
denominators.  No device work: this pins the index maps the header comments state, not the CUDA code."""
import numpy as np

from oracle import cnmf_oracle as orc

LANES = np.arange(32)
M8, C4 = LANES >> 2, LANES & 3


def sigma(i):
    return (i & 4) | ((i & 1) << 1) | ((i >> 1) & 1)


def rho(n):
    return (n >> 1) + 4 * (n & 1)


def mma(c, a, b):
    """d = A B + C on per-lane fragments: a, b are [32], c is [32, 2]."""
    A = np.zeros((8, 4)); B = np.zeros((4, 8)); C = np.zeros((8, 8))
    A[M8, C4] = a
    B[C4, M8] = b
    C[M8, 2 * C4] = c[:, 0]
    C[M8, 2 * C4 + 1] = c[:, 1]
    D = A @ B + C
    return np.stack([D[M8, 2 * C4], D[M8, 2 * C4 + 1]], axis=1)


def test_fragment_chaining_matches_the_oracle_contractions():
    rng = np.random.default_rng(5)
    for k in (3, 6, 8):
        ng = 5
        m = 8 * ng
        X, V = rng.random((8, m)), rng.random((8, m))
        W, H = rng.random((8, k)), rng.random((k, m))
        Hp = np.zeros((8, m)); Hp[:k] = H          # rows K..7 of H are zero in shared memory
        Wp = np.zeros((8, 8)); Wp[:, :k] = W

        # ---- role W: lane (m8, c4) works on row sigma(m8), columns 8 g + 2 c4 + j
        row = sigma(M8)
        num = np.zeros((32, 2)); den = np.zeros((32, 2))
        for g in range(ng):
            cols = 8 * g + 2 * C4
            x = np.stack([X[row, cols], X[row, cols + 1]], axis=1)
            v = np.stack([V[row, cols], V[row, cols + 1]], axis=1)
            wh = mma(np.zeros((32, 2)), Wp[row, C4], Hp[C4, 8 * g + M8])          # components 0..3
            wh = mma(wh, Wp[row, C4 + 4], Hp[C4 + 4, 8 * g + M8])               # components 4..7
            np.testing.assert_allclose(wh, np.stack([(W @ H)[row, cols], (W @ H)[row, cols + 1]], axis=1), rtol=1e-13)
            for j in range(2):                                                   # the column pair is the contracted index
                ht = Hp[sigma(M8), cols + j]
                num = mma(num, v[:, j] * x[:, j], ht)
                den = mma(den, v[:, j] * wh[:, j], ht)
        num_ref, den_ref = (V * X) @ H.T, (V * (W @ H)) @ H.T
        for j in range(2):
            a = sigma(2 * C4 + j)
            ok = a < k
            np.testing.assert_allclose(num[ok, j], num_ref[row[ok], a[ok]], rtol=1e-12)
            np.testing.assert_allclose(den[ok, j], den_ref[row[ok], a[ok]], rtol=1e-12)

        # ---- role H: lane (m8, c4) works on column 8 g + m8, rows rho(2 c4 + j) = c4 + 4 j
        Wn = Wp * 1.0
        for g in range(ng):
            col = 8 * g + M8
            r0, r1 = C4, C4 + 4
            assert np.array_equal(rho(2 * C4), r0) and np.array_equal(rho(2 * C4 + 1), r1)
            x = np.stack([X[r0, col], X[r1, col]], axis=1)
            v = np.stack([V[r0, col], V[r1, col]], axis=1)
            wh = mma(np.zeros((32, 2)), Hp[C4, col], Wn[rho(M8), C4])           # (W'H)^T = H^T W'^T
            wh = mma(wh, Hp[C4 + 4, col], Wn[rho(M8), C4 + 4])
            np.testing.assert_allclose(wh, np.stack([(W @ H)[r0, col], (W @ H)[r1, col]], axis=1), rtol=1e-13)
            nH = np.zeros((32, 2)); dH = np.zeros((32, 2))
            for t, r in enumerate((r0, r1)):                                     # the row pair is the contracted index
                nH = mma(nH, v[:, t] * x[:, t], Wn[r, M8])
                dH = mma(dH, v[:, t] * wh[:, t], Wn[r, M8])
            nref, dref = W.T @ (V * X), W.T @ (V * (W @ H))
            for j in range(2):
                a = 2 * C4 + j
                ok = a < k
                np.testing.assert_allclose(nH[ok, j], nref[a[ok], col[ok]], rtol=1e-12)
                np.testing.assert_allclose(dH[ok, j], dref[a[ok], col[ok]], rtol=1e-12)

        # the two sets of sums are what one oracle iteration divides (W update with H, H update with the same W here)
        Wo = W * num_ref / np.maximum(den_ref, orc.default_eps(np.float64))
        assert np.all(np.isfinite(Wo))


def test_row_and_component_permutations_are_bank_conflict_free():
    """What the permutations are for: with rows 8 (mod 32) words apart, the two rows of a quarter-warp (LDS.128 of role W) must be
    two pitches apart, and the four rows of a half-warp (LDS.64 of role H) four different pitches mod 4."""
    for q in range(4):
        a, b = sigma(2 * q), sigma(2 * q + 1)
        assert (a - b) % 4 == 2
    for j in range(2):
        rows = rho(2 * np.arange(4) + j)
        assert sorted(rows % 4) == [0, 1, 2, 3]
    assert sorted(sigma(np.arange(8))) == list(range(8)) and sorted(rho(np.arange(8))) == list(range(8))
