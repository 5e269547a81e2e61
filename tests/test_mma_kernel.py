"""GPU parity of the fp64 tensor-path stream kernel (cnmf_b200/csrc/mma_kernel.cuh) against the NumPy oracle,
through the C ABI, for every setting of its two product knobs (CNMF_MMA = 0 scalar kernel / 1 twelve-warp /
3 sixteen-warp form, CNMF_MMA_TW = column parts of a tile), on ragged shapes: row counts that are not multiples
of the 8-row tile (incl. an odd rows x k product, which takes the W rows off the bulk-copy path), widths that
are not multiples of the 8-column group, every k the kernel has an instance for, W-only / H-only passes, the
tolerance stop, a grid smaller than the device, and the default policy.  Tolerance: BASELINE.json's 1e-10 (fp64)."""
import numpy as np
import pytest

from oracle import cnmf_oracle as orc
from tests.test_gpu_parity import assert_parity, case

pytestmark = pytest.mark.gpu

MODES = [pytest.param("3", "2", id="16warps-2parts"), pytest.param("3", "1", id="16warps-1part"),
         pytest.param("3", "4", id="16warps-4parts"), pytest.param("1", "2", id="12warps-2parts"),
         pytest.param("1", "4", id="12warps-4parts")]


def setenv(monkeypatch, mode, tw):
    monkeypatch.setenv("CNMF_ONE_CTA", "0")  # the streaming plan, not the resident cluster kernel
    monkeypatch.setenv("CNMF_MMA", mode)
    monkeypatch.setenv("CNMF_MMA_TW", tw)


def variant_of(n, m, k):
    from cnmf_b200 import Plan
    plan = Plan(n, m, k, np.float64)
    v = plan.info["variant"]
    plan.close()
    return v


@pytest.mark.parametrize("mode,tw", MODES)
@pytest.mark.parametrize("n,m,k", [(64, 20, 2), (5003, 301, 6), (2001, 81, 4), (1237, 184, 8), (777, 77, 7), (40, 301, 3),
                                   (3, 17, 1), (1, 5, 1), (999, 320, 5), (8, 8, 8), (4097, 9, 6)])
def test_tensor_path_matches_the_oracle(n, m, k, mode, tw, monkeypatch):
    import cnmf
    setenv(monkeypatch, mode, tw)
    assert variant_of(n, m, k) // 1000 == 900 + int(mode)
    X, V, W0, H0 = case(n, m, k, seed=n + m + k, vmode="ivar_mask", p=0.1)
    got = cnmf.nmf(X, k, weights=V, W0=W0, H0=H0, max_iter=25, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 25), np.float64)


@pytest.mark.parametrize("mode,tw", MODES[:1] + MODES[3:4])
@pytest.mark.parametrize("update", ["W", "H"])
def test_tensor_path_single_factor_updates(update, mode, tw, monkeypatch):
    import cnmf
    setenv(monkeypatch, mode, tw)
    X, V, W0, H0 = case(1501, 301, 6, seed=17, vmode="ivar_mask", p=0.05)
    got = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=12, tol=0, update=update)
    ref = orc.solve(X, V, W0, H0, 12, update_W=update == "W", update_H=update == "H")
    assert_parity(got, ref, np.float64)
    fixed, start = (got[1], H0) if update == "W" else (got[0], W0)
    assert np.array_equal(fixed, start)


@pytest.mark.parametrize("mode,tw", MODES[:1] + MODES[3:4])
def test_tensor_path_tolerance_stop_and_repeatability(mode, tw, monkeypatch):
    import cnmf
    setenv(monkeypatch, mode, tw)
    X, V, W0, H0 = case(6001, 301, 6, seed=29, vmode="ivar_mask", p=0.05)
    a = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=300, tol=1e-2)
    ref = orc.solve(X, V, W0, H0, 300, tol=1e-2)
    assert len(a[2]) == len(ref[2]) < 301
    assert_parity(a, ref, np.float64)
    b = cnmf.nmf(X, 6, weights=V, W0=W0, H0=H0, max_iter=300, tol=1e-2)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)  # which warp takes which chunk varies; the order of every sum does not


@pytest.mark.parametrize("tw", ["4", "2"])
def test_tensor_path_repeats_bit_for_bit_under_load(tw, monkeypatch):
    """Race evidence in place of compute-sanitizer (closed on the pool): 400 000 x 301 rows on all SMs -- 338 tiles per CTA
    through the five-stage ring, every chunk of the shared queue taken by whichever warp is free -- give the same bits in
    three runs, and the scalar stream kernel's answer to rounding."""
    import torch
    from cnmf_b200 import Plan, _lib
    dev = torch.device("cuda", 0)
    n, m, k, iters = 400_000, 301, 6, 8
    ld = _lib.padded_ld(m, _lib.F64)
    X = torch.empty((n, ld), dtype=torch.float64, device=dev)
    V = torch.empty_like(X)
    from cnmf_b200 import synth
    hs = torch.from_numpy(synth.basis(k, m, 0)).to(dev)
    _lib.check(_lib.load().cnmf_synth(_lib.F64, X.data_ptr(), V.data_ptr(), n, 0, m, ld, k, hs.data_ptr(), 0, 3, 0.05, None))
    rng = np.random.default_rng(1)
    W0 = torch.from_numpy(rng.random((n, k))).to(dev)
    H0 = torch.from_numpy(rng.random((k, m))).to(dev)

    def run(mode):
        monkeypatch.setenv("CNMF_MMA", mode)
        monkeypatch.setenv("CNMF_MMA_TW", tw)
        plan = Plan(n, m, k, np.float64, ld=ld)
        assert (plan.info["variant"] >= 900000) == (mode != "0")
        W, H = W0.clone(), H0.clone()
        hist = torch.zeros(iters + 1, dtype=torch.float64, device=dev)
        assert plan.solve(X, V, W, H, hist, iters) == iters
        torch.cuda.synchronize()
        plan.close()
        return W, H, hist

    a = run("3")
    for _ in range(2):
        b = run("3")
        for u, v in zip(a, b):
            assert torch.equal(u, v)
    c = run("0")
    for u, v in zip(a, c):
        assert float((u - v).abs().max() / v.abs().max()) < 1e-12
    h = a[2].cpu().numpy()
    assert np.all(np.isfinite(h)) and np.all(np.diff(h) <= 0)


def test_tensor_path_on_a_small_grid(monkeypatch):
    """Many tiles per CTA through the 5-stage ring (3 SMs walk 1126 tiles), and 1 SM."""
    import torch
    from cnmf_b200 import Plan, prepare_device
    setenv(monkeypatch, "3", "2")
    X, V, W0, H0 = case(9001, 301, 6, seed=71, vmode="ivar_mask", p=0.05)
    dev = torch.device("cuda", 0)
    Xp, Vp = prepare_device(torch.from_numpy(X).to(dev), torch.from_numpy(V).to(dev))
    ref = orc.solve(X, V, W0, H0, 8)
    for lim in (3, 1):
        plan = Plan(9001, 301, 6, np.float64, ld=Xp.shape[1], sm_limit=lim)
        assert plan.info["variant"] // 1000 == 903 and plan.info["grid"] == lim
        W, H = torch.from_numpy(W0).to(dev), torch.from_numpy(H0).to(dev)
        hist = torch.zeros(9, dtype=torch.float64, device=dev)
        assert plan.solve(Xp, Vp, W, H, hist, 8) == 8
        torch.cuda.synchronize()
        assert_parity((W.cpu().numpy(), H.cpu().numpy(), hist.cpu().numpy()), ref, np.float64)
        plan.close()


@pytest.mark.parametrize("tw", ["4", "2"])
def test_tensor_path_wide_rows(tw, monkeypatch):
    """Up to 6 column groups per column-owner warp: 321..384 columns run on the 16-warp form (the 12-warp form stops at 320)."""
    import cnmf
    setenv(monkeypatch, "3", tw)
    assert variant_of(1203, 383, 7) // 1000 == 903
    X, V, W0, H0 = case(1203, 383, 7, seed=5, vmode="ivar_mask", p=0.1)
    got = cnmf.nmf(X, 7, weights=V, W0=W0, H0=H0, max_iter=15, tol=0)
    assert_parity(got, orc.solve(X, V, W0, H0, 15), np.float64)
    monkeypatch.setenv("CNMF_MMA", "1")
    assert variant_of(1203, 383, 7) < 900000


def test_default_policy(monkeypatch):
    """k = 7, 8 and wide k = 6 take the tensor path, everything else (and fp32, k > 8, rows wider than the
    column owners' accumulators) the scalar stream kernel; CNMF_MMA=0 switches it off."""
    from cnmf_b200 import Plan
    monkeypatch.delenv("CNMF_MMA", raising=False)
    monkeypatch.delenv("CNMF_MMA_TW", raising=False)
    assert variant_of(100000, 301, 6) // 1000 == 903
    assert variant_of(100000, 301, 8) // 1000 == 903
    assert variant_of(100000, 184, 7) // 1000 == 903
    assert variant_of(100000, 81, 7) < 900000     # a tie at 11 column groups: the scalar kernel keeps it
    assert variant_of(100000, 81, 6) < 900000
    assert variant_of(100000, 150, 6) == 903207   # two column parts below 32 column groups
    assert variant_of(100000, 301, 6) == 903407   # four from there up
    assert variant_of(100000, 301, 5) == 903407   # k = 5 from 36 column groups
    assert variant_of(100000, 200, 5) < 900000
    assert variant_of(100000, 383, 6) < 900000    # three stages only: not by policy (CNMF_MMA=3 still selects it)
    assert variant_of(100000, 383, 8) // 1000 == 903
    assert variant_of(100000, 301, 4) < 900000
    assert variant_of(100000, 301, 9) < 900000
    assert variant_of(100000, 400, 8) < 900000
    plan = Plan(100000, 301, 6, np.float32)
    assert plan.info["variant"] < 900000
    plan.close()
    monkeypatch.setenv("CNMF_MMA", "0")
    assert variant_of(100000, 301, 6) < 900000
