"""CPU: the oracle reproduces the committed golden vectors; the C-ABI library
loads and exports every symbol include/cnmf_b200.h declares; host logic."""
import os
import re

import numpy as np
import pytest

from oracle import cnmf_oracle as orc
from tests import make_golden as mg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_oracle_matches_golden(name):
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    X, V, W0, H0, k, iters = mg.inputs(name)
    assert np.isclose(X.sum(dtype=np.float64), g["x_sum"], rtol=1e-12)
    assert np.isclose(V.sum(dtype=np.float64), g["v_sum"], rtol=1e-12)
    W, H, hist = orc.solve(X, V, W0, H0, iters)
    tol = 1e-10 if X.dtype == np.float64 else 1e-4   # BLAS builds may differ in summation order
    assert np.max(np.abs(W - g["W"])) <= tol * np.max(np.abs(g["W"]))
    assert np.max(np.abs(H - g["H"])) <= tol * np.max(np.abs(g["H"]))
    assert np.max(np.abs(hist - g["hist"]) / g["hist"]) <= tol


def test_library_exports_every_declared_symbol():
    import ctypes
    from cnmf_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "cnmf_b200.h")).read()
    declared = set(re.findall(r"CNMF_API[^;(]*?\b(cnmf_\w+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS)
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in declared:
        assert hasattr(lib, s), s
    assert _lib.load().cnmf_version() == 100
    assert _lib.padded_ld(301, _lib.F32) == 304 and _lib.padded_ld(301, _lib.F64) == 302
    assert _lib.padded_ld(184, _lib.F32) == 184 and _lib.padded_ld(81, _lib.F32) == 84


def test_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import cnmf
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cnmf.nmf(np.ones((8, 5)), 2)
    from cnmf_b200 import _lib
    import ctypes as C
    d = _lib.Desc()
    d.n_rows, d.ld, d.m, d.k, d.dtype, d.batch = 100, 84, 81, 2, _lib.F32, 1
    h = C.c_void_p()
    rc = _lib.load().cnmf_plan_create(C.byref(h), C.byref(d))
    assert rc == _lib.ERR_NODEVICE
    assert b"no CPU path" in _lib.load().cnmf_last_error()


def test_argument_errors_before_any_device_work():
    import cnmf
    X = np.ones((8, 5))
    with pytest.raises(ValueError):
        cnmf.nmf(X, 0)
    with pytest.raises(ValueError):
        cnmf.nmf(X, 17)
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, weights=np.ones((4, 5)))
    with pytest.raises(ValueError):
        cnmf.nmf(X[0], 2)
    with pytest.raises(ValueError):
        cnmf.nmf(X, 2, update="X")
    from cnmf_b200 import _lib
    import ctypes as C
    d = _lib.Desc()
    d.n_rows, d.ld, d.m, d.k, d.dtype, d.batch = 100, 81, 81, 2, _lib.F32, 1
    h = C.c_void_p()
    assert _lib.load().cnmf_plan_create(C.byref(h), C.byref(d)) == _lib.ERR_ALIGN
    d.k = 99
    assert _lib.load().cnmf_plan_create(C.byref(h), C.byref(d)) == _lib.ERR_ARG


def test_shard_rows_partition():
    from cnmf_b200 import shard_rows
    for n in (1, 7, 1003, 10_000_000):
        for g in (1, 2, 4, 8):
            b = [shard_rows(n, g, r) for r in range(g)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(g - 1))
            assert b == orc.shard_bounds(n, g)


def test_init_factors_matches_oracle():
    from cnmf_b200 import init_factors
    from cnmf_b200 import synth
    X, V = synth.spectra(200, 81, 3, seed=2)
    W0, H0 = orc.init_factors(X, V, 3, seed=5)
    W1, H1 = init_factors(200, 81, 3, X[V > 0].mean(), seed=5)
    assert np.array_equal(W0, W1) and np.array_equal(H0, H1)


def test_result_file_round_trip(tmp_path):
    import cnmf
    rng = np.random.default_rng(0)
    W, H, hist = rng.random((7, 3)), rng.random((3, 11)), np.array([5.0, 4.0, 3.5])
    path = str(tmp_path / "fit.npz")
    cnmf.save_result(path, W, H, hist, wavelengths=np.arange(11) + 400.0, k=3, note="x")
    r = cnmf.load_result(path)
    assert np.array_equal(r["W"], W) and np.array_equal(r["H"], H) and np.array_equal(r["chi2_history"], hist)
    assert np.array_equal(r["wavelengths"], np.arange(11) + 400.0) and int(r["k"]) == 3
    with pytest.raises(ValueError):
        cnmf.save_result(path, W, H[:2], hist)
    with pytest.raises(ValueError):
        cnmf.save_result(path, W, H, hist, wavelengths=[1.0, 2.0])


def test_shim_and_restart_argument_errors():
    import cnmf
    with pytest.raises(ValueError):
        cnmf.nmf_restarts(np.ones((2, 3, 4)), 2)
    with pytest.raises(ValueError):
        cnmf.nmf_restarts(np.ones((3, 4)), 2, n_restarts=0)
    with pytest.raises(ValueError):
        cnmf.NMF(np.ones((3, 4)), n_components=2).SolveNMF(W_only=True, H_only=True)


def test_plan_dryrun_picks_the_configs_the_design_states():
    """Host logic only (no device work): the launch configuration of the bench shape, the bigger tiles of
    the 184-band scene, and the cluster sizes of the resident kernel for the small-problem shapes."""
    import ctypes as C
    from cnmf_b200 import _lib
    lib = _lib.load()

    def dry(n, m, k, dt, variant=0, batch=1):
        d, i = _lib.Desc(), _lib.Info()
        d.n_rows, d.m, d.k, d.dtype, d.ld, d.batch, d.variant = n, m, k, dt, _lib.padded_ld(m, dt), batch, variant
        return lib.cnmf_plan_dryrun(C.byref(d), C.byref(i)), i.as_dict()

    rc, i = dry(10_000_000, 301, 6, _lib.F64)  # the bench shape: the fp64 tensor-path kernel, 16 warps, 4 column parts, 7 row owners
    assert rc == 0 and (i["grid"], i["threads"], i["stages"], i["tile_rows"], i["variant"]) == (148, 512, 5, 8, 903407)
    assert i["bytes_per_iter"] == 10_000_000 * 301 * 2 * 8 and i["smem_bytes"] <= 232448
    rc, i = dry(10_000_000, 301, 8, _lib.F64)   # k = 8 still gets its five stages
    assert rc == 0 and (i["stages"], i["variant"]) == (5, 903407) and i["smem_bytes"] <= 232448
    rc, i = dry(10_000_000, 301, 6, _lib.F32)   # fp32 stays on the scalar (FFMA2) stream kernel
    assert rc == 0 and (i["grid"], i["threads"], i["tile_rows"]) == (148, 384, 16) and i["variant"] < 900000
    rc, i = dry(100_000, 301, 4, _lib.F64)      # config 2: k = 4 pads half of the tensor shape, scalar kernel
    assert rc == 0 and (i["threads"], i["stages"], i["tile_rows"]) == (384, 5, 8) and i["variant"] < 900000
    rc, i = dry(1_700_000, 184, 8, _lib.F32)
    assert rc == 0 and (i["stages"], i["tile_rows"]) == (4, 32)
    for shape, cs, rows in (((1000, 81, 2, _lib.F64), 8, 125), ((512, 81, 3, _lib.F32), 2, 256), ((512, 81, 3, _lib.F64), 4, 128),
                            ((1, 5, 1, _lib.F64), 1, 1)):
        rc, i = dry(*shape, variant=_lib.VARIANT_RESIDENT)
        assert rc == 0 and i["variant"] == 700000 + cs and i["grid"] == cs and i["tile_rows"] == rows and i["smem_bytes"] <= 232448
    for shape in ((10_000, 301, 6, _lib.F64), (300, 41, 9, _lib.F64)):  # too big for 8 SMs / k > 8: no resident plan
        assert dry(*shape, variant=_lib.VARIANT_RESIDENT)[0] == _lib.ERR_UNSUPPORTED


def test_plan_dryrun_tensor_path_policy_and_shared_memory():
    """Host logic of the fp64 tensor-path kernel (no device work): over every width and k the plan either takes the scalar
    stream kernel or a tensor-path configuration that fits a B200's 227 KB with at least three stages and at most 512
    threads; the policy's corners are where DESIGN.md section 4 says they are."""
    import ctypes as C
    from cnmf_b200 import _lib
    lib = _lib.load()

    def dry(m, k, dt=_lib.F64):
        d, i = _lib.Desc(), _lib.Info()
        d.n_rows, d.m, d.k, d.dtype, d.ld, d.batch, d.variant = 1_000_000, m, k, dt, _lib.padded_ld(m, dt), 1, 0
        assert lib.cnmf_plan_dryrun(C.byref(d), C.byref(i)) == 0, (m, k)
        return i.as_dict()

    taken = {}
    for k in range(1, 9):
        for m in list(range(1, 40)) + list(range(40, 420, 7)) + [88, 89, 136, 137, 280, 281, 301, 304, 305, 320, 383, 384, 385]:
            i = dry(m, k)
            assert i["smem_bytes"] <= 232448 and i["grid"] == 148
            if i["variant"] >= 900000:
                assert i["variant"] // 1000 == 903 and i["threads"] == 512 and i["tile_rows"] == 8 and i["stages"] >= 3
                assert (i["variant"] // 100) % 10 in (1, 2, 4)
                taken.setdefault(k, []).append(m)
            if k <= 4:
                assert i["variant"] < 900000
    assert min(taken[8]) == 89 and max(taken[8]) == 384 and min(taken[7]) == 89      # 12 .. 48 column groups
    assert min(taken[6]) == 137 and max(taken[6]) == 304 and 301 in taken[6]       # 18 groups up, while five stages fit
    assert min(taken[5]) == 281 and max(taken[5]) == 304
    assert dry(301, 6, _lib.F32)["variant"] < 900000 and dry(301, 8, _lib.F32)["variant"] < 900000
