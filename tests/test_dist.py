"""N > 1 path.  CPU (gloo, world_size 2): the sharding arithmetic and the partial-sum
algebra the ranks exchange.  GPU (nccl, needs >= 2 devices): cnmf.nmf with a Comm
against the single-process oracle."""
import os
import socket

import numpy as np
import pytest

from cnmf_b200 import synth
from oracle import cnmf_oracle as orc


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _case(n=1501, m=81, k=3, dtype=np.float64):
    X, V = synth.spectra(n, m, k, seed=77, vmode="ivar_mask", p_missing=0.1, dtype=dtype)
    W0, H0 = orc.init_factors(X, V, k, seed=78, dtype=dtype)
    return X, V, W0, H0


def _gloo_worker(rank, world, port, iters, out):
    import torch
    import torch.distributed as dist
    from cnmf_b200 import shard_rows
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    X, V, W0, H0 = _case()
    lo, hi = shard_rows(X.shape[0], world, rank)
    X, V, W, H = X[lo:hi], V[lo:hi], W0[lo:hi].copy(), H0.copy()
    eps = orc.default_eps(np.float64)
    k, m = H.shape

    def exchange():
        # the buffer the CUDA path all-reduces: [num(k*m) | den(k*m) | chi2]
        num, den = orc.h_partials(X, V, W, H)
        buf = torch.from_numpy(np.concatenate([num.ravel(), den.ravel(), [orc.chi2(X, V, W, H)]]))
        dist.all_reduce(buf)
        b = buf.numpy()
        return b[:k * m].reshape(k, m), b[k * m:2 * k * m].reshape(k, m), float(b[-1])

    num, den, c = exchange()
    hist = [c]
    for _ in range(iters):
        H = H * num / np.maximum(den, eps)
        W = orc.w_update(X, V, W, H, eps)
        num, den, c = exchange()
        hist.append(c)
    np.savez(out % rank, W=W, H=H, hist=np.asarray(hist), lo=lo, hi=hi)
    dist.destroy_process_group()


def test_gloo_two_ranks_match_single_process(tmp_path):
    import torch.multiprocessing as mp
    iters, world = 40, 2
    out = str(tmp_path / "r%d.npz")
    mp.spawn(_gloo_worker, args=(world, _free_port(), iters, out), nprocs=world, join=True)
    X, V, W0, H0 = _case()
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    parts = [np.load(out % r) for r in range(world)]
    assert parts[0]["lo"] == 0 and parts[-1]["hi"] == X.shape[0] and parts[0]["hi"] == parts[1]["lo"]
    W = np.concatenate([p["W"] for p in parts])
    for p in parts:
        assert np.max(np.abs(p["H"] - Hr)) <= 1e-11 * np.max(Hr)
        assert np.max(np.abs(p["hist"] - hr) / hr) <= 1e-11
    assert np.array_equal(parts[0]["H"], parts[1]["H"])      # identical on every rank: no broadcast needed
    assert np.max(np.abs(W - Wr)) <= 1e-11 * np.max(Wr)


def _nccl_worker(rank, world, port, dtype_name, iters, out, peer):
    import torch
    import torch.distributed as dist
    os.environ["CNMF_PEER"] = "1" if peer else "0"
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import cnmf
    from cnmf_b200 import shard_rows
    dtype = np.dtype(dtype_name)
    X, V, W0, H0 = _case(n=20011, m=301, k=6, dtype=dtype)
    lo, hi = shard_rows(X.shape[0], world, rank)
    comm = cnmf.Comm()
    W, H, hist = cnmf.nmf(X[lo:hi], 6, weights=V[lo:hi], W0=W0[lo:hi], H0=H0, max_iter=iters, tol=0, comm=comm)
    W2, H2, hist2 = cnmf.nmf(X[lo:hi], 6, weights=V[lo:hi], W0=W0[lo:hi], H0=H0, max_iter=500, tol=1e-3, comm=comm)
    np.savez(out % rank, W=W, H=H, hist=hist, W2=W2, H2=H2, hist2=hist2, peer=int(comm.peer))
    comm.close()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("peer", [1, 0], ids=["peer-windows", "nccl"])
@pytest.mark.parametrize("dtype_name", ["float64", "float32"])
def test_nccl_row_sharded_matches_oracle(tmp_path, dtype_name, peer):
    """Row-sharded solve on all GPUs of the box (2-8) against the oracle, with the per-iteration sum going through
    the peer windows (exchange_finalize_kernel) and through ncclAllReduce."""
    import torch
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    iters = 30
    out = str(tmp_path / "g%d.npz")
    mp.spawn(_nccl_worker, args=(world, _free_port(), dtype_name, iters, out, peer), nprocs=world, join=True)
    dtype = np.dtype(dtype_name)
    tol = 1e-10 if dtype == np.float64 else 1e-4
    X, V, W0, H0 = _case(n=20011, m=301, k=6, dtype=dtype)
    Wr, Hr, hr = orc.solve(X, V, W0, H0, iters)
    W2r, H2r, h2r = orc.solve(X, V, W0, H0, 500, tol=1e-3)
    parts = [np.load(out % r) for r in range(world)]
    assert all(int(p["peer"]) == peer for p in parts)         # the path under test is the one that ran
    W = np.concatenate([p["W"] for p in parts])
    assert np.max(np.abs(W - Wr)) <= tol * np.max(Wr)
    for p in parts:
        assert np.array_equal(p["H"], parts[0]["H"]) and np.array_equal(p["hist"], parts[0]["hist"])
        assert np.max(np.abs(p["H"] - Hr)) <= tol * np.max(Hr)
        assert np.max(np.abs(p["hist"] - hr) / hr) <= tol
        assert len(p["hist2"]) == len(h2r)                    # every rank stops at the same iteration
        assert np.max(np.abs(p["H2"] - H2r)) <= tol * np.max(H2r)
