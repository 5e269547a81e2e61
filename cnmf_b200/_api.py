"""Python host of the NMF entry point: `nmf(X, k, weights=, mask=, W0=, H0=,
max_iter=, tol=) -> (W, H, chi2_history)`.

The upstream package pins only its import name (/root/reference/setup.py:9;
/root/reference/cnmf/__init__.py is empty), so the keyword list is the one
BASELINE.json `north_star` states and SURVEY.md section 8(b) freezes.  All
arithmetic happens in libcnmf_b200.so (sm_100a CUDA) through ctypes; torch is
used only to hold device memory and streams.  No CPU fallback exists.
"""
from __future__ import annotations

import ctypes as C
import os

import time

import numpy as np

from . import _lib

_DT = {np.dtype(np.float32): _lib.F32, np.dtype(np.float64): _lib.F64}
_UPDATE = {"both": _lib.UPDATE_BOTH, "W": _lib.UPDATE_W, "H": _lib.UPDATE_H,
           "w": _lib.UPDATE_W, "h": _lib.UPDATE_H}


def _torch():
    import torch
    return torch


def _require_cuda():
    torch = _torch()
    if not torch.cuda.is_available():
        raise RuntimeError("cnmf_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch


def _resolve_dtype(x_dtype, dtype):
    if dtype is not None:
        dt = np.dtype(str(dtype).replace("torch.", "")) if not isinstance(dtype, np.dtype) else dtype
    else:
        dt = np.dtype(str(x_dtype).replace("torch.", "")) if not isinstance(x_dtype, np.dtype) else x_dtype
        if dt not in _DT:
            dt = np.dtype(np.float64)
    if dt not in _DT:
        raise ValueError("dtype must be float32 or float64")
    return dt


def shard_rows(n: int, world_size: int, rank: int):
    """Contiguous row block of `rank` (the first n % world_size ranks hold one more row)."""
    base, rem = divmod(int(n), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def init_factors(n, m, k, mean, seed=None, dtype=np.float64, lead=()):
    """Seeded U(0,1) start scaled so that W0 H0 has the mean of X (same rule as the oracle)."""
    rng = np.random.default_rng(seed)
    scale = np.sqrt(max(float(mean), 1e-300) / k)
    W0 = (rng.random(tuple(lead) + (n, k)) * scale).astype(dtype)
    H0 = (rng.random(tuple(lead) + (k, m)) * scale).astype(dtype)
    return W0, H0


def init_factors_sharded(n_local, m, k, mean, seed, rank, dtype=np.float64):
    """Row-sharded start: H0 is drawn from a stream every rank shares, W0 from a
    per-rank stream (so it differs from the single-GPU default start; pass W0/H0
    explicitly when the two must agree)."""
    s = 0 if seed is None else int(seed)
    scale = np.sqrt(max(float(mean), 1e-300) / k)
    H0 = (np.random.default_rng([s, 0x48]).random((k, m)) * scale).astype(dtype)
    W0 = (np.random.default_rng([s, 0x57, int(rank)]).random((n_local, k)) * scale).astype(dtype)
    return W0, H0


class Comm:
    """Library-owned NCCL communicator, bootstrapped through torch.distributed."""

    def __init__(self, group=None):
        import torch.distributed as dist
        torch = _require_cuda()
        lib = _lib.load()
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        uid = (C.c_uint8 * 128)()
        if self.rank == 0:
            _lib.check(lib.cnmf_comm_unique_id(uid))
        box = [bytes(uid)]
        src = dist.get_global_rank(group, 0) if group is not None else 0
        dist.broadcast_object_list(box, src=src, group=group)
        uid = (C.c_uint8 * 128).from_buffer_copy(box[0])
        h = C.c_void_p()
        _lib.check(lib.cnmf_comm_create(C.byref(h), uid, self.world_size, self.rank))
        self.handle = h
        self.group = group
        torch.cuda.synchronize()
        self.peer = False
        if self.world_size > 1 and os.environ.get("CNMF_PEER", "1") != "0":
            self._attach_peers(dist, lib)

    def _attach_peers(self, dist, lib):
        """Peer windows for the per-iteration exchange (include/cnmf_b200.h, cnmf_comm_peer_*):
        gather the IPC handles, import, and switch on only if EVERY rank could."""
        mine = (C.c_uint8 * 64)()
        ok = lib.cnmf_comm_peer_export(self.handle, mine) == 0
        got = [None] * self.world_size
        dist.all_gather_object(got, (ok, bytes(mine)), group=self.group)
        ok = all(g[0] for g in got)
        if ok:
            blob = b"".join(g[1] for g in got)
            buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
            ok = lib.cnmf_comm_peer_import(self.handle, buf) == 0
        got = [None] * self.world_size
        dist.all_gather_object(got, ok, group=self.group)
        if all(got):
            _lib.check(lib.cnmf_comm_peer_enable(self.handle, 1))
            self.peer = True

    def close(self, barrier=True):
        """Collective when peer windows are attached (every rank calls it): the peers' mappings of
        this rank's window must be gone before the window is freed."""
        if self.handle:
            if self.peer and barrier:
                # the peers' mappings of this rank's window must be gone before it is freed
                try:
                    import torch
                    import torch.distributed as dist
                    torch.cuda.synchronize()
                    if dist.is_initialized():
                        dist.barrier(group=self.group)
                except Exception:
                    pass
            _lib.load().cnmf_comm_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close(barrier=False)
        except Exception:
            pass


class _LocalRank:
    """One rank of a LocalGroup: what Plan / cnmf_nmf_host need of a communicator."""

    def __init__(self, handle, rank, world_size, device):
        self.handle, self.rank, self.world_size, self.device = handle, rank, world_size, device
        self.group, self.peer = None, world_size > 1


class LocalGroup:
    """The ranks of ONE process, one per entry of `devices` (cnmf_comm_create_local): the sharded
    schedule without torchrun -- the per-iteration exchange goes through peer windows reached with
    cudaDeviceEnablePeerAccess, no NCCL.  A device may be listed more than once (virtual shards).
    Every rank needs its own host thread; `run(fn)` starts them, calls fn(rank_comm) on each with the
    rank's device current, and re-raises the first error."""

    def __init__(self, devices, timeout=None):
        _require_cuda()
        lib = _lib.load()
        self.devices = [int(d) for d in devices]
        n = len(self.devices)
        if n < 1:
            raise ValueError("devices must name at least one GPU")
        hs = (C.c_void_p * n)()
        _lib.check(lib.cnmf_comm_create_local(hs, n, (C.c_int32 * n)(*self.devices)))
        self.ranks = [_LocalRank(C.c_void_p(hs[r]), r, n, self.devices[r]) for r in range(n)]
        if timeout is not None:
            for c in self.ranks:
                _lib.check(lib.cnmf_comm_set_timeout(c.handle, float(timeout)))

    def run(self, fn, only=None):
        import threading
        torch = _torch()
        out, err = [None] * len(self.ranks), [None] * len(self.ranks)

        def body(c):
            try:
                torch.cuda.set_device(c.device)
                out[c.rank] = fn(c)
            except BaseException as exc:  # re-raised on the calling thread
                err[c.rank] = exc
        ts = [threading.Thread(target=body, args=(c,)) for c in self.ranks if only is None or c.rank in only]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        for e in err:
            if e is not None:
                raise e
        return out

    def close(self):
        lib = _lib.load()
        for c in self.ranks:
            if c.handle:
                lib.cnmf_comm_destroy(c.handle)
                c.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Plan:
    """Thin owner of a cnmf_plan_t (device-resident operands; used by nmf() and bench.py)."""

    def __init__(self, n_rows, m, k, dtype, ld=None, comm=None, batch=1, eps=None, sm_limit=0, variant=0):
        _require_cuda()
        self.lib = _lib.load()
        self.dtype = np.dtype(dtype)
        code = _DT[self.dtype]
        self.ld = int(ld) if ld is not None else _lib.padded_ld(m, code)
        d = _lib.Desc()
        d.n_rows, d.ld, d.m, d.k, d.dtype, d.batch = int(n_rows), self.ld, int(m), int(k), code, int(batch)
        d.eps = float(eps) if eps else 0.0
        d.comm = comm.handle if comm is not None else None
        d.sm_limit, d.variant = int(sm_limit), int(variant)
        self._comm = comm
        self.handle = C.c_void_p()
        _lib.check(self.lib.cnmf_plan_create(C.byref(self.handle), C.byref(d)))
        self.n_rows, self.m, self.k, self.batch = int(n_rows), int(m), int(k), int(batch)

    @property
    def info(self):
        i = _lib.Info()
        _lib.check(self.lib.cnmf_plan_info(self.handle, C.byref(i)))
        return i.as_dict()

    def solve(self, X, V, W, H, hist, max_iter, tol=0.0, update=_lib.UPDATE_BOTH, check_every=0, stream=None):
        n_iter = C.c_int32(0)
        st = _stream_ptr(stream)
        _lib.check(self.lib.cnmf_solve(self.handle, X.data_ptr(), V.data_ptr(), W.data_ptr(), H.data_ptr(),
                                       hist.data_ptr(), int(max_iter), float(tol or 0.0), int(check_every),
                                       int(update), C.byref(n_iter), st))
        return int(n_iter.value)

    def kernel_time(self):
        """(total ms, launches) of the fused kernel in the last solve that carried SOLVE_TIME_KERNELS."""
        ms, n = C.c_double(0), C.c_int32(0)
        _lib.check(self.lib.cnmf_plan_kernel_time(self.handle, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    def solve_batched(self, X, V, W, H, hist, niter, max_iter, tol=0.0, update=_lib.UPDATE_BOTH, stream=None):
        st = _stream_ptr(stream)
        _lib.check(self.lib.cnmf_solve_batched(self.handle, X.data_ptr(), V.data_ptr(), W.data_ptr(), H.data_ptr(),
                                               hist.data_ptr(), niter.data_ptr(), int(max_iter), float(tol or 0.0),
                                               int(update), st))

    def chi2(self, X, V, W, H, out, stream=None):
        st = _stream_ptr(stream)
        _lib.check(self.lib.cnmf_chi2(self.handle, X.data_ptr(), V.data_ptr(), W.data_ptr(), H.data_ptr(),
                                      out.data_ptr(), st))

    def close(self):
        if self.handle:
            self.lib.cnmf_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _stream_ptr(stream):
    if stream is None:
        stream = _torch().cuda.current_stream()
    return C.c_void_p(int(stream.cuda_stream) if hasattr(stream, "cuda_stream") else int(stream))


def prepare_device(X, weights=None, mask=None, dtype=None):
    """Device-side fold of weights*mask into V, zero-fill of X under V == 0 and
    padding of the pitch to 16 bytes.  X: CUDA tensor (..., N, M).  Returns (Xp, Vp)
    of shape (..., N, ld); raises ValueError on negative / non-finite data."""
    torch = _require_cuda()
    lib = _lib.load()
    dt = _resolve_dtype(X.dtype, dtype)
    tdt = torch.float32 if dt == np.float32 else torch.float64
    X = X.to(tdt).contiguous()
    lead, (n, m) = tuple(X.shape[:-2]), X.shape[-2:]
    rows = int(np.prod(lead, dtype=np.int64)) * n if lead else n
    wp = None
    if weights is not None:
        weights = weights.to(device=X.device, dtype=tdt).contiguous()
        if tuple(weights.shape) != tuple(X.shape):
            raise ValueError("weights must have the shape of X")
        wp = weights.data_ptr()
    mp = None
    if mask is not None:
        mask = (mask.to(device=X.device) != 0).to(torch.uint8).contiguous()
        if tuple(mask.shape) != tuple(X.shape):
            raise ValueError("mask must have the shape of X")
        mp = mask.data_ptr()
    ld = _lib.padded_ld(m, _DT[dt])
    Xp = torch.empty(lead + (n, ld), dtype=tdt, device=X.device)
    Vp = torch.empty_like(Xp)
    status = torch.zeros(1, dtype=torch.int32, device=X.device)
    _lib.check(lib.cnmf_prepare(_DT[dt], X.data_ptr(), wp, mp, m, m, Xp.data_ptr(), Vp.data_ptr(), ld, rows, m,
                                status.data_ptr(), _stream_ptr(None)))
    s = int(status.item())
    if s:
        what = [t for b, t in ((1, "X is negative where the weight is positive"),
                               (2, "X is not finite where the weight is positive"),
                               (4, "weights are negative"), (8, "weights are not finite")) if s & b]
        raise ValueError("; ".join(what))
    return Xp, Vp


def _check_factors(W0, H0, lead, n, m, k):
    if tuple(W0.shape) != tuple(lead) + (n, k) or tuple(H0.shape) != tuple(lead) + (k, m):
        raise ValueError("W0/H0 shape mismatch: expected %s and %s" % (tuple(lead) + (n, k), tuple(lead) + (k, m)))


def _global_mean(sum_x, cnt, comm):
    if comm is None or comm.world_size == 1:
        return sum_x / max(cnt, 1.0)
    import torch.distributed as dist
    torch = _torch()
    t = torch.tensor([sum_x, cnt], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, group=comm.group)
    return float(t[0]) / max(float(t[1]), 1.0)


def nmf(X, k, *, weights=None, mask=None, W0=None, H0=None, max_iter=1000, tol=1e-5, dtype=None, seed=None,
        eps=None, update="both", comm=None, check_every=10, devices=None):
    """Weighted / masked multiplicative-update NMF on a B200.

    X : (N, M) spectra, or (B, N, M) for B independent problems; NumPy array or
        torch tensor (CPU or CUDA).  The same kind comes back.
    k : number of components (1..16).
    weights : per-element non-negative weights (inverse variance), shape of X.
    mask : True / non-zero = good; folded into the weights.
    W0, H0 : initial factors ((N, k), (k, M)); seeded U(0,1) * sqrt(mean(X)/k) if omitted.
    max_iter, tol : stop at max_iter, or when the relative chi2 decrease of one
        iteration is <= tol (tol = 0 or None: fixed iteration count).
    update : "both" | "W" | "H" (fit abundances for a fixed basis, or the reverse).
    comm : a cnmf_b200.Comm; X, weights, mask and W0 are then THIS rank's rows and
        only the k x M partials and chi2 are all-reduced each iteration.
    devices : GPU indices (e.g. range(8)) for a single-process multi-GPU solve of a host-resident
        (N, M) problem: contiguous row blocks go to one GPU each, every GPU uploads and solves its
        block on its own host thread, the per-iteration exchange runs through peer windows, and W
        comes back as one array.  The default start is the single-GPU one, so the result equals
        `devices=None` to rounding of the summation order.

    Returns (W, H, chi2_history) with chi2_history float64 of length n_iter + 1
    (entry 0 is chi2 at W0, H0).  Raises ValueError on bad arguments / data and
    RuntimeError on CUDA / NCCL failure or non-finite chi2.
    """
    if update not in _UPDATE:
        raise ValueError("update must be 'both', 'W' or 'H'")
    upd = _UPDATE[update]
    if int(k) != k or k < 1 or k > _lib.MAX_K:
        raise ValueError("k must be an integer in 1..%d" % _lib.MAX_K)
    k = int(k)
    if max_iter < 0:
        raise ValueError("max_iter must be >= 0")
    tol = 0.0 if tol is None else float(tol)
    is_torch = type(X).__module__.startswith("torch")
    ndim = X.ndim
    if ndim not in (2, 3):
        raise ValueError("X must be (N, M) or (B, N, M)")
    for name, arr in (("weights", weights), ("mask", mask)):
        if arr is not None and tuple(arr.shape) != tuple(X.shape):
            raise ValueError("%s must have the shape of X" % name)
    if ndim == 3 and (comm is not None or devices is not None):
        raise ValueError("batched problems are not sharded: split the batch across ranks instead")
    dt = _resolve_dtype(X.dtype, dtype)
    if devices is not None:
        if comm is not None:
            raise ValueError("give either comm (one process per GPU) or devices (one process), not both")
        if is_torch and X.is_cuda:
            raise ValueError("devices= shards host-resident data; a CUDA tensor already lives on one GPU")
        return _nmf_devices(X, k, weights, mask, W0, H0, int(max_iter), tol, dt, seed, eps, upd, list(devices), is_torch)
    if not is_torch and ndim == 2:
        return _nmf_host(X, k, weights, mask, W0, H0, int(max_iter), tol, dt, seed, eps, upd, comm)
    return _nmf_device(X, k, weights, mask, W0, H0, int(max_iter), tol, dt, seed, eps, upd, comm, is_torch,
                       int(check_every))


def _nmf_host(X, k, weights, mask, W0, H0, max_iter, tol, dt, seed, eps, upd, comm):
    """NumPy (N, M) input: one cnmf_nmf_host call does upload, prepare, solve, download."""
    _require_cuda()
    lib = _lib.load()
    t_enter = time.perf_counter()
    X = np.ascontiguousarray(X, dtype=dt)
    n, m = X.shape
    w = None if weights is None else np.ascontiguousarray(weights, dtype=dt)
    mk = None if mask is None else np.ascontiguousarray(np.asarray(mask) != 0, dtype=np.uint8)
    if W0 is None or H0 is None:
        good = np.ones(X.shape, bool) if w is None else (w > 0)
        if mk is not None:
            good &= mk != 0
        sx, cnt = float(X[good].sum(dtype=np.float64)), float(good.sum())
        mean = _global_mean(sx, cnt, comm)
        if comm is None:
            Wi, Hi = init_factors(n, m, k, mean, seed, dt)
        else:
            Wi, Hi = init_factors_sharded(n, m, k, mean, seed, comm.rank, dt)
        W0 = Wi if W0 is None else W0
        H0 = Hi if H0 is None else H0
    W = np.array(W0, dtype=dt, order="C", copy=True)
    H = np.array(H0, dtype=dt, order="C", copy=True)
    _check_factors(W, H, (), n, m, k)
    if (W < 0).any() or (H < 0).any():
        raise ValueError("W0 and H0 must be non-negative")
    hist = np.full(max_iter + 1, np.nan)
    n_iter = C.c_int32(0)
    up, down = C.c_int64(0), C.c_int64(0)
    vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    t_call = time.perf_counter()
    rc = lib.cnmf_nmf_host(_DT[dt], n, m, k, vp(X), vp(w), vp(mk), vp(W), vp(H), vp(hist), max_iter, tol,
                           float(eps) if eps else 0.0, upd, comm.handle if comm is not None else None,
                           C.byref(n_iter), C.byref(up), C.byref(down))
    if rc == _lib.ERR_NONFINITE:
        raise RuntimeError("chi2 became non-finite: " + lib.cnmf_last_error().decode())
    _lib.check(rc)
    hist = hist[:n_iter.value + 1]
    if not np.all(np.isfinite(hist)):
        raise RuntimeError("chi2 became non-finite")
    nmf.last_transfer = {"h2d_bytes": int(up.value), "d2h_bytes": int(down.value)}
    ph = (C.c_double * 6)()
    lib.cnmf_host_timing(ph, 6)
    nmf.last_timing = {"host_checks_and_copies_s": t_call - t_enter, "alloc_s": ph[0], "upload_prepare_s": ph[1], "plan_s": ph[2],
                       "solve_s": ph[3], "download_s": ph[4], "library_call_s": ph[5]}
    return W, H, hist


def _nmf_devices(X, k, weights, mask, W0, H0, max_iter, tol, dt, seed, eps, upd, devices, is_torch):
    """Host-resident (N, M) input sharded over the GPUs of this process: one host thread per GPU calls
    cnmf_nmf_host on its row block (upload, prepare, solve, download run in parallel across the GPUs)."""
    _require_cuda()
    lib = _lib.load()
    t_enter = time.perf_counter()
    to_np = lambda a: None if a is None else (a.numpy() if type(a).__module__.startswith("torch") else np.asarray(a))
    X = np.ascontiguousarray(to_np(X), dtype=dt)
    n, m = X.shape
    w = None if weights is None else np.ascontiguousarray(to_np(weights), dtype=dt)
    mk = None if mask is None else np.ascontiguousarray(to_np(mask) != 0, dtype=np.uint8)
    if n < len(devices):
        devices = devices[:max(n, 1)]
    if W0 is None or H0 is None:
        good = np.ones(X.shape, bool) if w is None else (w > 0)
        if mk is not None:
            good &= mk != 0
        mean = float(X[good].sum(dtype=np.float64)) / max(float(good.sum()), 1.0)
        Wi, Hi = init_factors(n, m, k, mean, seed, dt)
        W0 = Wi if W0 is None else W0
        H0 = Hi if H0 is None else H0
    W = np.array(to_np(W0), dtype=dt, order="C", copy=True)
    H0 = np.array(to_np(H0), dtype=dt, order="C", copy=True)
    _check_factors(W, H0, (), n, m, k)
    if (W < 0).any() or (H0 < 0).any():
        raise ValueError("W0 and H0 must be non-negative")
    if len(set(int(d) for d in devices)) != len(devices):
        raise ValueError("devices must name each GPU at most once (ranks that wait on each other cannot share a GPU; "
                         "solve_virtual_shards emulates that case)")
    grp = LocalGroup(devices)
    g = len(devices)
    vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)

    def rank_body(c):
        lo, hi = shard_rows(n, g, c.rank)
        Hr = H0.copy()
        hist = np.full(max_iter + 1, np.nan)
        n_iter, up, down = C.c_int32(0), C.c_int64(0), C.c_int64(0)
        sl = lambda a: None if a is None else a[lo:hi]
        rc = lib.cnmf_nmf_host(_DT[dt], hi - lo, m, k, vp(X[lo:hi]), vp(sl(w)), vp(sl(mk)), vp(W[lo:hi]), vp(Hr), vp(hist),
                               max_iter, tol, float(eps) if eps else 0.0, upd, c.handle if g > 1 else None,
                               C.byref(n_iter), C.byref(up), C.byref(down))
        text = lib.cnmf_last_error().decode("utf-8", "replace") if rc else ""
        ph = (C.c_double * 6)()
        lib.cnmf_host_timing(ph, 6)
        return rc, text, Hr, hist[:n_iter.value + 1], int(up.value), int(down.value), list(ph)
    try:
        t_call = time.perf_counter()
        res = grp.run(rank_body)
    finally:
        grp.close()
    for rc, text, *_ in res:
        if rc == _lib.ERR_NONFINITE:
            raise RuntimeError("chi2 became non-finite: " + text)
        if rc == _lib.ERR_DATA:
            raise ValueError(text)
        if rc:
            raise _lib.CnmfError(rc, text)
    _, _, H, hist, _, _, ph = res[0]
    if not np.all(np.isfinite(hist)):
        raise RuntimeError("chi2 became non-finite")
    nmf.last_transfer = {"h2d_bytes": sum(r[4] for r in res), "d2h_bytes": sum(r[5] for r in res)}
    nmf.last_timing = {"host_checks_and_copies_s": t_call - t_enter, "alloc_s": ph[0], "upload_prepare_s": ph[1],
                       "plan_s": ph[2], "solve_s": ph[3], "download_s": ph[4], "library_call_s": ph[5],
                       "upload_prepare_s_per_rank": [r[6][1] for r in res]}
    if is_torch:
        torch = _torch()
        return torch.from_numpy(W), torch.from_numpy(H), torch.from_numpy(hist)
    return W, H, hist


def solve_virtual_shards(X, k, world, *, weights=None, mask=None, W0, H0, max_iter=100, tol=0.0, dtype=None, eps=None,
                         update="both", device=None):
    """The row-sharded schedule with all `world` ranks on ONE GPU (cnmf_solve_group): every rank has its own
    plan, operands, H copy and history, and the per-iteration exchange of all ranks is one cooperative launch
    of the kernel body the multi-GPU path runs.  This is how a one-GPU box exercises the exchange arithmetic;
    it is not a fast path.  NumPy in; returns (W, [H of every rank], [history of every rank])."""
    torch = _require_cuda()
    lib = _lib.load()
    dt = _resolve_dtype(np.asarray(X).dtype, dtype)
    tdt = torch.float32 if dt == np.float32 else torch.float64
    dev = torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
    X = np.ascontiguousarray(X, dtype=dt)
    n, m = X.shape
    world = int(world)
    if world < 1 or world > n:
        raise ValueError("world must be in 1..n_rows")
    t = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    with torch.cuda.device(dev):
        Xp, Vp = prepare_device(t(X), t(None if weights is None else np.asarray(weights, dtype=dt)),
                                t(None if mask is None else np.asarray(mask)), dt)
        ld = Xp.shape[-1]
        grp = LocalGroup([dev.index] * world)
        plans, keep = [], []
        try:
            arr = lambda: (C.c_void_p * world)()
            aX, aV, aW, aH, ah, ap = arr(), arr(), arr(), arr(), arr(), arr()
            Ws, Hs, hs = [], [], []
            for r, c in enumerate(grp.ranks):
                lo, hi = shard_rows(n, world, r)
                plan = Plan(hi - lo, m, k, dt, ld=ld, comm=c, eps=eps)
                plans.append(plan)
                Wr = t(np.asarray(W0, dtype=dt)[lo:hi]).contiguous().clone()
                Hr = t(np.asarray(H0, dtype=dt)).contiguous().clone()
                hr = torch.full((max_iter + 1,), float("nan"), dtype=torch.float64, device=dev)
                Xr, Vr = Xp[lo:hi], Vp[lo:hi]
                keep.append((Xr, Vr))
                Ws.append(Wr), Hs.append(Hr), hs.append(hr)
                aX[r], aV[r], aW[r], aH[r], ah[r], ap[r] = (Xr.data_ptr(), Vr.data_ptr(), Wr.data_ptr(), Hr.data_ptr(),
                                                            hr.data_ptr(), plan.handle.value)
            n_iter = (C.c_int32 * world)()
            rc = lib.cnmf_solve_group(ap, world, aX, aV, aW, aH, ah, int(max_iter), float(tol or 0.0), _UPDATE[update], n_iter,
                                      _stream_ptr(None))
            torch.cuda.synchronize()
            if rc == _lib.ERR_NONFINITE:
                raise RuntimeError("chi2 became non-finite: " + lib.cnmf_last_error().decode())
            _lib.check(rc)
            W = torch.cat(Ws).cpu().numpy()
            return W, [h.cpu().numpy() for h in Hs], [hs[r][:n_iter[r] + 1].cpu().numpy() for r in range(world)]
        finally:
            for plan in plans:
                plan.close()
            grp.close()


def _nmf_device(X, k, weights, mask, W0, H0, max_iter, tol, dt, seed, eps, upd, comm, is_torch, check_every):
    """torch input (kept on the GPU) and every batched input."""
    torch = _require_cuda()
    tdt = torch.float32 if dt == np.float32 else torch.float64
    dev_in = X.device if is_torch else None
    as_t = lambda a: None if a is None else (a if type(a).__module__.startswith("torch") else torch.from_numpy(np.ascontiguousarray(a)))
    Xt = as_t(X)
    dev = Xt.device if Xt.is_cuda else torch.device("cuda", torch.cuda.current_device())
    Xt = Xt.to(dev)
    wt, mt = as_t(weights), as_t(mask)
    with torch.cuda.device(dev):
        bad = None
        try:
            Xp, Vp = prepare_device(Xt, wt, mt, dt)
        except ValueError as exc:
            bad = exc
        if isinstance(comm, Comm) and comm.world_size > 1:
            # the verdict on the input is collective: a rank that raised alone would leave its peers
            # waiting in the first exchange
            import torch.distributed as dist
            flag = torch.tensor([0.0 if bad is None else 1.0], dtype=torch.float64, device=dev)
            dist.all_reduce(flag, group=comm.group)
            if bad is None and float(flag.item()) > 0:
                bad = ValueError("invalid input on another rank of the communicator (its shard failed validation)")
        if bad is not None:
            raise bad
        lead, (n, m) = tuple(Xt.shape[:-2]), Xt.shape[-2:]
        if W0 is None or H0 is None:
            good = Vp > 0
            if lead:
                mean = float((Xp.sum(dtype=torch.float64) / good.sum().clamp(min=1)).item())
            else:
                mean = _global_mean(float(Xp.sum(dtype=torch.float64).item()), float(good.sum().item()), comm)
            if comm is None:
                Wi, Hi = init_factors(n, m, k, mean, seed, dt, lead)
            else:
                Wi, Hi = init_factors_sharded(n, m, k, mean, seed, comm.rank, dt)
            W0 = Wi if W0 is None else W0
            H0 = Hi if H0 is None else H0
        W = as_t(W0).to(device=dev, dtype=tdt).contiguous().clone()
        H = as_t(H0).to(device=dev, dtype=tdt).contiguous().clone()
        _check_factors(W, H, lead, n, m, k)
        if bool((W < 0).any()) or bool((H < 0).any()):
            raise ValueError("W0 and H0 must be non-negative")
        plan = None
        if not lead and comm is None and os.environ.get("CNMF_ONE_CTA", "") != "0":
            # a single problem that fits a thread-block cluster's shared memory: the resident cluster kernel,
            # all iterations in one launch (the same default as the NumPy entry point, host.cu)
            try:
                plan = Plan(n, m, k, dt, ld=Xp.shape[-1], eps=eps, variant=_lib.VARIANT_RESIDENT)
            except _lib.CnmfError:
                plan = None
        if plan is not None:
            hist = torch.full((1, max_iter + 1), float("nan"), dtype=torch.float64, device=dev)
            niter = torch.zeros(1, dtype=torch.int32, device=dev)
            plan.solve_batched(Xp, Vp, W, H, hist, niter, max_iter, tol, upd)
            hist = hist[0, :int(niter.item()) + 1]
            plan.close()
        elif not lead:
            plan = Plan(n, m, k, dt, ld=Xp.shape[-1], comm=comm, eps=eps)
            hist = torch.full((max_iter + 1,), float("nan"), dtype=torch.float64, device=dev)
            n_iter = plan.solve(Xp, Vp, W, H, hist, max_iter, tol, upd, check_every)
            hist = hist[:n_iter + 1]
            plan.close()
        else:
            b = int(np.prod(lead))
            plan = Plan(n, m, k, dt, ld=Xp.shape[-1], batch=b, eps=eps)
            hist = torch.full((b, max_iter + 1), float("nan"), dtype=torch.float64, device=dev)
            niter = torch.zeros(b, dtype=torch.int32, device=dev)
            plan.solve_batched(Xp, Vp, W, H, hist, niter, max_iter, tol, upd)
            n_max = int(niter.max().item())
            # NaN doubles as the padding behind a problem's last entry, so the finite check reads each
            # problem's LAST RECORDED chi2 (a kernel whose chi2 went NaN writes it and stops there)
            last = hist.gather(1, niter.to(torch.int64).clamp(0, max_iter)[:, None])[:, 0]
            bad = torch.nonzero(~torch.isfinite(last))[:, 0]
            if bad.numel():
                plan.close()
                raise RuntimeError("chi2 became non-finite in problem(s) %s of the batch" % bad[:8].tolist())
            hist = hist[:, :n_max + 1].reshape(lead + (n_max + 1,))
            plan.close()
        torch.cuda.current_stream().synchronize()
        finite = torch.isfinite(hist) if not lead else (torch.isfinite(hist) | torch.isnan(hist))
        if not bool(finite.all()) or (not lead and bool(torch.isnan(hist).any())):
            raise RuntimeError("chi2 became non-finite")
    if is_torch:
        return W.to(dev_in), H.to(dev_in), hist.to(dev_in)
    return W.cpu().numpy(), H.cpu().numpy(), hist.cpu().numpy()
