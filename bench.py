"""bench.py -- NMF iterations/s and achieved HBM GB/s at 10M x 301, k = 6
(BASELINE.json metric, configs[4]: rows sharded over N B200s, one exchange of
the k x M partials per iteration).

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference ...                      (CPU oracle arm)
    python bench.py --config 1..4                             (another BASELINE config as the line)

A step is ONE iteration of the weighted multiplicative-update loop (H update
from the reduced partials, then the fused kernel: W update, chi2, next
partials).  `value` times K steps with X, V, W, H resident in HBM; `e2e` is a
whole cnmf.nmf() solve from HOST arrays (upload, prepare, plan, K iterations,
download) through the C ABI.  The oracle (oracle/cnmf_oracle.py) is imported
only by the cpu_baseline / parity / reference legs.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ROWS, M, K_COMP = 10_000_000, 301, 6
METRIC = "NMF iterations/sec at 10M x 301, k=6 (weighted MU, rows sharded over N B200)"
SEED = 20261104
STREAM_W0 = 4           # synth hash stream of the bench's W0 (global row index: the start is the same for every N)

# BASELINE.json configs[0..3] (configs[4] is the headline above): name, rows, M, k, dtype, V mode, p_missing, batch
CONFIGS = {
    1: ("configs[0]: 2-component absorption spectra 1,000 x 81, k=2, fp64, V=1", 1000, 81, 2, "f64", "ones", 0.0, 1),
    2: ("configs[1]: synthetic IOP set 100,000 x 301, k=4, fp64, V = 1/sigma^2", 100_000, 301, 4, "f64", "ivar", 0.0, 1),
    3: ("configs[2]: hyperspectral scene 1,700,000 x 184, k=8, fp32, 10% masked", 1_700_000, 184, 8, "f32", "mask", 0.10, 1),
    4: ("configs[3]: 4096 problems of 512 x 81, k=3, batched, fp32, V = 1/sigma^2", 512, 81, 3, "f32", "ivar", 0.0, 4096),
}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


def workload(dtype_name, n_gpus):
    return {"workload": "configs[4]: synthetic IOP spectra 10,000,000 x 301 bands, k=6, V = 1/sigma^2 with 5% masked",
            "n_rows": N_ROWS, "m": M, "k": K_COMP, "dtype": dtype_name, "weights": "ivar_mask p=0.05",
            "sharding": "contiguous row blocks over %d GPU(s); one fp64 exchange of 2*k*ld+1 values per iteration" % n_gpus,
            "start": "W0 = U(0,1) hashed from the GLOBAL row index, H0 seeded: identical for every N",
            "l2": "inputs (24-48 GB per iteration) are far larger than the 126 MB L2; no flush needed"}


def np_dtype(name):
    return np.float64 if name == "f64" else np.float32


def w0_rows(lo, hi, k, dtype):
    """W0 rows [lo, hi): U(0,1) from the counter hash of the global element index (same for every sharding)."""
    from concurrent.futures import ThreadPoolExecutor
    from cnmf_b200 import synth
    out = np.empty((hi - lo, k), dtype=dtype)
    step = 250_000

    def block(r0):
        r1 = min(r0 + step, hi)
        idx = (np.arange(r0, r1, dtype=np.uint64)[:, None] * np.uint64(k) + np.arange(k, dtype=np.uint64)[None, :])
        out[r0 - lo:r1 - lo] = synth._uniform_pair(synth._hash(SEED, STREAM_W0, idx))[0].astype(dtype)
    with ThreadPoolExecutor(min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(block, range(lo, hi, step)))
    return out


def h0(k, m, dtype):
    return np.random.default_rng(SEED + 1).random((k, m)).astype(dtype)


class ClockSampler:
    """SM clock, power and throttle reasons DURING the timed region (B200_PROFILING.md recipe).

    NVML is polled from a thread of this process (pynvml, initialised before the timed region) so
    that no nvidia-smi child starts up inside the region; nvidia-smi is the fallback when pynvml
    is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = (("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, index, uuid=None, period=0.05):
        self.p = None
        self.h = None
        self.samples = []
        self.period = period
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid:
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)) if not str(uuid).startswith("GPU-") else str(uuid))
                except Exception:
                    h = None
            self.h = h if h is not None else pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._read()            # first call pays NVML's lazy setup, outside the timed region
            self.samples = []
        except Exception:
            self.h = None
            self.index = index

    def _read(self):
        nv = self.nv
        try:
            reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        self.samples.append((float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)),
                             nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0, int(reasons)))

    def start(self):
        if self.h is not None:
            import threading
            self.stop_flag = False

            def loop():
                while not self.stop_flag:
                    try:
                        self._read()
                    except Exception:
                        pass
                    time.sleep(self.period)
            self.t = threading.Thread(target=loop, daemon=True)
            self.t.start()
        else:
            try:
                self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                           "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                          stderr=subprocess.DEVNULL, text=True)
            except Exception:
                self.p = None
        return self

    def stop(self):
        if self.h is not None:
            self.stop_flag = True
            self.t.join(timeout=2)
            if not self.samples:
                try:
                    self._read()
                except Exception:
                    return None
            sm = [x[0] for x in self.samples]
            mask = 0
            for x in self.samples:
                mask |= x[2]
            return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": self.max_mhz, "power_w_max": float(max(x[1] for x in self.samples)),
                    "samples": len(sm), "reasons": sorted(n for n, b in self.BITS if mask & b), "source": "nvml thread"}
        if self.p is None:
            return None
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons, pw = [], [], set(), []
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons), "source": "nvidia-smi"}


# --------------------------------------------------------------------------- CPU legs
def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def cpu_sample(dtype, rows, m=M, k=K_COMP, vmode="ivar_mask", p=0.05):
    """Rows [0, rows) of the bench workload from the host restatement of the device generator
    (row blocks fanned over threads), with the bench's start (W0 hashed from the row index, seeded H0)."""
    from concurrent.futures import ThreadPoolExecutor
    from cnmf_b200 import synth
    X = np.empty((rows, m), dtype=dtype)
    V = np.empty((rows, m), dtype=dtype)
    hs = synth.basis(k, m, SEED)
    step = 25_000

    def block(r0):
        r1 = min(r0 + step, rows)
        X[r0:r1], V[r0:r1] = synth.spectra(r1 - r0, m, k, seed=SEED, vmode=vmode, p_missing=p, dtype=dtype, row0=r0, Hstar=hs)
    with ThreadPoolExecutor(os.cpu_count() or 1) as ex:
        list(ex.map(block, range(0, rows, step)))
    return X, V, w0_rows(0, rows, k, dtype), h0(k, m, dtype)


def cpu_arm(dtype, rows, steps, warmup, budget_s):
    """The oracle iteration on `rows` rows of the workload with all host cores.

    Two ways to use the cores are timed once each and the faster one runs the measurement:
      pool  -- row blocks fanned over a thread pool, BLAS held to ONE thread per call (the pool x BLAS
               oversubscription of round 1 made this arm 5-8x too slow);
      blas  -- one Python thread, BLAS threaded over all cores.
    Steps stop early when the time budget is spent (the count actually timed is reported)."""
    from concurrent.futures import ThreadPoolExecutor
    from threadpoolctl import threadpool_limits
    from oracle import cnmf_oracle as orc
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    X, V, W, H = cpu_sample(dtype, rows)
    gen_s = time.perf_counter() - t0
    eps = orc.default_eps(dtype)
    pool = ThreadPoolExecutor(cores)

    def step_pool(W, H):
        with threadpool_limits(limits=1, user_api="blas"):
            return orc.iteration_threaded(X, V, W, H, eps, pool, cores * 2)

    def step_blas(W, H):
        with threadpool_limits(limits=cores, user_api="blas"):
            return orc.iteration(X, V, W, H, eps)
    trial = {}
    state = (W.copy(), H.copy())
    for name, fn in (("pool", step_pool), ("blas", step_blas)):
        fn(W.copy(), H.copy())                       # page in / warm the threads
        t = time.perf_counter()
        fn(W.copy(), H.copy())
        trial[name] = time.perf_counter() - t
    mode = min(trial, key=trial.get)
    fn = step_pool if mode == "pool" else step_blas
    Wc, Hc = state
    hist = []
    for _ in range(max(warmup, 1)):
        Wc, Hc, c = fn(Wc, Hc)
        hist.append(c)
    t_start = time.perf_counter()
    done = 0
    for _ in range(max(steps, 1)):
        Wc, Hc, c = fn(Wc, Hc)
        hist.append(c)
        done += 1
        if time.perf_counter() - t_start > budget_s:
            break
    sec = (time.perf_counter() - t_start) / done
    pool.shutdown()
    info = {"mode": mode, "trial_s_per_step": trial, "pool_threads": cores if mode == "pool" else 1,
            "blas_threads": 1 if mode == "pool" else cores, "cores": cores, "cpu_model": cpu_model(), "rows_timed": rows,
            "scale_factor": N_ROWS / rows, "steps_timed": done, "warmup_done": max(warmup, 1), "generate_s": gen_s}
    return sec, info, (X, V, W, H, Wc, Hc, np.asarray(hist))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    dtype = np_dtype(args.dtype)
    rows = args.cpu_rows
    sec, info, _ = cpu_arm(dtype, rows, args.steps, args.warmup, args.cpu_budget)
    value = 1.0 / (sec * N_ROWS / rows)      # linear in rows: whole-workload iterations / s
    cfg = workload(args.dtype, args.gpus)
    sample = "%d of %d rows per step (%d steps after %d warm-up), timed directly, scaled x%g (cost is linear in rows)" % (
        rows, N_ROWS, info["steps_timed"], info["warmup_done"], N_ROWS / rows)
    cb = {"value": value, "unit": "iterations/s", "cores": info["cores"], "kind": "port", "sample": sample,
          "note": "NumPy oracle (the upstream snapshot holds no NMF code to time: SURVEY.md section 0)"}
    cb.update({k: info[k] for k in ("mode", "pool_threads", "blas_threads", "cpu_model", "rows_timed", "scale_factor",
                                    "steps_timed", "trial_s_per_step")})
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "iterations/s", "n_gpus": args.gpus,
            "steps": info["steps_timed"], "warmup": info["warmup_done"], "ms_per_step": 1e3 / value, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": cfg,
            "sample_ms_per_step": 1e3 * sec, "cpu_baseline": cb,
            "e2e": {"value": value, "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- other configs (one GPU)
def _operands(lib, dev, n, m, k, dtype_name, vmode, p, batch=1, seed=7):
    import torch
    from cnmf_b200 import _lib, synth
    code = _lib.F32 if dtype_name == "f32" else _lib.F64
    tdt = torch.float32 if dtype_name == "f32" else torch.float64
    ld = _lib.padded_ld(m, code)
    rows = n * batch
    X = torch.empty((rows, ld), dtype=tdt, device=dev)
    V = torch.empty_like(X)
    hs = torch.from_numpy(synth.basis(k, m, seed)).to(dev)
    _lib.check(lib.cnmf_synth(code, X.data_ptr(), V.data_ptr(), rows, 0, m, ld, k, hs.data_ptr(), seed,
                              synth.VMODES[vmode], p, None))
    rng = np.random.default_rng(seed + 1)
    dt = np_dtype(dtype_name)
    W = torch.from_numpy(rng.random((rows, k)).astype(dt)).to(dev)
    H = torch.from_numpy(rng.random((batch * k, m)).astype(dt)).to(dev)
    return X, V, W, H, ld


def _parity(got, ref):
    """max relative deviation of (W, H, chi2 trace) from the oracle: arrays as max|d| / max|ref|, the trace per entry"""
    W, H, hist = got
    Wr, Hr, hr = ref
    rel = lambda a, b: float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))) / np.max(np.abs(b)))
    return {"W": rel(W, Wr), "H": rel(H, Hr), "chi2": float(np.max(np.abs(np.asarray(hist) - hr) / hr))}


def measure_config(cfg_id, iters, peak, dev, dtype_override=None):
    """it/s, GB/s and roofline fraction of one BASELINE config on one GPU (device-resident operands, CUDA
    events around `iters` iterations), plus a reduced-size parity figure against the oracle."""
    import torch
    import cnmf
    from cnmf_b200 import Plan, _lib
    from oracle import cnmf_oracle as orc
    lib = _lib.load()
    name, n, m, k, dtype_name, vmode, p, batch = CONFIGS[cfg_id]
    dtype_name = dtype_override or dtype_name
    dt = np_dtype(dtype_name)
    X, V, W, H, ld = _operands(lib, dev, n, m, k, dtype_name, vmode, p, batch)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if batch > 1:
        hist = torch.zeros((batch, iters + 1), dtype=torch.float64, device=dev)
        niter = torch.zeros(batch, dtype=torch.int32, device=dev)
        plan = Plan(n, m, k, dt, ld=ld, batch=batch)
        plan.solve_batched(X, V, W.clone(), H.clone(), hist, niter, 3)
        torch.cuda.synchronize()
        e0.record()
        plan.solve_batched(X, V, W, H, hist, niter, iters)
        e1.record()
        torch.cuda.synchronize()
        h = hist.cpu().numpy()
        mono = bool(np.all(np.isfinite(h)) and np.all(np.diff(h, axis=1) <= 1e-5 * h[:, :-1]))
        route = "resident clusters" if plan.info["variant"] >= 700000 else "one CTA per problem (tiles re-read from L2: effective GB/s)"
    else:
        variant = _lib.VARIANT_RESIDENT if cfg_id == 1 else 0
        plan = Plan(n, m, k, dt, ld=ld, variant=variant)
        if variant:
            hist = torch.zeros((1, iters + 1), dtype=torch.float64, device=dev)
            niter = torch.zeros(1, dtype=torch.int32, device=dev)
            plan.solve_batched(X, V, W.clone(), H.clone(), hist, niter, 3)
            torch.cuda.synchronize()
            e0.record()
            plan.solve_batched(X, V, W, H, hist, niter, iters)
            e1.record()
            torch.cuda.synchronize()
            h = hist.cpu().numpy()[0]
            route = "resident cluster kernel, one launch for all iterations"
        else:
            hist = torch.zeros(iters + 16, dtype=torch.float64, device=dev)
            plan.solve(X, V, W, H, hist, 5)
            torch.cuda.synchronize()
            e0.record()
            plan.solve(X, V, W, H, hist, iters, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME)
            e1.record()
            torch.cuda.synchronize()
            h = hist.cpu().numpy()[:iters + 6]
            route = "streaming plan (fused row-streaming kernel)"
        mono = bool(np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]))
    ms = e0.elapsed_time(e1) / iters
    bytes_iter = plan.info["bytes_per_iter"]
    plan.close()
    del X, V, W, H
    # parity at a size the oracle finishes in a second: the same generator rows through cnmf.nmf
    from cnmf_b200 import synth
    pn, pb, pit = (min(n, 20_000), 1, 10) if batch == 1 else (n, 4, 10)
    Xs, Vs = synth.spectra(pn * pb, m, k, seed=7, vmode=vmode, p_missing=p, dtype=dt)
    rng = np.random.default_rng(8)
    if pb == 1:
        W0, H0 = rng.random((pn, k)).astype(dt), rng.random((k, m)).astype(dt)
        got = cnmf.nmf(Xs, k, weights=Vs, W0=W0, H0=H0, max_iter=pit, tol=0)
        par = _parity(got, orc.solve(Xs, Vs, W0, H0, pit))
    else:
        W0, H0 = rng.random((pb, pn, k)).astype(dt), rng.random((pb, k, m)).astype(dt)
        Xb, Vb = Xs.reshape(pb, pn, m), Vs.reshape(pb, pn, m)
        got = cnmf.nmf(Xb, k, weights=Vb, W0=W0, H0=H0, max_iter=pit, tol=0)
        ps = [_parity((got[0][b], got[1][b], got[2][b]), orc.solve(Xb[b], Vb[b], W0[b], H0[b], pit)) for b in range(pb)]
        par = {q: max(x[q] for x in ps) for q in ("W", "H", "chi2")}
    par.update({"rows": pn, "problems": pb, "iterations": pit, "tolerance": 1e-10 if dtype_name == "f64" else 1e-4})
    gbs = bytes_iter / ms / 1e6
    return {"config": name, "dtype": dtype_name, "ms_per_step": ms, "value": 1e3 / ms, "unit": "iterations/s", "gb_per_s": gbs,
            "roofline_frac": gbs / peak, "algorithmic_bytes_per_step": bytes_iter, "route": route, "chi2_monotone": mono,
            "parity_vs_oracle": par}


# --------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist
    from cnmf_b200 import Comm, Plan, _lib, launch_count, nmf, shard_rows, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and rank == 0:
        print("warning: --gpus %d but WORLD_SIZE %d; using WORLD_SIZE" % (args.gpus, world), file=sys.stderr)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        comm = Comm()
    lib = _lib.load()
    pk = peaks()
    peak = float(pk["hbm_gbs"]) if pk else 6650.0
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if pk else "fallback (B200_PROFILING.md)"

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def over_ranks(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(x):
        return over_ranks(x, dist.ReduceOp.MAX) if world > 1 else x

    def min_over_ranks(x):
        return over_ranks(x, dist.ReduceOp.MIN) if world > 1 else x

    if args.config != 5:
        # another BASELINE config as the line (one GPU; rank 0 only)
        if rank == 0:
            r = measure_config(args.config, max(args.steps, 10), peak, dev)
            line = {"metric": "NMF iterations/sec, " + r["config"], "value": r["value"], "unit": "iterations/s", "n_gpus": 1,
                    "steps": max(args.steps, 10), "warmup": 5, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                    "scaling": "strong", "vs_baseline": None, "dtype": r["dtype"], "data": "synthetic",
                    "config": {"workload": r["config"], "route": r["route"]},
                    "roofline": {"bound": "hbm", "achieved": r["gb_per_s"], "peak": peak, "unit": "GB/s", "frac": r["roofline_frac"],
                                 "traffic": None, "peak_source": peak_src, "note": "whole step (CUDA events around the iterations)"},
                    "parity_vs_oracle": r["parity_vs_oracle"], "gpu_launches": None}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.barrier()
            comm.close()
            dist.destroy_process_group()
        return

    def measure(dtype_name, steps, warmup, with_clocks, diag_steps=0):
        dtype = np_dtype(dtype_name)
        tdt = torch.float64 if dtype_name == "f64" else torch.float32
        code = _lib.F64 if dtype_name == "f64" else _lib.F32
        lo, hi = shard_rows(N_ROWS, world, rank)
        n = hi - lo
        ld = _lib.padded_ld(M, code)
        X = torch.empty((n, ld), dtype=tdt, device=dev)
        V = torch.empty_like(X)
        hs = torch.from_numpy(synth.basis(K_COMP, M, SEED)).to(dev)
        _lib.check(lib.cnmf_synth(code, X.data_ptr(), V.data_ptr(), n, lo, M, ld, K_COMP, hs.data_ptr(), SEED,
                                  synth.VMODES["ivar_mask"], 0.05, None))
        W0 = torch.from_numpy(w0_rows(lo, hi, K_COMP, dtype)).to(dev)
        H0 = torch.from_numpy(h0(K_COMP, M, dtype)).to(dev)
        W, H = W0.clone(), H0.clone()
        n_roof = 20
        hist = torch.full((warmup + steps + n_roof + diag_steps + 8,), float("nan"), dtype=torch.float64, device=dev)
        plan = Plan(n, M, K_COMP, dtype, ld=ld, comm=comm)
        info = plan.info
        stream = torch.cuda.Stream(device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks = None
        if with_clocks and rank == 0:
            try:
                uuid = torch.cuda.get_device_properties(local).uuid
            except Exception:
                uuid = None
            clocks = ClockSampler(local, uuid)
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            plan.solve(X, V, W, H, hist, warmup, stream=stream)                 # pass over (W0, H0) + warm-up steps
            # plan-time work of the timed solve (its CUDA graph is captured, instantiated and uploaded once per
            # plan and buffer set): outside the timed region, like the plan itself
            plan.solve(X, V, W, H, hist, steps, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME | _lib.SOLVE_PREPARE, stream=stream)
            barrier()
            if clocks is not None:
                clocks.start()
            l0 = launch_count()
            e0.record(stream)
            plan.solve(X, V, W, H, hist, steps, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME, stream=stream)
            e1.record(stream)
            barrier()
            launches = launch_count() - l0
            ms_mine = e0.elapsed_time(e1)
            ms = max_over_ranks(ms_mine)
            clk = clocks.stop() if clocks is not None else None
            # roofline pass: the fused kernel alone, CUDA events on its launching stream inside the library
            plan.solve(X, V, W, H, hist, n_roof, stream=stream,
                       update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME | _lib.SOLVE_TIME_KERNELS)
            k_ms, k_n = plan.kernel_time()
            k_ms /= max(k_n, 1)
            barrier()
            diag = None
            if diag_steps:
                # a longer untimed-for-the-headline run: tells rank skew from exchange cost (N > 1)
                e0.record(stream)
                plan.solve(X, V, W, H, hist, diag_steps, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME, stream=stream)
                e1.record(stream)
                barrier()
                d_ms = e0.elapsed_time(e1) / diag_steps
                diag = {"steps": diag_steps, "ms_per_step_max": max_over_ranks(d_ms), "ms_per_step_min": min_over_ranks(d_ms)}
        total = warmup + steps + n_roof + diag_steps
        h = hist.cpu().numpy()[:total + 1]
        ok = bool(np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]) and float(W.min()) >= 0
                  and float(H.min()) >= 0 and bool(torch.isfinite(W).all()))
        res = {"ms": ms, "launches": launches, "clk": clk, "kernel_ms": k_ms, "kernel_ms_max": max_over_ranks(k_ms),
               "kernel_ms_min": min_over_ranks(k_ms), "step_ms_min": min_over_ranks(ms_mine) / steps, "info": info, "hist": h,
               "properties_ok": ok, "n_local": n, "ld": ld, "diag": diag}
        plan.close()
        keep = (X, V, W0, H0) if dtype_name == args.dtype else None
        return res, keep

    sz = 8 if args.dtype == "f64" else 4
    main, keep = measure(args.dtype, args.steps, args.warmup, True, diag_steps=100 if world > 1 else 0)
    ms_step = main["ms"] / args.steps
    value = 1e3 / ms_step
    bytes_iter_total = N_ROWS * M * 2 * sz
    bytes_iter_local = main["n_local"] * M * 2 * sz
    achieved = bytes_iter_local / (main["kernel_ms"] * 1e-3) / 1e9
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get("%s_n%d%s" % (args.dtype, world, "_mma" if main["info"]["variant"] >= 900000 else ""))
    except Exception:
        pass
    h = main["hist"]
    # the chi2 trace against the committed one-GPU trace (same start for every N: only the summation order differs)
    parity_n1 = None
    try:
        ref = json.load(open(os.path.join(ROOT, "profiles", "chi2_trace_n1.json")))[args.dtype]
        nref = min(len(ref), len(h))
        parity_n1 = {"max_rel": float(np.max(np.abs(h[:nref] - np.asarray(ref[:nref])) / np.asarray(ref[:nref]))),
                     "entries": nref, "tolerance": 1e-10 if sz == 8 else 1e-4,
                     "source": "profiles/chi2_trace_n1.json (one-GPU run of this bench, committed)"}
    except Exception:
        pass
    line = {"metric": METRIC, "value": value, "unit": "iterations/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": workload(args.dtype, world),
            "exchange": ("none (1 GPU)" if comm is None else
                         "peer windows over NVLink, fused with the H update (exchange_finalize_kernel)" if comm.peer
                         else "ncclAllReduce + finalize_kernel"),
            "hbm_gbs_whole_job": bytes_iter_total / (ms_step * 1e-3) / 1e9,
            "roofline_frac_whole_step": bytes_iter_total / (ms_step * 1e-3) / 1e9 / (peak * world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_is": "kernel only (algorithmic bytes / fused-kernel time); whole step: roofline_frac_whole_step",
                         "traffic": traffic, "traffic_source": "committed ncu capture (profiles/traffic.json), not measured by this run",
                         "kernel": ("cnmf::mma_kernel (fp64 tensor path, DMMA m8n8k4)" if main["info"]["variant"] >= 900000
                                    else "cnmf::stream_kernel"), "kernel_ms": main["kernel_ms"],
                         "kernel_ms_min_max_over_ranks": [main["kernel_ms_min"], main["kernel_ms_max"]],
                         "algorithmic_bytes_per_launch": bytes_iter_local, "peak_source": peak_src,
                         "frac_of_8000_nominal": achieved / 8000.0,
                         "second_bound": ("FP64 pipe: 10 DMMA + 22 DFMA-class warp instructions per 64 elements = 204 pipe cycles per "
                                          "SM sub-partition (scalar form: 6k+7 = 43 DFMA per element = 172)"
                                          if main["info"]["variant"] >= 900000 else
                                          "FP%s FMA pipe: 6k+7 = 43 FMA per element" % ("64" if sz == 8 else "32"))},
            "gpu_launches": int(main["launches"]), "clocks": main["clk"], "launch": main["info"],
            "chi2": [float(h[0]), float(h[args.warmup + args.steps])], "chi2_after_steps": args.warmup + args.steps,
            "chi2_trace_head": [float(x) for x in h[:8]], "parity_vs_n1": parity_n1,
            "properties_ok": main["properties_ok"], "step_ms_min_max_over_ranks": [main["step_ms_min"], ms_step],
            "diag_long_run": main["diag"]}
    if args.dump_trace and rank == 0:
        line["chi2_trace"] = [float(x) for x in h]

    if not args.no_alt:
        alt_name = "f32" if args.dtype == "f64" else "f64"
        alt_steps = max(10, args.steps // 4)
        alt, _ = measure(alt_name, alt_steps, args.warmup, False)
        asz = 8 if alt_name == "f64" else 4
        a_ms = alt["ms"] / alt_steps
        line["alt"] = {"dtype": alt_name, "value": 1e3 / a_ms, "ms_per_step": a_ms,
                       "roofline_frac": alt["n_local"] * M * 2 * asz / (alt["kernel_ms"] * 1e-3) / 1e9 / peak,
                       "properties_ok": alt["properties_ok"]}
        if args.dump_trace and rank == 0:
            line["alt"]["chi2_trace"] = [float(x) for x in alt["hist"]]

    # ---- end to end: whole cnmf.nmf() solves from HOST arrays through the C ABI
    if not args.no_e2e:
        try:
            X, V, W0, H0 = keep
            n = X.shape[0]
            tdt = X.dtype
            Xh = torch.empty((n, M), dtype=tdt, pin_memory=True)
            Vh = torch.empty((n, M), dtype=tdt, pin_memory=True)
            Xh.copy_(X[:, :M])
            Vh.copy_(V[:, :M])
            W0h = W0.cpu().pin_memory()
            H0h = H0.cpu().numpy()
            del X, V, W0, keep
            torch.cuda.empty_cache()

            def call(steps, Xa, Va):
                barrier()
                t0 = time.perf_counter()
                _, _, hh = nmf(Xa, K_COMP, weights=Va, W0=W0h.numpy(), H0=H0h, max_iter=steps, tol=0, comm=comm)
                torch.cuda.synchronize()
                sec = max_over_ranks(time.perf_counter() - t0)
                return sec, hh, dict(nmf.last_transfer), dict(getattr(nmf, "last_timing", {}) or {})
            sec, hh, tr, ph = call(args.steps, Xh.numpy(), Vh.numpy())
            sec100, _, _, _ = call(100, Xh.numpy(), Vh.numpy())
            line["e2e"] = {"value": args.steps / sec, "unit": "iterations/s", "seconds_per_call": sec, "iterations": args.steps,
                           "h2d_bytes_per_step": tr["h2d_bytes"] / args.steps, "d2h_bytes_per_step": tr["d2h_bytes"] / args.steps,
                           "h2d_bytes_per_call": tr["h2d_bytes"], "phases_rank0": ph,
                           "upload_gbs_per_rank": tr["h2d_bytes"] / max(ph.get("upload_prepare_s", 0.0), 1e-9) / 1e9,
                           "at_100_iterations": {"seconds_per_call": sec100, "value": 100 / sec100},
                           "note": "one nmf() call from PINNED host arrays: upload + prepare + plan + iterations + download; the upload "
                                   "is a fixed cost per call, so it/s at a short step count is an amortisation figure -- "
                                   "seconds_per_call and the 100-iteration call say the rest; bytes are per rank",
                           "chi2_last": float(hh[-1])}
            if not args.no_pageable:
                # the same call from PAGEABLE arrays (what a plain NumPy user hands over)
                Xp, Vp = np.array(Xh.numpy(), copy=True), np.array(Vh.numpy(), copy=True)
                del Xh, Vh
                secp, _, _, php = call(args.steps, Xp, Vp)
                line["e2e"]["pageable"] = {"seconds_per_call": secp, "value": args.steps / secp, "iterations": args.steps,
                                           "upload_prepare_s": php.get("upload_prepare_s")}
                del Xp, Vp
        except Exception as exc:  # host memory too small for the pinned copy, etc.
            line["e2e"] = {"value": None, "unit": "iterations/s", "error": repr(exc)[:200]}

    if rank == 0 and world == 1:
        torch.cuda.empty_cache()
        if not args.no_other:
            # the other BASELINE configs on this GPU, each with a reduced-size parity figure against the oracle
            others = []
            for cid, dt_over in ((1, None), (2, None), (3, None), (4, None), (4, "f64")):
                try:
                    others.append(measure_config(cid, 40 if cid >= 3 else 200, peak, dev, dt_over))
                except Exception as exc:
                    others.append({"config": CONFIGS[cid][0], "error": repr(exc)[:200]})
            line["other_configs"] = others
        if not args.no_cpu:
            dtype = np_dtype(args.dtype)
            sec, info, st = cpu_arm(dtype, args.cpu_rows, 3, 1, args.cpu_budget)
            v = 1.0 / (sec * N_ROWS / args.cpu_rows)
            cb = {"value": v, "unit": "iterations/s", "cores": info["cores"], "kind": "port",
                  "sample": "%d of %d rows x %d iterations, scaled x%g (cost is linear in rows)" % (
                      args.cpu_rows, N_ROWS, info["steps_timed"], N_ROWS / args.cpu_rows)}
            cb.update({k: info[k] for k in ("mode", "pool_threads", "blas_threads", "cpu_model", "rows_timed", "scale_factor",
                                            "steps_timed", "trial_s_per_step")})
            line["cpu_baseline"] = cb
            # parity of the CUDA path on the very rows the CPU leg iterated: same start, same iteration count
            try:
                Xs, Vs, W0s, H0s, Wc, Hc, hc = st
                nit = len(hc)
                got = nmf(Xs, K_COMP, weights=Vs, W0=W0s, H0=H0s, max_iter=nit, tol=0)
                # the CPU leg's history has no entry for the start: compare the iterations
                par = _parity((got[0], got[1], got[2][1:]), (Wc, Hc, hc))
                par.update({"rows": int(Xs.shape[0]), "iterations": nit, "tolerance": 1e-10 if sz == 8 else 1e-4,
                            "what": "cnmf.nmf on the CPU leg's sample against the oracle iterations the leg ran"})
                line["parity_vs_oracle"] = par
            except Exception as exc:
                line["parity_vs_oracle"] = {"error": repr(exc)[:200]}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        if comm is not None:
            comm.close()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--config", type=int, default=5, choices=[1, 2, 3, 4, 5],
                    help="BASELINE.json config the line measures (5 = configs[4], the headline)")
    ap.add_argument("--cpu-rows", type=int, default=1_000_000)
    ap.add_argument("--cpu-budget", type=float, default=60.0, help="seconds of timed CPU iterations at most")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-pageable", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-alt", action="store_true")
    ap.add_argument("--no-other", action="store_true")
    ap.add_argument("--dump-trace", action="store_true", help="add the whole chi2 trace to the line (profiles/chi2_trace_n1.json)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner
    # to stdout when NCCL_DEBUG is set), so file descriptor 1 points at stderr while the run lasts and
    # the JSON line goes to the real stdout.
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
