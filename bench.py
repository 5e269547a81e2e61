"""bench.py -- NMF iterations/s and achieved HBM GB/s at 10M x 301, k = 6
(BASELINE.json metric, configs[4]: rows sharded over N B200s, one NCCL
all-reduce of the k x M partials per iteration).

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference ...                      (CPU oracle arm)

A step is ONE iteration of the weighted multiplicative-update loop (H update
from the reduced partials, then the fused kernel: W update, chi2, next
partials).  `value` times K steps with X, V, W, H resident in HBM; `e2e` is a
whole cnmf.nmf() solve from pinned HOST arrays (upload, prepare, plan, K
iterations, download) through the C ABI.  The oracle (oracle/cnmf_oracle.py)
is imported only by the cpu_baseline / reference legs.
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


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


def workload(dtype_name, n_gpus):
    return {"workload": "configs[4]: synthetic IOP spectra 10,000,000 x 301 bands, k=6, V = 1/sigma^2 with 5% masked",
            "n_rows": N_ROWS, "m": M, "k": K_COMP, "dtype": dtype_name, "weights": "ivar_mask p=0.05",
            "sharding": "contiguous row blocks over %d GPU(s); one fp64 all-reduce of 2*k*ld+1 values per iteration" % n_gpus,
            "l2": "inputs (24-48 GB per iteration) are far larger than the 126 MB L2; no flush needed"}


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
def cpu_sample(dtype, rows):
    """Rows [0, rows) of the bench workload, generated by the host restatement of the device generator."""
    from cnmf_b200 import synth
    X, V = synth.spectra(rows, M, K_COMP, seed=SEED, vmode="ivar_mask", p_missing=0.05, dtype=dtype)
    rng = np.random.default_rng(SEED + 1)
    W = rng.random((rows, K_COMP)).astype(dtype)
    H = rng.random((K_COMP, M)).astype(dtype)
    return X, V, W, H


def cpu_iterations(dtype, rows, steps, warmup):
    """Oracle iteration on `rows` rows with all host threads; returns (s per step on the sample, cores)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import cnmf_oracle as orc
    cores = os.cpu_count() or 1
    X, V, W, H = cpu_sample(dtype, rows)
    eps = orc.default_eps(dtype)
    with ThreadPoolExecutor(cores) as pool:
        for _ in range(warmup):
            W, H, _c = orc.iteration_threaded(X, V, W, H, eps, pool, cores * 2)
        t0 = time.perf_counter()
        for _ in range(steps):
            W, H, _c = orc.iteration_threaded(X, V, W, H, eps, pool, cores * 2)
        dt = (time.perf_counter() - t0) / max(steps, 1)
    return dt, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    dtype = np.float64 if args.dtype == "f64" else np.float32
    rows = args.cpu_rows
    steps, warmup = max(1, min(args.steps, 5)), max(1, min(args.warmup, 2))
    sec, cores = cpu_iterations(dtype, rows, steps, warmup)
    value = 1.0 / (sec * N_ROWS / rows)      # linear in rows: whole-workload iterations / s
    cfg = workload(args.dtype, args.gpus)
    sample = "%d of %d rows per step, timed directly, scaled x%d (cost is linear in rows)" % (rows, N_ROWS, N_ROWS // rows)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "iterations/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1e3 / value, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": "iterations/s", "cores": cores, "kind": "port", "sample": sample,
                             "note": "NumPy oracle (the upstream snapshot holds no NMF code to time: SURVEY.md section 0)"},
            "e2e": {"value": value, "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def measure(dtype_name, steps, warmup, with_clocks):
        dtype = np.float64 if dtype_name == "f64" else np.float32
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
        W0 = torch.from_numpy(np.random.default_rng([SEED, rank]).random((n, K_COMP)).astype(dtype)).to(dev)
        H0 = torch.from_numpy(np.random.default_rng(SEED + 1).random((K_COMP, M)).astype(dtype)).to(dev)
        W, H = W0.clone(), H0.clone()
        n_roof = 20
        hist = torch.full((warmup + steps + n_roof + 8,), float("nan"), dtype=torch.float64, device=dev)
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
            barrier()
            if clocks is not None:
                clocks.start()
            l0 = launch_count()
            e0.record(stream)
            plan.solve(X, V, W, H, hist, steps, update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME, stream=stream)
            e1.record(stream)
            barrier()
            launches = launch_count() - l0
            ms = max_over_ranks(e0.elapsed_time(e1))
            clk = clocks.stop() if clocks is not None else None
            # roofline pass: the fused kernel alone, CUDA events on its launching stream inside the library
            plan.solve(X, V, W, H, hist, n_roof, stream=stream,
                       update=_lib.UPDATE_BOTH | _lib.SOLVE_RESUME | _lib.SOLVE_TIME_KERNELS)
            k_ms, k_n = plan.kernel_time()
            barrier()
        h = hist.cpu().numpy()
        h = h[:warmup + steps + n_roof + 1]
        ok = bool(np.all(np.isfinite(h)) and np.all(np.diff(h) <= 1e-9 * h[:-1]) and float(W.min()) >= 0
                  and float(H.min()) >= 0 and bool(torch.isfinite(W).all()))
        res = {"ms": ms, "launches": launches, "clk": clk, "kernel_ms": k_ms / max(k_n, 1), "info": info,
               "chi2_first": float(h[0]), "chi2_last": float(h[-1]), "properties_ok": ok, "n_local": n, "ld": ld}
        plan.close()
        keep = (X, V, W0, H0) if dtype_name == args.dtype else None
        return res, keep

    pk = peaks()
    peak = float(pk["hbm_gbs"]) if pk else 6650.0
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if pk else "fallback (B200_PROFILING.md)"
    sz = 8 if args.dtype == "f64" else 4
    main, keep = measure(args.dtype, args.steps, args.warmup, True)
    ms_step = main["ms"] / args.steps
    value = 1e3 / ms_step
    bytes_iter_total = N_ROWS * M * 2 * sz
    bytes_iter_local = main["n_local"] * M * 2 * sz
    achieved = bytes_iter_local / (main["kernel_ms"] * 1e-3) / 1e9
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get("%s_n%d" % (args.dtype, world))
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
                         "traffic": traffic, "kernel": "cnmf::stream_kernel", "kernel_ms": main["kernel_ms"],
                         "algorithmic_bytes_per_launch": bytes_iter_local, "peak_source": peak_src,
                         "frac_of_8000_nominal": achieved / 8000.0,
                         "second_bound": "FP%s FMA pipe: 6k+7 = 43 FMA per element" % ("64" if sz == 8 else "32")},
            "gpu_launches": int(main["launches"]), "clocks": main["clk"], "launch": main["info"],
            "chi2": [main["chi2_first"], main["chi2_last"]], "properties_ok": main["properties_ok"]}

    if not args.no_alt:
        alt_name = "f32" if args.dtype == "f64" else "f64"
        alt, _ = measure(alt_name, max(10, args.steps // 4), args.warmup, False)
        asz = 8 if alt_name == "f64" else 4
        a_ms = alt["ms"] / max(10, args.steps // 4)
        line["alt"] = {"dtype": alt_name, "value": 1e3 / a_ms, "ms_per_step": a_ms,
                       "roofline_frac": alt["n_local"] * M * 2 * asz / (alt["kernel_ms"] * 1e-3) / 1e9 / peak,
                       "properties_ok": alt["properties_ok"]}

    # ---- end to end: one cnmf.nmf() solve from pinned host arrays through the C ABI
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
            barrier()
            t0 = time.perf_counter()
            Wn, Hn, hh = nmf(Xh.numpy(), K_COMP, weights=Vh.numpy(), W0=W0h.numpy(), H0=H0h, max_iter=args.steps,
                             tol=0, comm=comm)
            torch.cuda.synchronize()
            sec = max_over_ranks(time.perf_counter() - t0)
            tr = nmf.last_transfer
            line["e2e"] = {"value": args.steps / sec, "unit": "iterations/s", "seconds": sec, "iterations": args.steps,
                           "h2d_bytes_per_step": tr["h2d_bytes"] / args.steps, "d2h_bytes_per_step": tr["d2h_bytes"] / args.steps,
                           "phases_rank0": getattr(nmf, "last_timing", None),
                           "note": "one nmf() call from pinned host arrays: upload + prepare + plan + iterations + download; "
                                   "bytes are per rank, amortised over the iterations of the call",
                           "chi2_last": float(hh[-1])}
        except Exception as exc:  # host memory too small for the pinned copy, etc.
            line["e2e"] = {"value": None, "unit": "iterations/s", "error": repr(exc)[:200]}

    if rank == 0:
        if world == 1 and not args.no_cpu:
            dtype = np.float64 if args.dtype == "f64" else np.float32
            sec, cores = cpu_iterations(dtype, args.cpu_rows, 3, 1)
            v = 1.0 / (sec * N_ROWS / args.cpu_rows)
            line["cpu_baseline"] = {"value": v, "unit": "iterations/s", "cores": cores, "kind": "port",
                                    "sample": "%d of %d rows x 3 iterations, scaled x%d (cost is linear in rows)" % (
                                        args.cpu_rows, N_ROWS, N_ROWS // args.cpu_rows)}
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
    ap.add_argument("--cpu-rows", type=int, default=200_000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-alt", action="store_true")
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
