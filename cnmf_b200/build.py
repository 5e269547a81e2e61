"""Builds libcnmf_b200.so in-tree with nvcc for sm_100a (no torch C++ dependency).

One object per (dtype, k) instance of the stream kernel so the build fans out
over the host cores; objects are cached by source mtime under csrc/_obj/.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libcnmf_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
DEFAULT_G = 8   # lanes per row of role W in the product library (4 = column-split role W)
FLAGS = ["-DCNMF_DEFAULT_G=%d" % DEFAULT_G, "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
MAX_K = 16


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    out.append(os.path.join(HERE, "..", "include", "cnmf_b200.h"))
    out.append(os.path.abspath(__file__))
    return out


# tuning / diagnostics builds carry the sweep variants only for these k (every other k keeps the product
# instances: the libraries stay small enough to travel to the GPU box)
TUNE_KS = (4, 6, 8)


def _jobs():
    jobs = []
    for src in ("api.cu", "host.cu", "reduce.cu", "util.cu", "batched.cu", "resident.cu", "static_chi2.cu", "comm.cpp"):
        jobs.append((os.path.join(CSRC, src), os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o"), []))
    for tn, t in (("f32", "float"), ("f64", "double")):
        for k in range(1, MAX_K + 1):
            jobs.append((os.path.join(CSRC, "stream_inst.cu"), os.path.join(OBJ, f"stream_{tn}_k{k}.o"),
                         [f"-DCNMF_T={t}", f"-DCNMF_TN={tn}", f"-DCNMF_K={k}"] + ([] if k in TUNE_KS else ["-UCNMF_TUNABLE"])))
    for k in range(1, 9):   # fp64 tensor-path kernel (mma_kernel.cuh), k <= 8
        jobs.append((os.path.join(CSRC, "mma_inst.cu"), os.path.join(OBJ, f"mma_k{k}.o"), [f"-DCNMF_K={k}"]))
        # the fp32 tensor-path kernel (TF32 x 3, mma32_kernel.cuh) measured slower than the FFMA2 stream kernel: tuning builds only
        jobs.append((os.path.join(CSRC, "mma32_inst.cu"), os.path.join(OBJ, f"mma32_k{k}.o"), [f"-DCNMF_K={k}"]))
    return jobs


def _compile(job):
    src, obj, defs = job
    cmd = [NVCC] + FLAGS + defs + (["-x", "cu"] if src.endswith(".cpp") else []) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return obj


def build(force: bool = False, verbose: bool = False, prof: bool = False, tune: bool = False) -> str:
    """The product library by default.  tune=True: libcnmf_b200_tune.so (-DCNMF_TUNABLE), which also reads
    the CNMF_TUNE sweep knobs and carries the kernel variants only they can select (16 lanes per row,
    512 / 640-thread caps, register-resident role W).  prof=True: libcnmf_b200_prof.so (-DCNMF_PROF
    -DCNMF_TUNABLE), the tuning build plus per-role clock64 phase timers printed by CTA 0 and the
    resource-probe bits.  Select either at run time with CNMF_LIB=<path>; neither is ever shipped."""
    global OBJ, LIB, FLAGS
    base_flags = [f for f in FLAGS if f not in ("-DCNMF_PROF", "-DCNMF_TUNABLE")]
    OBJ, LIB, FLAGS = os.path.join(CSRC, "_obj"), os.path.join(HERE, "libcnmf_b200.so"), base_flags
    if prof:
        OBJ = os.path.join(CSRC, "_obj_prof")
        LIB = os.path.join(HERE, "libcnmf_b200_prof.so")
        FLAGS = base_flags + ["-DCNMF_PROF", "-DCNMF_TUNABLE"]
    elif tune:
        OBJ = os.path.join(CSRC, "_obj_tune")
        LIB = os.path.join(HERE, "libcnmf_b200_tune.so")
        FLAGS = base_flags + ["-DCNMF_TUNABLE"]
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs = _jobs()
    todo = [j for j in jobs if force or _newer(j[1], [j[0]] + hdrs)]
    if todo:
        if verbose:
            print("cnmf_b200: compiling %d objects for sm_100a" % len(todo), file=sys.stderr)
        with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
            list(ex.map(_compile, todo))
    objs = [j[1] for j in jobs]
    if todo or _newer(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-Xcompiler", "-fPIC", "-ldl", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, prof="--prof" in sys.argv, tune="--tune" in sys.argv))
