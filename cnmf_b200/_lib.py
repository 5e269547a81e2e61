"""ctypes binding of libcnmf_b200.so (include/cnmf_b200.h).

There is no CPU path: if the shared library is missing, or a call needs a GPU
and none is present, this raises -- it never falls back to NumPy.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CNMF_LIB") or os.path.join(HERE, "libcnmf_b200.so")  # CNMF_LIB: diagnostics build

F32, F64 = 0, 1
MAX_K = 16
UPDATE_W, UPDATE_H, UPDATE_BOTH = 1, 2, 3
VARIANT_ONE_CTA, VARIANT_RESIDENT = 1, 2  # cnmf_desc_t.variant
SOLVE_RESUME, SOLVE_TIME_KERNELS, SOLVE_PREPARE = 4, 8, 16
ERR_ARG, ERR_ALIGN, ERR_UNSUPPORTED, ERR_NODEVICE, ERR_NCCL, ERR_DATA, ERR_NONFINITE, ERR_NOMEM = \
    -1, -2, -3, -4, -5, -6, -7, -8
BAD_X_NEGATIVE, BAD_X_NONFINITE, BAD_W_NEGATIVE, BAD_W_NONFINITE = 1, 2, 4, 8

# every symbol include/cnmf_b200.h declares (tests check the library exports them all)
SYMBOLS = (
    "cnmf_version", "cnmf_last_error", "cnmf_padded_ld", "cnmf_plan_create", "cnmf_plan_destroy",
    "cnmf_plan_info", "cnmf_plan_dryrun", "cnmf_plan_kernel_time", "cnmf_chi2", "cnmf_solve", "cnmf_solve_batched", "cnmf_prepare", "cnmf_nmf_host",
    "cnmf_synth", "cnmf_comm_unique_id", "cnmf_comm_create", "cnmf_comm_destroy", "cnmf_comm_info",
    "cnmf_launch_count", "cnmf_host_timing", "cnmf_comm_peer_export", "cnmf_comm_peer_import", "cnmf_comm_peer_enable",
    "cnmf_comm_peer_active", "cnmf_comm_create_local", "cnmf_comm_set_timeout", "cnmf_solve_group",
    "cnmf_chi2_static",
)
CHI2_STATIC_MAX_ROWS, CHI2_STATIC_MAX_H_BYTES = 1 << 25, 49152


class Desc(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("ld", C.c_int64), ("m", C.c_int32), ("k", C.c_int32),
                ("dtype", C.c_int32), ("batch", C.c_int32), ("eps", C.c_double), ("comm", C.c_void_p),
                ("sm_limit", C.c_int32), ("variant", C.c_int32)]


class Info(C.Structure):
    _fields_ = [("grid", C.c_int32), ("threads", C.c_int32), ("smem_bytes", C.c_int32), ("stages", C.c_int32),
                ("tile_rows", C.c_int32), ("lanes_per_row", C.c_int32), ("h_streams", C.c_int32),
                ("partial_blocks", C.c_int32), ("variant", C.c_int32), ("launches_per_iter", C.c_int32),
                ("bytes_per_iter", C.c_int64), ("scratch_bytes", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class CnmfError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("libcnmf_b200 error %d: %s" % (code, text))
        self.code = code


_lib = None


def load():
    """Load the library (building nothing: run cnmf_b200/build.py or __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `python cnmf_b200/build.py` "
                          "(cnmf_b200 has no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    lib.cnmf_version.restype = C.c_int
    lib.cnmf_last_error.restype = C.c_char_p
    lib.cnmf_padded_ld.restype = i64
    lib.cnmf_padded_ld.argtypes = [i32, i32]
    lib.cnmf_plan_create.argtypes = [C.POINTER(vp), C.POINTER(Desc)]
    lib.cnmf_plan_destroy.argtypes = [vp]
    lib.cnmf_plan_info.argtypes = [vp, C.POINTER(Info)]
    lib.cnmf_plan_dryrun.argtypes = [C.POINTER(Desc), C.POINTER(Info)]
    lib.cnmf_plan_kernel_time.argtypes = [vp, C.POINTER(dbl), C.POINTER(i32)]
    lib.cnmf_chi2.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.cnmf_chi2_static.argtypes = [i32, vp, vp, vp, vp, i64, i64, i32, i32, vp, vp]
    lib.cnmf_solve.argtypes = [vp, vp, vp, vp, vp, vp, i32, dbl, i32, i32, C.POINTER(i32), vp]
    lib.cnmf_solve_batched.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, dbl, i32, vp]
    lib.cnmf_prepare.argtypes = [i32, vp, vp, vp, i64, i64, vp, vp, i64, i64, i32, vp, vp]
    lib.cnmf_nmf_host.argtypes = [i32, i64, i32, i32, vp, vp, vp, vp, vp, vp, i32, dbl, dbl, i32, vp,
                                  C.POINTER(i32), C.POINTER(i64), C.POINTER(i64)]
    lib.cnmf_synth.argtypes = [i32, vp, vp, i64, i64, i32, i64, i32, vp, C.c_uint64, i32, dbl, vp]
    lib.cnmf_comm_unique_id.argtypes = [vp]
    lib.cnmf_comm_create.argtypes = [C.POINTER(vp), vp, i32, i32]
    lib.cnmf_comm_destroy.argtypes = [vp]
    lib.cnmf_comm_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
    lib.cnmf_comm_peer_export.argtypes = [vp, vp]
    lib.cnmf_comm_peer_import.argtypes = [vp, vp]
    lib.cnmf_comm_peer_enable.argtypes = [vp, i32]
    lib.cnmf_comm_peer_active.argtypes = [vp]
    lib.cnmf_comm_create_local.argtypes = [C.POINTER(vp), i32, C.POINTER(i32)]
    lib.cnmf_comm_set_timeout.argtypes = [vp, dbl]
    lib.cnmf_solve_group.argtypes = [C.POINTER(vp), i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                     C.POINTER(vp), i32, dbl, i32, C.POINTER(i32), vp]
    lib.cnmf_launch_count.restype = i64
    lib.cnmf_host_timing.argtypes = [C.POINTER(dbl), i32]
    _lib = lib
    return lib


def check(code):
    if code != 0:
        text = load().cnmf_last_error().decode("utf-8", "replace")
        if code == ERR_DATA:
            raise ValueError(text)
        raise CnmfError(code, text)


def launch_count() -> int:
    return int(load().cnmf_launch_count())


def padded_ld(m: int, dtype_code: int) -> int:
    return int(load().cnmf_padded_ld(m, dtype_code))
