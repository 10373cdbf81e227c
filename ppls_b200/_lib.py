"""ctypes binding of libppls_b200.so (the C ABI declared in include/ppls_b200.h).

Fails loudly: a missing library is an ImportError-grade RuntimeError, there is no Python
or CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libppls_b200.so")

PPLS_OK = 0
PPLS_E_INVALID, PPLS_E_CUDA, PPLS_E_NCCL, PPLS_E_OOM, PPLS_E_NONFINITE, PPLS_E_UNSUPPORTED, PPLS_E_DEGENERATE = -1, -2, -3, -4, -5, -6, -7
PPLS_MAX_R = 8
PPLS_NCCL_ID_BYTES = 128
PPLS_ORTH_SVD, PPLS_ORTH_QR = 0, 1
PPLS_F_TIMING = 1

c_double_p = C.POINTER(C.c_double)


class Opts(C.Structure):
    _fields_ = [("r", C.c_int32), ("max_steps", C.c_int32), ("orth_type", C.c_int32),
                ("check_every", C.c_int32), ("atol", C.c_double), ("flags", C.c_uint32),
                ("reserved", C.c_uint32)]


class Params(C.Structure):
    _fields_ = [("W", c_double_p), ("C", c_double_p), ("B", c_double_p), ("sigT", c_double_p),
                ("sigE", C.c_double), ("sigF", C.c_double), ("sigH", C.c_double)]


class Constraints(C.Structure):
    _fields_ = [("mask", C.c_uint32), ("reserved", C.c_uint32), ("W", c_double_p), ("C", c_double_p),
                ("B", C.c_double), ("sigE", C.c_double), ("sigF", C.c_double), ("sigH", C.c_double), ("sigT", C.c_double)]


FIX_BITS = {"W": 1, "C": 2, "B": 4, "sigE": 8, "sigF": 16, "sigH": 32, "sigT": 64}


class Timing(C.Structure):
    _fields_ = [("stream_ms", C.c_double), ("stream_launches", C.c_int64), ("loop_ms", C.c_double),
                ("kernel_launches", C.c_int64), ("passes", C.c_int64), ("bytes_per_pass", C.c_int64),
                ("tail_ms", C.c_double), ("peer_exchange", C.c_int64), ("stream_phase_ms", C.c_double), ("scalar_ms", C.c_double)]


# every exported symbol of include/ppls_b200.h: name -> (restype, argtypes)
_vp = C.c_void_p
_vpp = C.POINTER(C.c_void_p)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_ip = C.POINTER(C.c_int)
SIGNATURES = {
    "ppls_version": (C.c_char_p, []),
    "ppls_last_error": (C.c_char_p, []),
    "ppls_ctx_create": (C.c_int, [_vpp, C.c_int]),
    "ppls_nccl_unique_id": (C.c_int, [_vp, C.c_size_t]),
    "ppls_ctx_create_dist": (C.c_int, [_vpp, C.c_int, C.c_int, C.c_int, _vp, C.c_size_t]),
    "ppls_ctx_create_multi": (C.c_int, [_vpp, _ip, C.c_int]),
    "ppls_ctx_destroy": (C.c_int, [_vp]),
    "ppls_ctx_info": (C.c_int, [_vp, _ip, _ip, _ip, _ip]),
    "ppls_ctx_sync": (C.c_int, [_vp]),
    "ppls_probe": (C.c_int, [_vp, _vp, C.c_int32, c_double_p]),
    "ppls_data_from_host": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _vpp]),
    "ppls_data_synthetic": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_uint64,
                                      C.POINTER(Params), _vpp]),
    "ppls_data_from_host_cols": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _vpp]),
    "ppls_data_synthetic_cols": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_uint64,
                                           C.POINTER(Params), _vpp]),
    "ppls_data_export": (C.c_int, [_vp, _vp, C.c_int64, C.c_int64, _vp, C.c_int64, _vp, C.c_int64]),
    "ppls_data_info": (C.c_int, [_vp, _i64p, _i64p, _i64p, _i64p, c_double_p, c_double_p]),
    "ppls_data_destroy": (C.c_int, [_vp]),
    "ppls_fit_create": (C.c_int, [_vp, _vp, C.POINTER(Opts), C.POINTER(Params), _vpp]),
    "ppls_fit_iterate": (C.c_int, [_vp, C.c_int32, _i32p, _i32p]),
    "ppls_fit_result": (C.c_int, [_vp, C.POINTER(Params), _vp, _i32p, _vp, _vp, _vp]),
    "ppls_fit_timing": (C.c_int, [_vp, C.POINTER(Timing)]),
    "ppls_fit_destroy": (C.c_int, [_vp]),
    "ppls_sequential_run": (C.c_int, [_vp, _vp, C.c_int32, C.c_int32, C.c_double, _vp, C.POINTER(Params), _vp, _vp, _i32p, _vp, _i32p]),
    "ppls_sequential_run_ex": (C.c_int, [_vp, _vp, C.c_int32, C.c_int32, C.c_double, _vp, _vp, C.c_int32, C.POINTER(Params), _vp, _vp,
                                         _i32p, _vp, _i32p]),
    "ppls_initial_guess": (C.c_int, [_vp, C.c_int32, C.c_int64, C.c_int64, C.c_uint64, C.c_int32, C.POINTER(Params)]),
    "ppls_loglik": (C.c_int, [_vp, _vp, C.c_int32, C.POINTER(Params), _vp]),
    "ppls_orth": (C.c_int, [_vp, _vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, _vp, C.c_int64]),
    "ppls_orth_noalloc": (C.c_int, [_vp, _vp, C.c_int64, C.c_int32, C.c_int32, _vp]),
    "ppls_variances": (C.c_int, [_vp, _vp, C.c_int32, C.c_int32, _vp, C.c_int64, _vp, C.c_double, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ppls_data_simulc": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_uint64,
                                   C.POINTER(Params), _vp, _vp, _vp, _vp, _vpp]),
    "ppls_meta_run": (C.c_int, [_vp, _vp, C.c_int32, _vp, C.c_int32, C.c_double, C.POINTER(Params), _vp, _vp, _vp, _vp, _i32p]),
    "ppls_simult_run_batched": (C.c_int, [_vp, _vp, C.POINTER(Opts), C.c_int32, _vp, _vp, _vp, _i32p, C.POINTER(Timing)]),
    "ppls_simult_run": (C.c_int, [_vp, _vp, C.POINTER(Opts), C.POINTER(Params), C.POINTER(Params), _vp, _i32p,
                                  _vp, _vp, _vp]),
}

_lib = None


class PPLSError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"ppls_b200 error {code}: {message}")
        self.code = code


def lib() -> C.CDLL:
    """Load libppls_b200.so (built in-tree by ppls_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m ppls_b200.build` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != PPLS_OK:
        raise PPLSError(rc, lib().ppls_last_error().decode("utf-8", "replace"))
