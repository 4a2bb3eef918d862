"""ctypes binding of ``libba_b200.so`` (C ABI in ``include/ba_b200.h``).

The shared object is built in-tree by ``__graft_entry__.build()`` / ``python -m
bundle_adjustment_b200.build``.  There is no fallback: a missing library or a machine without a
B200-class GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libba_b200.so")

ROBUST = {"none": 0, "irls": 1, "triggs": 2}
PRECOND = {"jacobi": 0, "schur_jacobi": 1, "two_level": 2, "auto": 3}
PRECOND_NAME = {v: k for k, v in PRECOND.items()}
SCHUR_PATH = {"tiles": 0, "two_pass": 1, "tiles_rows": 2}
TRIANGULATE_STATIC_MAX = 16384   # BA_TRIANGULATE_STATIC_MAX in include/ba_b200.h
STATUS_TEXT = {0: "iteration limit", 1: "gtol", 2: "ftol", 3: "xtol", 4: "ftol and xtol", 5: "damping limit"}


class BaError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"libba_b200 error {code}: {text}")
        self.code = code


class Options(C.Structure):
    _fields_ = [("max_iters", C.c_int32), ("gtol", C.c_double), ("ftol", C.c_double), ("xtol", C.c_double),
                ("huber_delta", C.c_double), ("robust_mode", C.c_int32), ("initial_lambda", C.c_double),
                ("lambda_min", C.c_double), ("lambda_max", C.c_double), ("precond", C.c_int32),
                ("pcg_max_iters", C.c_int32), ("pcg_tol", C.c_double), ("pcg_check_every", C.c_int32),
                ("verbose", C.c_int32), ("schur_path", C.c_int32), ("pcg_tail", C.c_int32),
                ("device_loop", C.c_int32), ("coarse_aggregate", C.c_int32), ("coarse_after", C.c_int32),
                ("point_linesearch", C.c_int32), ("dense_schur", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("status", C.c_int32), ("iterations", C.c_int32), ("accepted", C.c_int32), ("rejected", C.c_int32),
                ("pcg_iterations", C.c_int32), ("launches", C.c_int32), ("initial_raw_cost", C.c_double),
                ("final_raw_cost", C.c_double), ("initial_cost", C.c_double), ("final_cost", C.c_double),
                ("gradient_max", C.c_double), ("final_lambda", C.c_double), ("solve_ms", C.c_double),
                ("precond_used", C.c_int32), ("coarse_builds", C.c_int32), ("coarse_matvecs", C.c_int32),
                ("coarse_unknowns", C.c_int32), ("dense_solves", C.c_int32), ("reserved_i", C.c_int32), ("reserved", C.c_double * 5)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("reserved")}
        d["status_text"] = STATUS_TEXT.get(self.status, "?")
        d["precond_text"] = PRECOND_NAME.get(self.precond_used, "?")
        return d


_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)

# name -> (restype, argtypes); every symbol include/ba_b200.h declares
SYMBOLS = {
    "ba_default_options": (None, [C.POINTER(Options)]),
    "ba_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "ba_nccl_unique_id": (C.c_int, [C.c_char_p]),
    "ba_create_distributed": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    "ba_destroy": (C.c_int, [C.c_void_p]),
    "ba_set_problem": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32,
                                 C.c_int32, _dp, _dp, _dp, _dp, _dp, _ip, _ip, _dp]),
    "ba_set_params": (C.c_int, [C.c_void_p, _dp, _dp, _dp]),
    "ba_get_params": (C.c_int, [C.c_void_p, _dp, _dp, _dp]),
    "ba_restore_initial": (C.c_int, [C.c_void_p]),
    "ba_timer_start": (C.c_int, [C.c_void_p]),
    "ba_timer_stop": (C.c_int, [C.c_void_p, _dp]),
    "ba_fp64_peak": (C.c_int, [C.c_void_p, _dp]),
    "ba_residuals": (C.c_int, [C.c_void_p, _dp]),
    "ba_cost": (C.c_int, [C.c_void_p, C.c_double, _dp, _dp]),
    "ba_jacobian_blocks": (C.c_int, [C.c_void_p, _dp, _dp]),
    "ba_sparsity_csr": (C.c_int, [C.c_void_p, _lp, _lp, _lp]),
    "ba_normal_blocks": (C.c_int, [C.c_void_p, C.POINTER(Options), _dp, _dp, _dp, _dp]),
    "ba_schur_matvec": (C.c_int, [C.c_void_p, C.POINTER(Options), C.c_double, _dp, _dp]),
    "ba_solve": (C.c_int, [C.c_void_p, C.POINTER(Options), C.POINTER(Stats)]),
    "ba_solve_trace": (C.c_int, [C.c_void_p, C.c_int32, _dp, _dp, _dp, _ip, _ip, _dp, _ip]),
    "ba_kernel_count": (C.c_int, []),
    "ba_kernel_name": (C.c_char_p, [C.c_int]),
    "ba_kernel_stats": (C.c_int, [C.c_void_p, _lp, _dp]),
    "ba_layout_info": (C.c_int, [C.c_void_p, _lp]),
    "ba_tile_layout": (C.c_int, [C.c_void_p, _ip, _ip, _ip, _ip, _dp, _ip, _ip, _ip, _ip, _ip]),
    "ba_tile_items": (C.c_int, [C.c_void_p, _lp, _ip, _ip]),
    "ba_triangulate": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp, C.c_int32, _dp, _dp, _dp, _dp, C.POINTER(C.c_uint8)]),
    "ba_triangulate_static": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp, C.c_int32, _dp, _dp, _dp, _dp, C.POINTER(C.c_uint8)]),
    "ba_map_clear": (C.c_int, [C.c_void_p]),
    "ba_map_append_observations": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _ip, _dp]),
    "ba_map_set_points": (C.c_int, [C.c_void_p, C.c_int32, _ip, _dp]),
    "ba_map_forget_points": (C.c_int, [C.c_void_p, C.c_int32, _ip]),
    "ba_map_window": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, _ip, _dp, _dp, _ip, _ip]),
    "ba_map_window_points": (C.c_int, [C.c_void_p, _ip]),
    "ba_map_commit": (C.c_int, [C.c_void_p]),
    "ba_map_stats": (C.c_int, [C.c_void_p, _lp]),
    "ba_last_error": (C.c_char_p, []),
}
MAP_MAX_ID = 1 << 24   # dense id tables of the resident map (include/ba_b200.h)


def lib():
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise BaError(-2, f"{LIB_PATH} is missing: build it with `python -m bundle_adjustment_b200.build` "
                              "(nvcc, sm_100a).  There is no CPU fallback.")
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code: int):
    if code < 0:
        raise BaError(code, lib().ba_last_error().decode("utf-8", "replace"))
    return code


def dptr(a: np.ndarray):
    return a.ctypes.data_as(_dp)


def iptr(a: np.ndarray):
    return a.ctypes.data_as(_ip)


def lptr(a: np.ndarray):
    return a.ctypes.data_as(_lp)


def f64(a, shape=None) -> np.ndarray:
    out = np.ascontiguousarray(a, dtype=np.float64)
    return out if shape is None else out.reshape(shape)


def i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)
