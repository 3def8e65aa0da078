"""ctypes binding of libepsgpu.so (the C ABI declared in include/epsgpu.h).

This is the stub a maintainer of the reference would write in their own host language
(cgo / CPython C-API / JNI); here it is Python because the tests and bench.py are.
The library is loaded from the package directory (built in-tree by epsilon_b200.build);
there is no CPU fallback: if the library is missing, or no B200 is present when a context
is created, the call raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libepsgpu.so")

EPS_OK = 0


class EpsError(RuntimeError):
    """Mirrors `_solve.error` of the reference binding (solvemodule.cc:282-289)."""

    def __init__(self, code: int, msg: str):
        super().__init__("libepsgpu error %d: %s" % (code, msg))
        self.code = code


class Params(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32),
        ("rho", C.c_double),
        ("abs_tol", C.c_double),
        ("rel_tol", C.c_double),
        ("epoch_iterations", C.c_int32),
        ("warm_start", C.c_int32),
        ("kkt_mode", C.c_int32),
        ("cg_tol", C.c_double),
        ("cg_max_iterations", C.c_int32),
        ("use_graph", C.c_int32),
        ("profile", C.c_int32),
        ("gram_max_n", C.c_int64),
        ("solver", C.c_int32),
    ]


class Status(C.Structure):
    _fields_ = [
        ("state", C.c_int32),
        ("num_iterations", C.c_int32),
        ("r_norm", C.c_double),
        ("s_norm", C.c_double),
        ("epsilon_primal", C.c_double),
        ("epsilon_dual", C.c_double),
        ("iterations_run", C.c_int32),
        ("m", C.c_int32),
        ("n", C.c_int32),
        ("N", C.c_int32),
        ("init_ms", C.c_double),
        ("loop_ms", C.c_double),
        ("cg_iterations", C.c_int64),
        ("cg_unconverged", C.c_int64),
    ]

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_}


TRACE_FN = C.CFUNCTYPE(C.c_int, C.POINTER(Status), C.c_void_p)

_P = C.c_void_p
_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int32)

# name -> (restype, argtypes); every symbol include/epsgpu.h declares
SIGNATURES = {
    "eps_ctx_create": (C.c_int, [C.c_int, _P, C.POINTER(_P)]),
    "eps_ctx_destroy": (None, [_P]),
    "eps_last_error": (C.c_char_p, []),
    "eps_ctx_kernel_launches": (C.c_int64, [_P]),
    "eps_ctx_bytes_allocated": (C.c_int64, [_P]),
    "eps_ctx_synchronize": (C.c_int, [_P]),
    "eps_nccl_unique_id": (C.c_int, [_P]),
    "eps_ctx_init_comm": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "eps_map_dense": (C.c_int, [_P, C.c_int, C.c_int, _DP, C.POINTER(_P)]),
    "eps_map_dense_dev": (C.c_int, [_P, C.c_int, C.c_int, _P, C.POINTER(_P)]),
    "eps_map_csc": (C.c_int, [_P, C.c_int, C.c_int, C.c_int64, _IP, _IP, _DP, C.POINTER(_P)]),
    "eps_map_diagonal": (C.c_int, [_P, C.c_int, _DP, C.POINTER(_P)]),
    "eps_map_scalar": (C.c_int, [_P, C.c_int, C.c_double, C.POINTER(_P)]),
    "eps_map_kronecker": (C.c_int, [_P, _P, _P, C.POINTER(_P)]),
    "eps_map_transpose": (C.c_int, [_P, C.POINTER(_P)]),
    "eps_map_inverse": (C.c_int, [_P, C.POINTER(_P)]),
    "eps_map_dims": (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "eps_map_apply": (C.c_int, [_P, _DP, _DP]),
    "eps_map_apply_dev": (C.c_int, [_P, _P, _P]),
    "eps_map_time": (C.c_int, [_P, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
    "eps_map_destroy": (None, [_P]),
    "eps_dense_apply_noalloc": (C.c_int, [_P, C.c_int, C.c_int, _P, _P, _P]),
    "eps_problem_create": (C.c_int, [_P, C.POINTER(_P)]),
    "eps_problem_destroy": (None, [_P]),
    "eps_problem_add_variable": (C.c_int, [_P, C.c_char_p, C.c_int, C.POINTER(C.c_int)]),
    "eps_problem_add_constraint": (C.c_int, [_P, C.c_int, C.POINTER(C.c_int)]),
    "eps_problem_add_prox": (C.c_int, [_P, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int)]),
    "eps_problem_prox_sum_largest": (C.c_int, [_P, C.c_int, C.c_int]),
    "eps_problem_prox_soc_dims": (C.c_int, [_P, C.c_int, C.c_int, C.c_int]),
    "eps_problem_prox_mat_dims": (C.c_int, [_P, C.c_int, C.c_int, C.c_int]),
    "eps_problem_prox_zone": (C.c_int, [_P, C.c_int, C.c_double, _DP, _DP, C.c_int]),
    "eps_problem_prox_axis": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int]),
    "eps_problem_prox_add_arg": (C.c_int, [_P, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "eps_problem_prox_arg_sharded": (C.c_int, [_P, C.c_int, C.c_int, C.c_int64]),
    "eps_problem_constraint_sharded": (C.c_int, [_P, C.c_int, C.c_int64]),
    "eps_problem_add_term": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_int, _DP]),
    "eps_params_default": (None, [C.POINTER(Params)]),
    "eps_solver_create": (C.c_int, [_P, C.POINTER(Params), C.POINTER(_P)]),
    "eps_solver_destroy": (None, [_P]),
    "eps_solver_solve": (C.c_int, [_P, C.POINTER(Status), TRACE_FN, _P]),
    "eps_solver_set_constant": (C.c_int, [_P, C.c_int, C.c_int, _DP]),
    "eps_solver_copy_state": (C.c_int, [_P, _P]),
    "eps_solver_iterate": (C.c_int, [_P, C.c_int, C.POINTER(C.c_float)]),
    "eps_solver_residuals": (C.c_int, [_P, C.POINTER(Status)]),
    "eps_solver_get_variable": (C.c_int, [_P, C.c_char_p, _DP]),
    "eps_solver_profile": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int64), _DP, _DP]),
    "eps_solver_describe": (C.c_char_p, [_P]),
    "eps_prox_eval": (C.c_int, [_P, C.c_double, _DP, _DP]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen the in-tree library and set the prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EpsError(-3, "%s is not built: run `python -m epsilon_b200.build` "
                           "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int) -> None:
    if code != EPS_OK:
        raise EpsError(code, load().eps_last_error().decode("utf-8", "replace"))


def dptr(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags.c_contiguous or a.flags.f_contiguous
    return a.ctypes.data_as(_DP)


def iptr(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags.c_contiguous
    return a.ctypes.data_as(_IP)
