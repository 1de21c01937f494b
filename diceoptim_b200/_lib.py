"""ctypes binding of libdk_b200.so (the C ABI of include/dk_b200.h).

There is no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdk_b200.so")

c_double_p = C.POINTER(C.c_double)
c_int64_p = C.POINTER(C.c_int64)


class DkError(RuntimeError):
    """A dk_* entry point returned a non-zero status."""


class dk_model_desc(C.Structure):
    _fields_ = [
        ("n", C.c_int), ("d", C.c_int), ("p", C.c_int),
        ("kernel", C.c_int), ("trend", C.c_int),
        ("sd2", C.c_double),
        ("theta", c_double_p), ("shape", c_double_p),
        ("nugget", C.c_double), ("nugget_flag", C.c_int),
        ("noise_var", c_double_p),
        ("T", c_double_p), ("z", c_double_p), ("M", c_double_p), ("X", c_double_p),
        ("y", c_double_p), ("F", c_double_p), ("beta", c_double_p),
        ("device", C.c_int),
    ]


class dk_uniform_spec(C.Structure):
    _fields_ = [("seed", C.c_ulonglong), ("lower", c_double_p), ("upper", c_double_p)]


class dk_crit_spec(C.Structure):
    _fields_ = [("criterion", C.c_int), ("type", C.c_int), ("minimization", C.c_int), ("proxy", C.c_int),
                ("plugin", C.c_double), ("qnorm_beta", C.c_double), ("new_noise_var", C.c_double)]


class dk_lbfgs_options(C.Structure):
    _fields_ = [("maxit", C.c_int), ("history", C.c_int), ("pgtol", C.c_double), ("factr", C.c_double),
                ("max_backtracks", C.c_int)]


# every symbol include/dk_b200.h declares: name -> (restype, argtypes)
_I, _L, _D, _P, _V = C.c_int, C.c_int64, C.c_double, c_double_p, C.c_void_p
SIGNATURES = {
    "dk_last_error": (C.c_char_p, []),
    "dk_version": (C.c_char_p, []),
    "dk_model_create": (_I, [C.POINTER(dk_model_desc), C.POINTER(_V)]),
    "dk_model_destroy": (None, [_V]),
    "dk_model_append": (_I, [_V, _V, _D, _D]),
    "dk_model_n": (_I, [_V]),
    "dk_model_small_kernel": (_I, [_V]),
    "dk_model_device_bytes": (C.c_size_t, [_V]),
    "dk_model_launches": (C.c_longlong, [_V]),
    "dk_model_set_timing": (_I, [_V, _I]),
    "dk_model_last_kernel_ms": (_D, [_V]),
    "dk_model_last_phase_ms": (_D, [_V]),
    "dk_model_set_tile_mode": (_I, [_V, _I]),
    "dk_model_prepare": (_I, [_V, _I]),
    "dk_predict": (_I, [_V, _V, _L, _I, _V, _V]),
    "dk_predict_dev": (_I, [_V, _V, _L, _I, _V, _V, _V]),
    "dk_ei": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V, _V]),
    "dk_ei_dev": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V, _V, _V]),
    "dk_ei_grad": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V]),
    "dk_ei_grad_dev": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V, _V]),
    "dk_quantile": (_I, [_V, _V, _L, _D, _I, _V, _V]),
    "dk_quantile_dev": (_I, [_V, _V, _L, _D, _I, _V, _V, _V]),
    "dk_screen_ei": (_I, [_V, _V, _I, _L, _D, _I, _I, _I, _I, _V, _V, _V]),
    "dk_screen": (_I, [_V, _V, _I, C.POINTER(dk_uniform_spec), _L, C.POINTER(dk_crit_spec), _I, _V, _V, _V, _V]),
    "dk_multi_screen": (_I, [_V, _V, C.POINTER(dk_uniform_spec), _L, C.POINTER(dk_crit_spec), _I, _V, _V, _V]),
    "dk_polish": (_I, [_V, C.POINTER(dk_crit_spec), _I, _V, _I, _V, _V, C.POINTER(dk_lbfgs_options), _V, _V, _V]),
    "dk_sample_from_ei": (_I, [_V, _V, _L, _D, _I, _I, C.c_ulonglong, _V]),
    "dk_sample_weights": (_I, [_V, _V, _L, _D, _I, _V]),
    "dk_sample_weights_dev": (_I, [_V, _V, _L, _D, _I, _V, _V]),
    "dk_feas_dev": (_I, [_V, _V, _L, _I, _I, _D, _I, _V, _V]),
    "dk_check_distance_dev": (_I, [_V, _V, _L, _I, _I, _I, _V, _V]),
    "dk_check_predict": (_I, [_V, _I, _V, _L, _I, _I, _V]),
    "dk_feasible_improvement": (_I, [_V, _V, _I, _V, _L, _D, _I, _I, _V, _V, _V]),
    "dk_crit_sur_cst": (_I, [_V, _V, _I, _V, _L, _V, _L, _V, _D, _I, _V, _V, _V]),
    "dk_aei": (_I, [_V, _V, _L, _D, _D, _I, _V]),
    "dk_aei_dev": (_I, [_V, _V, _L, _D, _D, _I, _V, _V]),
    "dk_eqi": (_I, [_V, _V, _L, _D, _D, _D, _I, _V]),
    "dk_eqi_dev": (_I, [_V, _V, _L, _D, _D, _D, _I, _V, _V]),
    "dk_aei_grad": (_I, [_V, _V, _L, _D, _D, _I, _V]),
    "dk_aei_grad_dev": (_I, [_V, _V, _L, _D, _D, _I, _V, _V, _V]),
    "dk_eqi_grad": (_I, [_V, _V, _L, _D, _D, _D, _I, _V]),
    "dk_eqi_grad_dev": (_I, [_V, _V, _L, _D, _D, _D, _I, _V, _V, _V]),
    "dk_akg": (_I, [_V, _V, _L, _D, _I, _V]),
    "dk_akg_dev": (_I, [_V, _V, _L, _D, _I, _V, _V]),
    "dk_akg_grad": (_I, [_V, _V, _L, _D, _I, _V, _V]),
    "dk_akg_grad_dev": (_I, [_V, _V, _L, _D, _I, _V, _V, _V]),
    "dk_qei": (_I, [_V, _V, _L, _I, _D, _I, _I, _V]),
    "dk_qei_dev": (_I, [_V, _V, _L, _I, _D, _I, _I, _V, _V]),
    "dk_qei_grad": (_I, [_V, _V, _L, _I, _D, _I, _I, _V]),
    "dk_qei_grad_dev": (_I, [_V, _V, _L, _I, _D, _I, _I, _V, _V]),
    "dk_qei_value_grad_dev": (_I, [_V, _V, _L, _I, _D, _I, _I, _V, _V, _V]),
    "dk_mvn_cdf": (_I, [_I, _L, _V, _V, _V]),
    "dk_multi_create": (_I, [C.POINTER(dk_model_desc), _V, _I, C.POINTER(_V)]),
    "dk_multi_destroy": (None, [_V]),
    "dk_multi_count": (_I, [_V]),
    "dk_multi_model": (_V, [_V, _I]),
    "dk_model_clone": (_I, [_V, _I, C.POINTER(_V)]),
    "dk_multi_append": (_I, [_V, _V, _D, _D]),
    "dk_multi_predict": (_I, [_V, _V, _L, _I, _V, _V]),
    "dk_multi_ei": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V, _V]),
    "dk_multi_ei_grad": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V]),
    "dk_multi_quantile": (_I, [_V, _V, _L, _D, _I, _V, _V]),
    "dk_multi_akg": (_I, [_V, _V, _L, _D, _I, _V]),
    "dk_multi_qei": (_I, [_V, _V, _L, _I, _D, _I, _I, _V, _V]),
    "dk_multi_screen_ei": (_I, [_V, _V, _L, _D, _I, _I, _I, _I, _V, _V]),
    "dk_ei_dev_noalloc": (_I, [_V, _V, _L, _D, _I, _I, _I, _V, _V, _V, _V]),
}
DEBUG_SIGNATURES = {
    "dk_debug_plain": (_I, [_V, _V, _L, _I, _D, _D, _I, _I, _I, _V, _V, _V]),
}

_lib = None


def load():
    """Load libdk_b200.so (built by __graft_entry__.build()); raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DkError(f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in {**SIGNATURES, **DEBUG_SIGNATURES}.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().dk_last_error().decode("utf-8", "replace")
        raise DkError(f"{what} failed (status {rc}): {msg}")
