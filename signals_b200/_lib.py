"""ctypes loader for libsignals_b200.so (the C ABI of include/signals_b200.h).

The shared object is built in-tree by `python -m signals_b200.build` (nvcc, sm_100a).
There is no fallback: if the library is missing or CUDA is unusable, callers get an error.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libsignals_b200.so")

MODEL_HSCN, MODEL_ASCN = 0, 1
LIK_BINOMIAL, LIK_BETABINOMIAL = 0, 1
DECODE_REFERENCE, DECODE_CHAINED = 0, 1
T_U8, T_U16, T_I32, T_F64 = 1, 2, 4, 8


class SgbParams(ctypes.Structure):
    _fields_ = [("model", ctypes.c_int32), ("likelihood", ctypes.c_int32), ("n_states", ctypes.c_int32),
                ("decode", ctypes.c_int32), ("loherror", ctypes.c_double), ("eps", ctypes.c_double),
                ("rho", ctypes.c_double), ("Abias", ctypes.c_double), ("selftransitionprob", ctypes.c_double)]


class SignalsB200Error(RuntimeError):
    pass


# every symbol include/signals_b200.h declares, with its ctypes signature
_vp, _i32, _i64, _dbl = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_double
_PP = ctypes.POINTER(SgbParams)
SIGNATURES = {
    "sgb_init": (ctypes.c_int, [ctypes.c_int]),
    "sgb_shutdown": (ctypes.c_int, []),
    "sgb_last_error": (ctypes.c_char_p, []),
    "sgb_device_count": (ctypes.c_int, []),
    "sgb_hmm_decode": (ctypes.c_int, [_PP, _i64, _vp, _vp, _vp, _i32, _i64, _vp, _vp, _vp]),
    "sgb_hmm_decode_multi": (ctypes.c_int, [_PP, _i32, _i64, _vp, _vp, _vp, _i32, _i64, _vp, _vp]),
    "sgb_hmm_decode_typed": (ctypes.c_int, [_PP, _i64, _vp, _i32, _vp, _i32, _vp, _i32, _i64, _vp, _vp, _i32, _i32]),
    "sgb_cols_upload": (ctypes.c_int, [_i64, _vp, _i32, _vp, _i32, _vp, _i32, _i64, _vp, ctypes.POINTER(_vp)]),
    "sgb_cols_decode": (ctypes.c_int, [_vp, _PP, _vp, _i32]),
    "sgb_cols_free": (ctypes.c_int, [_vp]),
    "sgb_plan_create": (ctypes.c_int, [_i64, _i64, _vp, _i32, _i32, ctypes.POINTER(_vp)]),
    "sgb_plan_destroy": (ctypes.c_int, [_vp]),
    "sgb_plan_reserve_packed": (ctypes.c_int, [_vp]),
    "sgb_plan_reserve_split": (ctypes.c_int, [_vp]),
    "sgb_plan_split_info": (ctypes.c_int, [_vp, _vp, ctypes.POINTER(_i64), ctypes.POINTER(_i64)]),
    "sgb_plan_read_timing3": (ctypes.c_int, [_vp, _i32, _vp, _vp, _vp, ctypes.POINTER(_i32)]),
    "sgb_hmm_decode_device": (ctypes.c_int, [_PP, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "sgb_plan_check": (ctypes.c_int, [_vp, _vp]),
    "sgb_plan_scratch_bytes": (_i64, [_vp]),
    "sgb_plan_last_staging": (ctypes.c_int, [_vp]),
    "sgb_plan_enable_timing": (ctypes.c_int, [_vp, _i32]),
    "sgb_plan_read_timing": (ctypes.c_int, [_vp, _i32, _vp, _vp, ctypes.POINTER(_i32)]),
    "sgb_viterbi": (ctypes.c_int, [_vp, _i64, _vp, _vp, _i64, _i32, _i32, _i32, _vp]),
    "sgb_viterbi_chained_noalloc": (ctypes.c_int, [_vp, _i64, _vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    "sgb_logspace_add": (ctypes.c_int, [_dbl, _dbl, ctypes.POINTER(_dbl)]),
    "sgb_logspace_add_vec": (ctypes.c_int, [_vp, _vp, _i64, _vp]),
    "sgb_postprocess": (ctypes.c_int, [_i64, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "sgb_pass_stats": (ctypes.c_int, [_i64, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _vp]),
    "sgb_table_order": (ctypes.c_int, [_i64, _vp, _vp, _vp, _i32, _i32, _vp, ctypes.POINTER(_i64), _vp]),
    "sgb_combine_bafcn": (ctypes.c_int, [_i64] + [_vp] * 7 + [_i32, _i32, _dbl, _i32, _i32, _i32, ctypes.POINTER(_i64)] + [_vp] * 8 +
                          [ctypes.POINTER(_i64), _vp]),
    "sgb_table_last_ms": (ctypes.c_int, [ctypes.POINTER(_dbl)]),
    "sgb_pass_stats2": (ctypes.c_int, [_i64, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "sgb_qc_per_cell": (ctypes.c_int, [_i64, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp]),
    "sgb_fix_assignments": (ctypes.c_int, [_i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _dbl, _i32, _dbl, _vp, _vp]),
    "sgb_math_eval": (ctypes.c_int, [_i32, _vp, _vp, _i64, _vp]),
    "sgb_fp64_peak": (ctypes.c_int, [_i32, ctypes.POINTER(_dbl)]),
}

_lib = None


def load():
    """dlopen the library and bind every declared symbol; raises if the .so is absent"""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("SGB_LIB", SO_PATH)       # SGB_LIB: an experimental build made by signals_b200.build --variant
    if not os.path.exists(path):
        raise SignalsB200Error(
            "libsignals_b200.so is not built (%s). Run `python -m signals_b200.build`; "
            "there is no CPU fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if a declared symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().sgb_last_error()
        raise SignalsB200Error(msg.decode("utf-8", "replace") if msg else "signals_b200 call failed")


def make_params(model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12, rho=0.0, Abias=0.0,
                selftransitionprob=0.999, decode="reference"):
    if likelihood not in ("binomial", "betabinomial"):
        raise ValueError("Likelihood model must be one of binomial or betabinomial")   # R/callHSCN.R:778-780
    if model not in ("hscn", "ascn"):
        raise ValueError("model must be 'hscn' or 'ascn'")
    if decode not in ("reference", "chained"):
        raise ValueError("decode must be 'reference' or 'chained'")
    return SgbParams(MODEL_ASCN if model == "ascn" else MODEL_HSCN,
                     LIK_BETABINOMIAL if likelihood == "betabinomial" else LIK_BINOMIAL, int(n_states),
                     DECODE_CHAINED if decode == "chained" else DECODE_REFERENCE, float(loherror), float(eps),
                     float(rho), float(Abias), float(selftransitionprob))
