"""ctypes binding of oracle/liboracle.so (the C restatement, oracle/oracle.c).

TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.
"""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

MATH_LIBM, MATH_DET, MATH_RMATH = 0, 1, 2
MODEL_HSCN, MODEL_ASCN = 0, 1
LIK_BINOMIAL, LIK_BETABINOMIAL = 0, 1
DECODE_REFERENCE, DECODE_CHAINED = 0, 1


class OrcParams(ctypes.Structure):
    _fields_ = [("model", ctypes.c_int32), ("likelihood", ctypes.c_int32), ("S", ctypes.c_int32),
                ("decode", ctypes.c_int32), ("loherror", ctypes.c_double), ("eps", ctypes.c_double),
                ("rho", ctypes.c_double), ("Abias", ctypes.c_double),
                ("selftransitionprob", ctypes.c_double)]


def build(force=False):
    """compile liboracle.so with the committed Makefile (gcc only)"""
    src = os.path.join(_HERE, "oracle.c")
    deps = [src, os.path.join(_HERE, "..", "signals_b200", "csrc", "sgb_math.h"),
            os.path.join(_HERE, "..", "signals_b200", "csrc", "sgb_tables.h")]
    if force or not os.path.exists(_SO) or any(
            os.path.exists(d) and os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _SO


# ---- oracle/_ref: the reference's own src/code.cpp compiled verbatim against oracle/rcpp_mock (see oracle/Makefile)
_REF_SO = os.path.join(_HERE, "_ref", "libref_code.so")
_REF_SRC = "/root/reference/src/code.cpp"
_ref = None


def build_ref(force=False):
    """compile oracle/_ref/libref_code.so when the reference tree is present (it is not on the GPU box, where the
    prebuilt file travels with the repository snapshot); returns the path or None"""
    if os.path.exists(_REF_SRC):
        deps = [_REF_SRC, os.path.join(_HERE, "ref_driver.cpp"), os.path.join(_HERE, "rcpp_mock", "Rcpp.h")]
        if force or not os.path.exists(_REF_SO) or any(os.path.getmtime(d) > os.path.getmtime(_REF_SO) for d in deps):
            subprocess.check_call(["make", "-C", _HERE, "-B", "_ref"], stdout=subprocess.DEVNULL)
    return _REF_SO if os.path.exists(_REF_SO) else None


def ref_lib():
    """the compiled reference routines, or None when neither the prebuilt file nor the reference tree exists"""
    global _ref
    if _ref is None and build_ref():
        L = ctypes.CDLL(_REF_SO)
        L.ref_logspace_addcpp.restype = ctypes.c_double
        L.ref_logspace_addcpp.argtypes = [ctypes.c_double, ctypes.c_double]
        L.ref_viterbi.restype = ctypes.c_int
        _ref = L
    return _ref


def ref_viterbi(emission_mat, transition_mat, observations):
    """the reference's viterbi (src/code.cpp:11-45) itself: R-style (bins x states) emission, 1-based observations"""
    e = np.asfortranarray(emission_mat, dtype=np.float64)
    tr = np.asfortranarray(transition_mat, dtype=np.float64)
    obs = np.ascontiguousarray(observations, dtype=np.float64)
    out = np.empty(obs.size, dtype=np.float64)
    rc = ref_lib().ref_viterbi(_p(e), int(e.shape[0]), int(tr.shape[0]), _p(tr), _p(obs), int(obs.size), _p(out))
    if rc:
        raise ValueError("ref_viterbi rc=%d" % rc)
    return out


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        L.orc_logspace_add.restype = ctypes.c_double
        L.orc_logspace_add.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.c_int]
        L.orc_selftest.restype = ctypes.c_int
        if L.orc_selftest() != 0:
            raise RuntimeError("liboracle.so was built with fp contraction on; rebuild with oracle/Makefile")
        _lib = L
    return _lib


def params(model="hscn", likelihood="binomial", S=12, loherror=0.02, eps=1e-12, rho=0.0, Abias=0.0,
           selftransitionprob=0.999, decode="reference"):
    return OrcParams(MODEL_ASCN if model == "ascn" else MODEL_HSCN,
                     LIK_BETABINOMIAL if likelihood == "betabinomial" else LIK_BINOMIAL, int(S),
                     DECODE_CHAINED if decode == "chained" else DECODE_REFERENCE, float(loherror),
                     float(eps), float(rho), float(Abias), float(selftransitionprob))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def logspace_add(a, b, math=MATH_LIBM):
    return lib().orc_logspace_add(float(a), float(b), int(math))


def math_eval(fn, x, y=None, math=MATH_DET):
    """fn: 0 exp, 1 log, 2 log1p on [0,1], 3 log1p, 4 lgamma, 5 logspace_add(x,y), 6 log(x!)"""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, dtype=np.float64)
    out = np.empty_like(x)
    lib().orc_math_eval(int(fn), _p(x), _p(y), _p(out), ctypes.c_int64(x.size), int(math))
    return out


def emission(n, x, cn, P, math=MATH_LIBM):
    """bins x states log-likelihood matrix (the `l` of HaplotypeHMM / alleleHMM)"""
    n, x, cn = _i32(n), _i32(x), _i32(cn)
    T = n.size
    out = np.empty((P.S, T), dtype=np.float64)      # column-major T x S
    lib().orc_emission(_p(n), _p(x), _p(cn), ctypes.c_int64(T), ctypes.byref(P), int(math), _p(out))
    return out.T


def transition(P, math=MATH_LIBM):
    out = np.empty((P.S, P.S), dtype=np.float64)
    lib().orc_transition(ctypes.byref(P), int(math), _p(out))
    return out.T          # [from, to]


def viterbi(emission_mat, transition_mat, observations, init_mode=0, decode=DECODE_REFERENCE, math=MATH_LIBM):
    """src/code.cpp:11-45 on an R-style (bins x states) emission matrix; returns 0-based states"""
    e = np.asfortranarray(emission_mat, dtype=np.float64)
    tr = np.asfortranarray(transition_mat, dtype=np.float64)
    obs = _i32(observations)
    T, S = obs.size, tr.shape[0]
    out = np.empty(T, dtype=np.float64)
    rc = lib().orc_viterbi(_p(e), ctypes.c_int64(e.shape[0]), _p(tr), _p(obs), ctypes.c_int64(T), int(S),
                           int(init_mode), int(decode), int(math), _p(out))
    if rc:
        raise MemoryError("orc_viterbi")
    return out


def hmm_sequence(n, x, cn, P, math=MATH_LIBM, want_l=False):
    n, x, cn = _i32(n), _i32(x), _i32(cn)
    T = n.size
    path = np.empty(T, dtype=np.float64)
    l = np.empty((P.S, T), dtype=np.float64) if want_l else None
    rc = lib().orc_hmm_sequence(_p(n), _p(x), _p(cn), ctypes.c_int64(T), ctypes.byref(P), int(math), _p(path),
                                _p(l) if want_l else None)
    if rc:
        raise MemoryError("orc_hmm_sequence")
    return (path, l.T) if want_l else path


def decode_batch(n, x, cn, seg_off, P, math=MATH_LIBM, nthreads=1):
    n, x, cn = _i32(n), _i32(x), _i32(cn)
    seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
    out = np.zeros(n.size, dtype=np.uint8)
    rc = lib().orc_decode_batch(ctypes.byref(P), int(math), ctypes.c_int64(n.size), _p(n), _p(x), _p(cn),
                                ctypes.c_int64(seg_off.size - 1), _p(seg_off), _p(out), int(nthreads))
    if rc:
        raise MemoryError("orc_decode_batch")
    return out


def postprocess(cn, path):
    cn = _i32(cn)
    path = np.ascontiguousarray(path, dtype=np.uint8)
    N = cn.size
    A = np.empty(N, np.int32); B = np.empty(N, np.int32); smin = np.empty(N, np.int32)
    loh = np.empty(N, np.uint8); phase = np.empty(N, np.uint8); sp = np.empty(N, np.uint8)
    baf = np.empty(N, np.float64)
    lib().orc_postprocess(ctypes.c_int64(N), _p(cn), _p(path), _p(A), _p(B), _p(smin), _p(loh), _p(phase), _p(sp),
                          _p(baf))
    return dict(A=A, B=B, state_min=smin, LOH=loh, phase=phase, state_phase=sp, state_BAF=baf)


def fix_assignments(state, A, B, alleleA, alleleB, seg_off, loherror, likelihood="binomial", rho=0.0, math=MATH_DET):
    """fix_assignments as written (R/callHSCN.R:1072-1286); returns (A, B) after both blocks"""
    state, A, B, alleleA, alleleB = (_i32(v) for v in (state, A, B, alleleA, alleleB))
    seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
    N = state.size
    Ao = np.zeros(N, np.int32); Bo = np.zeros(N, np.int32)
    rc = lib().orc_fix_assignments(ctypes.c_int64(N), _p(state), _p(A), _p(B), _p(alleleA), _p(alleleB),
                                   ctypes.c_int64(seg_off.size - 1), _p(seg_off), ctypes.c_double(loherror),
                                   int(likelihood == "betabinomial" and rho > 0), ctypes.c_double(rho), int(math), _p(Ao), _p(Bo))
    if rc:
        raise MemoryError("orc_fix_assignments")
    return Ao, Bo
