"""The R-side binding (rpkg/src/r_shim.c) compiled UNMODIFIED against a mock of R's C API (tests/r_mock/) and
driven from C the way `.Call` would drive it.  R itself is not installed here, so this is what stands between
"source only" and a real package build: registration table, argument coercion, error unwinding and -- on a GPU
box -- the numeric results through the shim equal the ones through the Python mirror."""
import ctypes
import math
import os
import subprocess

import numpy as np
import pytest

import cases
from signals_b200 import _lib, hmm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MOCK = os.path.join(ROOT, "tests", "r_mock")
SO = os.path.join(MOCK, "libshimtest.so")


@pytest.fixture(scope="module")
def shim():
    if not os.path.exists(_lib.SO_PATH):
        pytest.skip("libsignals_b200.so is not built")
    srcs = [os.path.join(ROOT, "rpkg", "src", "r_shim.c"), os.path.join(MOCK, "r_mock.c"), os.path.join(MOCK, "driver.c")]
    deps = srcs + [os.path.join(MOCK, "Rinternals.h"), os.path.join(ROOT, "include", "signals_b200.h"), _lib.SO_PATH]
    if not os.path.exists(SO) or any(os.path.getmtime(p) > os.path.getmtime(SO) for p in deps):
        libdir = os.path.dirname(_lib.SO_PATH)
        subprocess.check_call(["gcc", "-O1", "-Wall", "-shared", "-fPIC", "-I", MOCK, "-I", os.path.join(ROOT, "include")] + srcs +
                              ["-L", libdir, "-lsignals_b200", "-Wl,-rpath," + libdir, "-o", SO])
    lib = ctypes.CDLL(SO)
    lib.drv_registered.restype = ctypes.c_int
    lib.drv_logspace_add.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.POINTER(ctypes.c_double), ctypes.c_char_p]
    lib.drv_viterbi.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_decode_batch.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long,
                                     ctypes.c_void_p, ctypes.c_long, ctypes.c_char_p, ctypes.c_int] + [ctypes.c_double] * 5 + \
                                    [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_char_p]
    return lib


def _decode(shim, model, n, x, cn, seg, lik="binomial", maxCN=11, loh=0.02, eps=1e-12, rho=0.0, abias=0.0, stp=0.95,
            want_l=False, as_double=False):
    n = np.ascontiguousarray(n, np.int32); x = np.ascontiguousarray(x, np.int32); cn = np.ascontiguousarray(cn, np.int32)
    segd = np.ascontiguousarray(seg, np.float64)
    path = np.zeros(max(n.size, 1), np.float64)
    ll = np.zeros(max(n.size * (maxCN + 1), 1), np.float64)
    err = ctypes.create_string_buffer(1024)
    rc = shim.drv_decode_batch(model, n.ctypes.data, x.ctypes.data, cn.ctypes.data, n.size, segd.ctypes.data, segd.size - 1,
                               lik.encode(), maxCN, loh, eps, rho, abias, stp, int(want_l), int(as_double),
                               path.ctypes.data, ll.ctypes.data, err)
    return rc, err.value.decode(), path[: n.size], ll[: n.size * (maxCN + 1)]


def test_shim_compiles_and_registers_the_reference_routines(shim):
    names = ctypes.create_string_buffer(512)
    nargs = (ctypes.c_int * 8)()
    dyn = ctypes.c_int(-1)
    k = shim.drv_registered(names, 512, nargs, 8, ctypes.byref(dyn))
    got = dict(zip(names.value.decode().strip(";").split(";"), list(nargs)[:k]))
    # src/RcppExports.cpp:39-48 registers these two names with these arities; R/RcppExports.R calls them by name
    assert got["_signals_logspace_addcpp"] == 2 and got["_signals_viterbi"] == 3
    assert got["_signals_hmm_decode_batch"] == 13
    assert dyn.value == 0                         # R_useDynamicSymbols(dll, FALSE), as the reference does


def test_shim_argument_errors_unwind_like_rf_error(shim):
    n = np.array([10, 12], np.int32); seg = np.array([0, 2])
    rc, msg, _, _ = _decode(shim, 0, n, n // 2, [2, 2], seg, lik="poisson")
    assert rc == 1 and "binomial or betabinomial" in msg          # R/callHSCN.R:778-780
    out = ctypes.c_double(0.0)
    err = ctypes.create_string_buffer(1024)
    rc = shim.drv_logspace_add(0.0, 0.0, ctypes.byref(out), err)
    if rc == 0:                                   # a GPU is present
        assert abs(out.value - math.log(2.0)) < 1e-15
    else:                                         # no device: the library's error text arrives through Rf_error
        assert err.value.decode().startswith("signals_b200: ")


@pytest.mark.gpu
def test_shim_results_equal_the_python_mirror(shim):
    d = cases.small_cohort("betabinomial", n_cells=12, seed=31)
    n, x, cn, seg = d["n"], d["x"], d["cn"], d["seg_off"]
    for model, lik, rho in ((0, "binomial", 0.0), (1, "betabinomial", 0.02)):
        ref, lref = hmm.decode_arrays(n, x, cn, seg, model="ascn" if model else "hscn", likelihood=lik, rho=rho,
                                      n_states=12, loherror=0.02, eps=1e-12, selftransitionprob=0.95, want_loglik=True)
        for as_double in (False, True):
            rc, msg, path, ll = _decode(shim, model, n, x, cn, seg, lik=lik, rho=rho, want_l=True, as_double=as_double)
            assert rc == 0, msg
            assert np.array_equal(path, ref.astype(np.float64))
            assert np.array_equal(np.isinf(ll), np.isinf(lref)) and np.array_equal(ll[np.isfinite(lref)], lref[np.isfinite(lref)])
    # non-integer counts are refused before anything reaches the device (the shim's own check)
    # -- and device-side validation comes back through Rf_error
    bad = n.copy(); bad[3] = -1
    rc, msg, _, _ = _decode(shim, 0, bad, x, cn, seg)
    assert rc == 1 and msg.startswith("signals_b200: ") and "negative" in msg
    # logspace_addcpp and viterbi under their original .Call names
    out = ctypes.c_double(0.0); err = ctypes.create_string_buffer(1024)
    assert shim.drv_logspace_add(-3.0, math.log(1e-12), ctypes.byref(out), err) == 0
    assert out.value == hmm.logspace_addcpp(-3.0, math.log(1e-12))
    rng = np.random.default_rng(5)
    T, S = 40, 12
    em = np.asfortranarray(np.log(rng.random((T, S))))
    tr = np.asfortranarray(np.log(rng.dirichlet(np.ones(S), size=S)))
    obs = np.arange(1, T + 1, dtype=np.float64)
    path = np.zeros(T, np.float64)
    assert shim.drv_viterbi(em.ctypes.data, T, S, tr.ctypes.data, obs.ctypes.data, T, path.ctypes.data, err) == 0, err.value
    assert np.array_equal(path, np.asarray(hmm.viterbi(em, tr, np.arange(1, T + 1)), dtype=np.float64))
