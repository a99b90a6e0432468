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
                                    [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_logspace_add_vec.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_postprocess.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_pass_stats.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_long, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_device_count.argtypes = [ctypes.POINTER(ctypes.c_int), ctypes.c_char_p]
    lib.drv_cols_upload.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_long,
                                    ctypes.c_int, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_double), ctypes.c_char_p]
    lib.drv_cols_decode.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_int, ctypes.c_char_p, ctypes.c_int] + \
                                   [ctypes.c_double] * 5 + [ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_cols_free.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
    lib.drv_cols_decode_bad_handle.argtypes = [ctypes.c_char_p]
    lib.drv_gc.restype = ctypes.c_int
    return lib


def _decode(shim, model, n, x, cn, seg, lik="binomial", maxCN=11, loh=0.02, eps=1e-12, rho=0.0, abias=0.0, stp=0.95,
            want_l=False, as_double=False, ngpus=1):
    n = np.ascontiguousarray(n, np.int32); x = np.ascontiguousarray(x, np.int32); cn = np.ascontiguousarray(cn, np.int32)
    segd = np.ascontiguousarray(seg, np.float64)
    path = np.zeros(max(n.size, 1), np.float64)
    ll = np.zeros(max(n.size * (maxCN + 1), 1), np.float64)
    err = ctypes.create_string_buffer(1024)
    rc = shim.drv_decode_batch(model, n.ctypes.data, x.ctypes.data, cn.ctypes.data, n.size, segd.ctypes.data, segd.size - 1,
                               lik.encode(), maxCN, loh, eps, rho, abias, stp, int(want_l), int(as_double), int(ngpus),
                               path.ctypes.data, ll.ctypes.data, err)
    return rc, err.value.decode(), path[: n.size], ll[: n.size * (maxCN + 1)]


def test_shim_compiles_and_registers_the_reference_routines(shim):
    names = ctypes.create_string_buffer(512)
    nargs = (ctypes.c_int * 16)()
    dyn = ctypes.c_int(-1)
    k = shim.drv_registered(names, 512, nargs, 16, ctypes.byref(dyn))
    got = dict(zip(names.value.decode().strip(";").split(";"), list(nargs)[:k]))
    # src/RcppExports.cpp:39-48 registers these two names with these arities; R/RcppExports.R calls them by name
    assert got["_signals_logspace_addcpp"] == 2 and got["_signals_viterbi"] == 3
    assert got["_signals_hmm_decode_batch"] == 14
    assert got["_signals_logspace_add_vec"] == 2 and got["_signals_postprocess"] == 2 and got["_signals_pass_stats"] == 5
    assert got["_signals_device_count"] == 0
    assert got["_signals_cols_upload"] == 4 and got["_signals_cols_decode"] == 10 and got["_signals_cols_free"] == 1
    assert dyn.value == 0                         # R_useDynamicSymbols(dll, FALSE), as the reference does


def test_shim_argument_errors_unwind_like_rf_error(shim):
    err = ctypes.create_string_buffer(1024)
    assert shim.drv_cols_decode_bad_handle(err) == 1 and "handle must come from the table upload" in err.value.decode()
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
def test_shim_results_equal_the_oracle(shim, oracle):
    """every .Call routine of the shim, driven the way R would drive it, against the CPU oracle (not against the Python
    mirror, which goes through the same library)"""
    d = cases.small_cohort("betabinomial", n_cells=12, seed=31)
    n, x, cn, seg = d["n"], d["x"], d["cn"], d["seg_off"]
    cn32 = cn.astype(np.int32)
    for model, lik, rho in ((0, "binomial", 0.0), (1, "betabinomial", 0.02)):
        P = oracle.params(model="ascn" if model else "hscn", likelihood=lik, S=12, loherror=0.02, eps=1e-12, rho=rho,
                          selftransitionprob=0.95)
        ref = oracle.decode_batch(n, x, cn32, seg, P, oracle.MATH_DET).astype(np.float64)
        for as_double in (False, True):
            # the batched path (INTSXP or REALSXP columns handed over as they are, numeric states back)
            rc, msg, path, _ = _decode(shim, model, n, x, cn, seg, lik=lik, rho=rho, as_double=as_double)
            assert rc == 0, msg
            assert np.array_equal(path, ref), (model, as_double)
            # the `l` matrix of HaplotypeHMM / alleleHMM, per sequence column-major, bit-exact vs the deterministic oracle
            rc, msg, path, ll = _decode(shim, model, n, x, cn, seg, lik=lik, rho=rho, want_l=True, as_double=as_double)
            assert rc == 0, msg
            assert np.array_equal(path, ref)
            for s_ in range(0, seg.size - 1, 37):
                a, b = int(seg[s_]), int(seg[s_ + 1])
                lm = oracle.emission(n[a:b], x[a:b], cn32[a:b], P, oracle.MATH_DET)
                blk = ll[a * 12: b * 12].reshape(12, b - a).T
                assert np.array_equal(np.isinf(blk), np.isinf(lm)) and np.array_equal(blk[np.isfinite(lm)], lm[np.isfinite(lm)])
        # ncores / options(signals.gpus) beyond what the box has is capped, not an error
        rc, msg, path, _ = _decode(shim, model, n, x, cn, seg, lik=lik, rho=rho, ngpus=64)
        assert rc == 0 and np.array_equal(path, ref), msg
    # device-side validation comes back through Rf_error; NA_integer_ is INT_MIN and named in the message
    bad = n.copy(); bad[3] = -1
    rc, msg, _, _ = _decode(shim, 0, bad, x, cn, seg)
    assert rc == 1 and msg.startswith("signals_b200: ") and "negative" in msg
    bad[3] = -2 ** 31
    rc, msg, _, _ = _decode(shim, 0, bad, x, cn, seg)
    assert rc == 1 and "NA_integer_" in msg
    # logspace_addcpp, its vector twin and viterbi
    err = ctypes.create_string_buffer(1024)
    out = ctypes.c_double(0.0)
    assert shim.drv_logspace_add(-3.0, math.log(1e-12), ctypes.byref(out), err) == 0
    assert out.value == oracle.logspace_add(-3.0, math.log(1e-12), oracle.MATH_DET)
    rng = np.random.default_rng(5)
    a_ = -rng.uniform(0, 300, 5000); b_ = -rng.uniform(0, 40, 5000)
    o_ = np.zeros(5000)
    assert shim.drv_logspace_add_vec(a_.ctypes.data, b_.ctypes.data, 5000, o_.ctypes.data, err) == 0, err.value
    assert np.array_equal(o_, oracle.math_eval(5, a_, b_, oracle.MATH_DET))
    T, S = 40, 12
    em = np.asfortranarray(np.log(rng.random((T, S))))
    tr = np.asfortranarray(np.log(rng.dirichlet(np.ones(S), size=S)))
    obs = np.arange(1, T + 1, dtype=np.float64)
    path = np.zeros(T, np.float64)
    assert shim.drv_viterbi(em.ctypes.data, T, S, tr.ctypes.data, obs.ctypes.data, T, path.ctypes.data, err) == 0, err.value
    assert np.array_equal(path, oracle.viterbi(em, tr, np.arange(1, T + 1), math=oracle.MATH_DET))
    # post-decode columns and between-pass sums
    P = oracle.params(model="hscn", S=12, selftransitionprob=0.95)
    pth = oracle.decode_batch(n, x, cn32, seg, P, oracle.MATH_DET)
    ref = oracle.postprocess(cn32, pth)
    N = n.size
    ints = np.zeros(6 * N, np.int32); baf = np.zeros(N)
    pd_ = pth.astype(np.float64)
    assert shim.drv_postprocess(cn32.ctypes.data, pd_.ctypes.data, N, ints.ctypes.data, baf.ctypes.data, err) == 0, err.value
    for k, key in enumerate(("A", "B", "state_min", "LOH", "phase", "state_phase")):
        assert np.array_equal(ints[k * N:(k + 1) * N], ref[key].astype(np.int32)), key
    assert np.array_equal(baf.view(np.uint64), ref["state_BAF"].view(np.uint64))
    sums = np.zeros(3); hist = np.zeros(144)
    assert shim.drv_pass_stats(n.ctypes.data, x.ctypes.data, cn32.ctypes.data, pd_.ctypes.data, N, 11, sums.ctypes.data,
                               hist.ctypes.data, err) == 0, err.value
    ahom = (ref["B"] == 0) & (ref["A"] > 0)
    assert sums.tolist() == [float(x[ahom].sum()), float(n[ahom].sum()), float(ahom.sum())]
    h = np.zeros((12, 12)); sel = ref["B"] > 0
    np.add.at(h, (ref["A"][sel], ref["B"][sel]), 1)
    assert np.array_equal(hist.reshape(12, 12).T, h)          # R matrix [A+1, B+1], column-major
    k = ctypes.c_int(0)
    assert shim.drv_device_count(ctypes.byref(k), err) == 0 and k.value >= 1


@pytest.mark.gpu
def test_shim_table_handle_two_passes_one_upload(shim, oracle):
    """_signals_cols_upload / _decode / _free: the workflow of callHaplotypeSpecificCN (pass 1 binomial at the default
    loherror, pass 2 at the inferred one / beta-binomial, R/callHSCN.R:883-935) on ONE upload of the three columns, INTSXP
    and REALSXP alike; every pass equals the oracle; a released handle is refused; an un-freed handle is collected by the
    finalizer."""
    d = cases.small_cohort("betabinomial", n_cells=25, seed=19)
    n, x, cn, so = d["n"], d["x"], d["cn"].astype(np.int32), d["seg_off"]
    segd = so.astype(np.float64)
    passes = [cases.PARAM_SETS[0], cases.PARAM_SETS[2], cases.PARAM_SETS[4]]
    err = ctypes.create_string_buffer(1024)
    for as_double in (0, 1):
        h = ctypes.c_void_p(); nr = ctypes.c_double(0.0)
        assert shim.drv_cols_upload(n.ctypes.data, x.ctypes.data, cn.ctypes.data, n.size, segd.ctypes.data, segd.size - 1, as_double,
                                    ctypes.byref(h), ctypes.byref(nr), err) == 0, err.value
        assert nr.value == n.size and h.value
        for ps in passes:
            pd = cases.param_dict(ps)
            ref = oracle.decode_batch(n, x, cn, so, oracle.params(**pd), oracle.MATH_DET, nthreads=4)
            path = np.zeros(n.size, np.float64)
            rc = shim.drv_cols_decode(h, nr.value, 1 if pd["model"] == "ascn" else 0, pd["likelihood"].encode(), pd["S"] - 1, pd["loherror"],
                                      pd["eps"], pd["rho"], pd["Abias"], pd["selftransitionprob"], path.ctypes.data, err)
            assert rc == 0, err.value
            assert np.array_equal(path, ref.astype(np.float64)), ps[0]
        if as_double == 0:
            assert shim.drv_cols_free(h, err) == 0
            path = np.zeros(n.size, np.float64)
            assert shim.drv_cols_decode(h, nr.value, 0, b"binomial", 11, 0.02, 1e-12, 0.0, 0.0, 0.95, path.ctypes.data, err) == 1
            assert "released" in err.value.decode()
    assert shim.drv_gc() >= 1          # the second handle was never freed: its finalizer runs at collection
    # a bad column fails the upload with the library's message
    bad = n.copy(); bad[3] = -2147483648      # NA_integer_
    h = ctypes.c_void_p(); nr = ctypes.c_double(0.0)
    assert shim.drv_cols_upload(bad.ctypes.data, x.ctypes.data, cn.ctypes.data, n.size, segd.ctypes.data, segd.size - 1, 1,
                                ctypes.byref(h), ctypes.byref(nr), err) == 1
    assert "signals_b200" in err.value.decode()
