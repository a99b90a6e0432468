"""The per-lane arithmetic of the CUDA decode kernel (signals_b200/csrc/hmm_core.cuh), compiled for the host
by g++ and compared bit-for-bit with the oracle's deterministic flavour -- CPU only.
This checks the O(S) structured-transition recursion, the reference decode rule, masks and special cases
of the code the GPU runs, before any GPU time is spent."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import cases

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostemu", "emu.cpp")
SO = os.path.join(HERE, "hostemu", "libemu.so")


@pytest.fixture(scope="module")
def emu():
    deps = [SRC] + [os.path.join(HERE, "..", "signals_b200", "csrc", f) for f in ("hmm_core.cuh", "sgb_math.h", "sgb_tables.h")]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-mfma", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-x", "c++",
                               SRC, "-o", SO])
    return ctypes.CDLL(SO)


def _emu_decode(E, d, kw, decode):
    n, x, cn, so = d
    out = np.zeros(n.size, np.uint8)
    D = ctypes.c_double
    for s in range(so.size - 1):
        a, b = int(so[s]), int(so[s + 1])
        if b == a:
            continue
        if decode in ("split", "split_block"):
            # the split schedule (k_emit + k_recur): plane-padded emissions + register-resident recursion, one step at a
            # time or in k_recur's branch-free four-bin blocks with the redo of flagged chunks
            fn = E.emu_decode_sequence_split if decode == "split" else E.emu_decode_sequence_split_block
            rc = fn(1 if kw["model"] == "ascn" else 0, 1 if kw["likelihood"] == "betabinomial" else 0,
                                             kw["S"], D(kw["loherror"]), D(kw["eps"]), D(kw["rho"]), D(kw["Abias"]),
                                             D(kw["selftransitionprob"]), ctypes.c_long(b - a),
                                             n[a:b].ctypes.data_as(ctypes.c_void_p), x[a:b].ctypes.data_as(ctypes.c_void_p),
                                             cn[a:b].ctypes.data_as(ctypes.c_void_p), out[a:b].ctypes.data_as(ctypes.c_void_p))
            assert rc == 0
            continue
        rc = E.emu_decode_sequence(1 if kw["model"] == "ascn" else 0, 1 if kw["likelihood"] == "betabinomial" else 0,
                                   kw["S"], decode, D(kw["loherror"]), D(kw["eps"]), D(kw["rho"]), D(kw["Abias"]),
                                   D(kw["selftransitionprob"]), ctypes.c_long(b - a),
                                   n[a:b].ctypes.data_as(ctypes.c_void_p), x[a:b].ctypes.data_as(ctypes.c_void_p),
                                   cn[a:b].ctypes.data_as(ctypes.c_void_p), out[a:b].ctypes.data_as(ctypes.c_void_p), None)
        assert rc == 0
    return out


@pytest.mark.parametrize("ps", cases.PARAM_SETS, ids=[p[0] for p in cases.PARAM_SETS])
def test_core_bit_exact_vs_oracle(oracle, emu, ps):
    kw = cases.param_dict(ps)
    seqs = cases.edge_sequences()
    n, x, cn, seg = cases.concat_sequences(seqs)
    d = cases.small_cohort("betabinomial" if kw["likelihood"] == "betabinomial" else "binomial", n_cells=4, seed=3)
    n = np.concatenate((n, d["n"])); x = np.concatenate((x, d["x"])); cn = np.concatenate((cn, d["cn"].astype(np.int32)))
    seg = np.concatenate((seg, seg[-1] + d["seg_off"][1:]))
    for decode, dname in ((0, "reference"), (1, "chained"), ("split", "reference"), ("split_block", "reference")):
        P = oracle.params(**kw, decode=dname)
        ref = oracle.decode_batch(n, x, cn, seg, P, oracle.MATH_DET)
        got = _emu_decode(emu, (n, x, cn, seg), kw, decode)
        assert np.array_equal(ref, got), (ps[0], decode, int((ref != got).sum()))


@pytest.mark.parametrize("seed", range(cases.N_RANDOM_CASES))
def test_core_bit_exact_on_randomised_cases(oracle, emu, seed):
    """the same randomised models / inputs the GPU parity test uses (tests/cases.py::random_case), on the CPU"""
    n, x, cn, seg, g = cases.random_case(seed)
    kw = dict(model=g["model"], likelihood=g["likelihood"], S=g["n_states"], loherror=g["loherror"], eps=g["eps"],
              rho=g["rho"], Abias=g["Abias"], selftransitionprob=g["selftransitionprob"])
    cn = cn.astype(np.int32)
    for decode, dname in ((0, "reference"), (1, "chained"), ("split", "reference"), ("split_block", "reference")):
        ref = oracle.decode_batch(n, x, cn, seg, oracle.params(**kw, decode=dname), oracle.MATH_DET)
        got = _emu_decode(emu, (n, x, cn, seg), kw, decode)
        assert np.array_equal(ref, got), (kw, decode, np.nonzero(ref != got)[0][:8])
