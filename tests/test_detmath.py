"""Deterministic fp64 math (signals_b200/csrc/sgb_math.h) against 50-digit mpmath -- CPU only."""
import math

import mpmath as mp
import numpy as np
import pytest

mp.mp.dps = 50


def _ulps(got, ref_fn, xs):
    worst = 0.0
    for x, g in zip(xs, got):
        t = ref_fn(mp.mpf(float(x)))
        tf = float(t)
        u = float(abs(mp.mpf(float(g)) - t)) / (np.spacing(abs(tf)) if tf != 0 else 5e-324)
        worst = max(worst, u)
    return worst


@pytest.mark.parametrize("fn,lo,hi,ref,tol", [
    (0, -708.0, 0.0, mp.exp, 1.01),
    (0, -40.0, 0.0, mp.exp, 1.01),
    (2, 0.0, 1.0, mp.log1p, 1.6),
])
def test_uniform_ranges(oracle, fn, lo, hi, ref, tol):
    rng = np.random.default_rng(fn + 1)
    xs = rng.uniform(lo, hi, 4000)
    got = oracle.math_eval(fn, xs)
    assert _ulps(got, ref, xs) <= tol


def test_log_all_ranges(oracle):
    rng = np.random.default_rng(3)
    for xs in (np.exp(rng.uniform(-700, 700, 3000)), rng.uniform(0.5, 2, 3000), 1 + rng.uniform(-0.07, 0.07, 3000)):
        assert _ulps(oracle.math_eval(1, xs), mp.log, xs) <= 1.3


def test_log1p_general(oracle):
    rng = np.random.default_rng(4)
    xs = np.concatenate((rng.uniform(-0.999999, 1, 3000), -rng.uniform(0, 1, 2000), [-0.98, -0.02, -0.5]))
    assert _ulps(oracle.math_eval(3, xs), mp.log1p, xs) <= 1.6


def test_log1p_small_arguments(oracle):
    rng = np.random.default_rng(5)
    xs = np.exp(-rng.uniform(0, 700, 3000))
    assert _ulps(oracle.math_eval(2, xs), mp.log1p, xs) <= 1.5


def test_lgamma_absolute_accuracy(oracle):
    rng = np.random.default_rng(6)
    xs = np.concatenate((rng.uniform(1e-6, 10, 3000), np.exp(rng.uniform(math.log(10), math.log(2e4), 3000))))
    got = oracle.math_eval(4, xs)
    for x, g in zip(xs, got):
        t = mp.loggamma(mp.mpf(float(x)))
        err = float(abs(mp.mpf(float(g)) - t))
        assert err <= 1e-14 + 6e-16 * abs(float(t)), (x, g, float(t))


def test_logfact_table_exact(oracle):
    ks = np.arange(0, 2048, dtype=np.float64)
    got = oracle.math_eval(6, ks)
    for k in (0, 1, 2, 3, 10, 170, 171, 1000, 2047):
        assert got[k] == float(mp.loggamma(mp.mpf(k + 1))), k
    big = oracle.math_eval(6, np.array([2048.0, 5000.0, 1e6]))
    for k, g in zip((2048, 5000, 10 ** 6), big):
        t = float(mp.loggamma(mp.mpf(k + 1)))
        assert abs(g - t) <= 4e-16 * t


def test_special_values(oracle):
    inf, nan = math.inf, math.nan
    assert oracle.math_eval(0, [-inf])[0] == 0.0
    assert oracle.math_eval(0, [-800.0])[0] == 0.0
    assert math.isnan(oracle.math_eval(0, [nan])[0])
    assert oracle.math_eval(0, [0.0])[0] == 1.0
    assert oracle.math_eval(1, [0.0])[0] == -inf
    assert oracle.math_eval(1, [1.0])[0] == 0.0
    assert math.isnan(oracle.math_eval(1, [-1.0])[0])
    assert oracle.math_eval(1, [inf])[0] == inf
    assert oracle.math_eval(2, [0.0])[0] == 0.0
    assert oracle.math_eval(2, [1.0])[0] == math.log(2.0)
    assert oracle.math_eval(4, [0.0])[0] == inf
    assert abs(oracle.math_eval(4, [1.0])[0]) < 1e-14 and abs(oracle.math_eval(4, [2.0])[0]) < 1e-14


def test_logspace_add_semantics(oracle):
    """src/code.cpp:5-8: same IEEE behaviour in both flavours"""
    inf, nan = math.inf, math.nan
    for m in (oracle.MATH_LIBM, oracle.MATH_DET):
        assert oracle.logspace_add(-inf, -27.6, m) == -27.6
        assert oracle.logspace_add(-3.0, -inf, m) == -3.0
        assert math.isnan(oracle.logspace_add(-inf, -inf, m))
        assert math.isnan(oracle.logspace_add(nan, -1.0, m))
        assert abs(oracle.logspace_add(0.0, 0.0, m) - math.log(2.0)) < 1e-15
        assert abs(oracle.logspace_add(-3.0, -27.631021115928547, m) - math.log(math.exp(-3.0) + 1e-12)) < 1e-15
    rng = np.random.default_rng(8)
    a = -rng.uniform(0, 600, 5000); b = -rng.uniform(0, 60, 5000)
    det = oracle.math_eval(5, a, b, oracle.MATH_DET); lm = oracle.math_eval(5, a, b, oracle.MATH_LIBM)
    assert np.max(np.abs(det - lm) / np.abs(lm)) < 1e-15
