"""The CPU oracle (oracle/oracle.c) pinned against its numpy twin, 50-digit mpmath densities,
hand-computed answers and the committed golden fixtures -- CPU only.

Parity status (see oracle/oracle.c): the reference ships no golden vectors for this path and
cannot run here, so the oracle is pinned by independent restatements instead."""
import math
import os

import numpy as np
import pytest

import cases
from oracle import oracle_np as onp

GOLD = os.path.join(os.path.dirname(__file__), "golden", "hmm_golden.npz")


def test_fp_contract_off(oracle):
    assert oracle.lib().orc_selftest() == 0


def test_hand_computed_emissions(oracle):
    # c=2, m=1 -> p = 0.5 exactly (no loherror shift); n=2, x=1 -> dbinom = log(2 * .25) = log .5
    P = oracle.params(S=3, eps=0.0)
    for m in (oracle.MATH_LIBM, oracle.MATH_DET):
        l = oracle.emission([2], [1], [2], P, m)[0]
        assert abs(l[1] - math.log(0.5)) < 1e-15
        # m=0 -> p = 0.02 ; m=2 -> p = 0.98 ; symmetric for x = n/2
        assert abs(l[0] - math.log(2 * 0.02 * 0.98)) < 1e-15 and abs(l[0] - l[2]) < 1e-15
    # eps floor: logspace_add(l, log eps)
    P = oracle.params(S=3, eps=1e-12)
    l = oracle.emission([2], [1], [2], P, oracle.MATH_LIBM)[0]
    assert abs(l[1] - math.log(0.5 + 1e-12)) < 1e-15
    # states above the total copy number are forbidden (R/callHSCN.R:80)
    P = oracle.params(S=5)
    l = oracle.emission([10], [3], [2], P, oracle.MATH_DET)[0]
    assert np.isinf(l[3]) and np.isinf(l[4]) and np.isfinite(l[:3]).all()
    # ASCN forbids m > c/2 (R/callASCN.R:81) and CN 0 kills every state (no 0 -> 1 remap)
    P = oracle.params(model="ascn", S=5)
    l = oracle.emission([10, 4], [3, 1], [4, 0], P, oracle.MATH_DET)
    assert np.isfinite(l[0, :3]).all() and np.isinf(l[0, 3:]).all() and np.isinf(l[1]).all()
    # HSCN remaps CN 0 to 1 (R/callHSCN.R:49): states 0 and 1 live
    P = oracle.params(model="hscn", S=5)
    l = oracle.emission([4], [1], [0], P, oracle.MATH_DET)[0]
    assert np.isfinite(l[:2]).all() and np.isinf(l[2:]).all()


def test_hand_computed_viterbi(oracle):
    """2 states, 3 bins, worked by hand with the reference's decode rule (src/code.cpp:35-42)"""
    e = np.log(np.array([[0.9, 0.1], [0.2, 0.8], [0.3, 0.7]]))
    tr = np.log(np.array([[0.7, 0.3], [0.3, 0.7]]))
    obs = [1, 2, 3]
    # T1[:,0] = log(.5)+e0 ; bin1: state0 max(.45*.7,.05*.3)*.2 <- from 0 ; state1 max(.45*.3,.05*.7)*.8 <- from 0
    # column 1 best = state1 (.108 vs .063) whose predecessor is 0 -> path[0] = 0
    # bin2: state0 max(.063*.7,.108*.3)*.3 <- 0 ; state1 max(.063*.3,.108*.7)*.7 <- 1 ; best state1 -> path[1] = 1
    # path[2] = path[1]
    for m in (oracle.MATH_LIBM, oracle.MATH_DET):
        assert oracle.viterbi(e, tr, obs, math=m).tolist() == [0.0, 1.0, 1.0]
    assert onp.viterbi(e, tr, obs).tolist() == [0.0, 1.0, 1.0]
    assert onp.viterbiR(e, tr, obs).tolist() == [0.0, 1.0, 1.0]
    # chained (textbook) decode on the same input: best final state 1, back 1 <- 1 <- 0
    assert oracle.viterbi(e, tr, obs, decode=oracle.DECODE_CHAINED).tolist() == [0.0, 1.0, 1.0]


def test_viterbi_edge_cases(oracle):
    tr = np.log(np.full((3, 3), 1.0 / 3))
    # T = 1 -> path[0] = 0 (T2 is zero-initialised, src/code.cpp:23,37)
    assert oracle.viterbi(np.log([[0.1, 0.8, 0.1]]), tr, [1]).tolist() == [0.0]
    # an all -Inf column: every later column is -Inf and first-max gives state 0
    e = np.log(np.array([[0.2, 0.5, 0.3], [0.0, 0.0, 0.0], [0.1, 0.1, 0.8]]))
    assert oracle.viterbi(e, tr, [1, 2, 3]).tolist() == [0.0, 0.0, 0.0]
    # exact ties: first index wins in both argmaxes
    e = np.log(np.full((4, 3), 1.0 / 3))
    assert oracle.viterbi(e, tr, [1, 2, 3, 4]).tolist() == [0.0, 0.0, 0.0, 0.0]
    # observations index the emission rows
    e = np.log(np.array([[0.9, 0.05, 0.05], [0.05, 0.9, 0.05]]))
    trs = np.log(np.array([[0.8, 0.1, 0.1], [0.1, 0.8, 0.1], [0.1, 0.1, 0.8]]))
    assert oracle.viterbi(e, trs, [2, 2, 1, 1, 2]).tolist() == onp.viterbi(e, trs, [2, 2, 1, 1, 2]).tolist()
    # last element always repeats the one before (reference rule)
    rng = np.random.default_rng(0)
    e = np.log(rng.dirichlet(np.ones(3), size=50))
    p = oracle.viterbi(e, trs, np.arange(1, 51))
    assert p[-1] == p[-2]


def test_c_oracle_matches_numpy_twin(oracle):
    rng = np.random.default_rng(21)
    for trial in range(6):
        S = [3, 12, 12, 7, 12, 12][trial]
        T = [40, 120, 200, 60, 90, 150][trial]
        e = np.log(rng.dirichlet(np.ones(S) * 0.5, size=T))
        e[rng.random((T, S)) < 0.1] = -np.inf
        stp = [0.0, 0.95, 0.999, 0.5, 1.0, 0.9][trial]
        tr = onp._transition(np.arange(S), stp)
        obs = np.arange(1, T + 1)
        assert np.array_equal(oracle.viterbi(e, tr, obs), onp.viterbi(e, tr, obs))
        assert np.array_equal(oracle.viterbi(e, tr, obs, init_mode=1), onp.viterbiR(e, tr, obs))
        assert np.array_equal(oracle.viterbi(e, tr, obs, decode=oracle.DECODE_CHAINED), onp.viterbi_chained(e, tr, obs))


@pytest.mark.parametrize("ps", cases.PARAM_SETS, ids=[p[0] for p in cases.PARAM_SETS])
def test_emission_and_path_vs_numpy_twin(oracle, ps):
    kw = cases.param_dict(ps)
    P = oracle.params(**kw)
    minor = np.arange(kw["S"])
    for name, n, x, cn in cases.edge_sequences():
        if kw["model"] == "hscn":
            r = onp.HaplotypeHMM(n, x, cn, minor, kw["loherror"], kw["selftransitionprob"], kw["eps"],
                                 kw["likelihood"], kw["rho"], kw["Abias"])
        else:
            r = onp.alleleHMM(n, x, cn, minor, kw["loherror"], kw["selftransitionprob"], kw["eps"], kw["rho"],
                              kw["likelihood"])
        for m in (oracle.MATH_LIBM, oracle.MATH_DET):
            path, l = oracle.hmm_sequence(n, x, cn, P, m, want_l=True)
            assert np.array_equal(np.isinf(l), np.isinf(r["l"])), (name, m)
            fin = np.isfinite(l)
            # 1e-9 relative; 1e-12 absolute floor for values that are ~0 (e.g. log(0.98 + 0.02) + log1p(eps))
            assert np.allclose(l[fin], r["l"][fin], rtol=1e-9, atol=1e-12), (name, m)
        # the libm flavour and the numpy twin share libm-class arithmetic: same paths unless a structural tie
        path = oracle.hmm_sequence(n, x, cn, P, oracle.MATH_LIBM)
        # (sequences built around exact ties -- mirror-image likelihoods at CN 0, n = 1, x = n/2 -- are decided
        #  by last-ulp rounding, which numpy's SIMD log/exp and glibc do not share: compare the others)
        if name in ("T1", "T2", "cn_gt_maxstate", "cn1_only", "piecewise300"):
            assert np.array_equal(path, r["minorcn"]), name


def test_densities_vs_mpmath(oracle):
    """emission log-likelihoods within 1e-9 relative (north-star tolerance); measured ~1e-13"""
    rng = np.random.default_rng(33)
    worst = {0: 0.0, 1: 0.0}
    for lik, rho in (("binomial", 0.0), ("betabinomial", 0.02)):
        P = oracle.params(likelihood=lik, rho=rho, S=12, eps=0.0, loherror=0.02)
        for _ in range(150):
            c = int(rng.integers(1, 12))
            n = int(max(1, round(c * rng.gamma(1.76, 1 / 0.09))))
            b = int(rng.integers(0, c + 1))
            x = int(rng.binomial(n, min(max(b / c, 0.02), 0.98)))
            for m in (oracle.MATH_LIBM, oracle.MATH_DET):
                l = oracle.emission([n], [x], [c], P, m)[0]
                for s in range(c + 1):
                    p = s / c + (0.02 if 2 * s < c else -0.02 if 2 * s > c else 0.0)
                    t = onp.dbinom_log_mp(x, n, p) if lik == "binomial" else onp.dbetabinom_log_mp(x, n, p, rho)
                    rel = abs(l[s] - float(t)) / abs(float(t))
                    worst[m] = max(worst[m], rel)
    assert worst[0] < 1e-9 and worst[1] < 1e-9
    assert worst[1] < 5e-12      # keep an eye on the deterministic flavour's real accuracy


def test_lgamma_form_matches_r_style_saddle_point_dbinom(oracle):
    """stats::dbinom is Loader's saddle-point algorithm (bd0 / stirlerr), not an lgamma formula: the oracle's
    restatement of it (flavour RMATH) and the lgamma forms used everywhere else agree far inside 1e-9"""
    rng = np.random.default_rng(44)
    P = oracle.params(S=12, eps=0.0, loherror=0.02)
    worst_rel = 0.0
    for _ in range(400):
        c = int(rng.integers(1, 12))
        n = int(max(1, round(c * rng.gamma(1.76, 1 / 0.09))))
        x = int(rng.binomial(n, rng.uniform(0.02, 0.98)))
        lr = oracle.emission([n], [x], [c], P, oracle.MATH_RMATH)[0]
        for m in (oracle.MATH_LIBM, oracle.MATH_DET):
            l = oracle.emission([n], [x], [c], P, m)[0]
            fin = np.isfinite(lr)
            assert np.array_equal(np.isfinite(l), fin)
            worst_rel = max(worst_rel, float(np.max(np.abs(l[fin] - lr[fin]) / np.maximum(np.abs(lr[fin]), 1e-3))))
        # and the saddle-point restatement itself against 50-digit arithmetic
        for s in range(min(c, 11) + 1):
            p = s / c + (0.02 if 2 * s < c else -0.02 if 2 * s > c else 0.0)
            t = float(onp.dbinom_log_mp(x, n, p))
            assert abs(lr[s] - t) <= 4e-14 * max(1.0, abs(t)), (n, x, c, s)
    assert worst_rel < 1e-11
    # paths decoded from the R-style densities differ from the lgamma ones only at structural ties (CN 0)
    d = cases.small_cohort("binomial", n_cells=20, seed=6)
    cn = d["cn"].astype(np.int32)
    Pd = oracle.params(S=12, selftransitionprob=0.95)
    a = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], Pd, oracle.MATH_RMATH, nthreads=4)
    b = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], Pd, oracle.MATH_DET, nthreads=4)
    diff = np.nonzero(a != b)[0]
    assert diff.size < 1e-3 * a.size and (cn[diff] == 0).all()


def test_libm_and_det_flavours_agree_except_structural_ties(oracle):
    d = cases.small_cohort("binomial", n_cells=30, seed=5)
    P = oracle.params(S=12, selftransitionprob=0.95)
    cn = d["cn"].astype(np.int32)
    a = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_LIBM)
    b = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_DET)
    diff = np.nonzero(a != b)[0]
    # differences only in homozygous-deletion bins (CN 0 -> states 0/1 have mirror-image likelihoods,
    # an exact tie in real arithmetic that rounding decides; documented in DESIGN.md "tie-breaks")
    assert diff.size < 1e-3 * a.size
    assert (cn[diff] == 0).all()


def test_threads_do_not_change_results(oracle):
    d = cases.small_cohort("binomial", n_cells=12, seed=9)
    P = oracle.params(S=12, selftransitionprob=0.95)
    cn = d["cn"].astype(np.int32)
    a = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_LIBM, nthreads=1)
    b = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_LIBM, nthreads=4)
    assert np.array_equal(a, b)


def test_golden_fixtures(oracle):
    g = np.load(GOLD)
    n, x, cn, seg = g["edge_n"], g["edge_x"], g["edge_cn"], g["edge_seg"]
    for ps in cases.PARAM_SETS:
        P = oracle.params(**cases.param_dict(ps))
        for m, tag in ((oracle.MATH_DET, "det"), (oracle.MATH_LIBM, "libm")):
            assert np.array_equal(oracle.decode_batch(n, x, cn, seg, P, m), g["edge_path_%s_%s" % (ps[0], tag)]), ps[0]
    for lik in ("binomial", "betabinomial"):
        d = cases.small_cohort(lik)
        for ps in cases.PARAM_SETS[:5]:
            if (ps[2] == "betabinomial") != (lik == "betabinomial"):
                continue
            P = oracle.params(**cases.param_dict(ps))
            got = oracle.decode_batch(d["n"], d["x"], d["cn"].astype(np.int32), d["seg_off"], P, oracle.MATH_DET)
            assert np.array_equal(got, g["cohort_path_%s_det" % ps[0]]), ps[0]


def test_acceptance_accuracy_like_reference_tests(oracle):
    """tests/testthat/test-hscn.R:33 and test-ascn.R:23,50 ask for > 99 % state recovery on a clonal cohort"""
    d = cases.small_cohort("binomial", n_cells=25, seed=123, ragged=0.0)
    cn = d["cn"].astype(np.int32)
    P = oracle.params(model="hscn", S=12, selftransitionprob=0.95)
    got = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_LIBM, nthreads=2)
    B = np.minimum(got, d["cn"])          # the table-level clamp B <= state (R/callHSCN.R:289) handles CN 0
    assert (B == d["truth_b"]).mean() > 0.99
    db = cases.small_cohort("betabinomial", n_cells=25, seed=124, ragged=0.0)
    P = oracle.params(model="ascn", likelihood="betabinomial", rho=0.02, S=12, selftransitionprob=0.95)
    got = oracle.decode_batch(db["n"], db["x"], db["cn"].astype(np.int32), db["seg_off"], P, oracle.MATH_LIBM, nthreads=2)
    minor = np.minimum(db["truth_b"], db["cn"] - db["truth_b"])
    assert (got == minor).mean() > 0.99


def test_acceptance_loherror_like_reference_tests(oracle):
    """tests/testthat/test-hscn.R:34 and test-ascn.R:24,51: expect_equal(round(results$loherror, 2), loherror) --
    the between-pass estimate weighted.mean(BAF, totalcounts) over the A-Hom bins (R/callHSCN.R:898-902) recovers the
    simulated loherror to two decimals"""
    for seed in (123, 321):
        d = cases.small_cohort("binomial", n_cells=40, seed=seed, ragged=0.0)
        cn = d["cn"].astype(np.int32)
        P = oracle.params(model="hscn", S=12, selftransitionprob=0.95, loherror=0.02)
        path = oracle.decode_batch(d["n"], d["x"], cn, d["seg_off"], P, oracle.MATH_LIBM, nthreads=2)
        tab = oracle.postprocess(cn, path)
        ahom = (tab["B"] == 0) & (tab["A"] > 0)
        assert ahom.sum() > 1000
        infloherror = float(d["x"][ahom].sum() / d["n"][ahom].sum())      # = weighted.mean(alleleB/totalcounts, totalcounts)
        assert round(infloherror, 2) == 0.02, infloherror


def test_postprocess_matches_r_semantics(oracle):
    cn = np.array([2, 2, 3, 0, 1, 5, 4, 4], dtype=np.int32)
    path = np.array([1, 0, 3, 1, 1, 2, 2, 0], dtype=np.uint8)
    r = oracle.postprocess(cn, path)
    assert r["A"].tolist() == [1, 2, 0, 0, 0, 3, 2, 4]
    assert r["B"].tolist() == [1, 0, 3, 0, 1, 2, 2, 0]      # CN 0: B clamped to state (R/callHSCN.R:289)
    assert r["state_min"].tolist() == [1, 0, 0, 0, 0, 2, 2, 0]
    assert r["phase"].tolist() == [0, 1, 2, 0, 2, 1, 0, 1]
    assert r["state_phase"].tolist() == [0, 3, 4, 0, 4, 1, 0, 3]
    assert np.allclose(r["state_BAF"], [0.5, 0.0, 1.0, 0.5, 1.0, 0.4, 0.5, 0.0])


# ---------------------------------------------------------------------------------------------------------------
# oracle/_ref: the reference's own src/code.cpp compiled verbatim (against oracle/rcpp_mock).  These tests pin the
# restated viterbi / logspace_add of oracle.c -- and through it every GPU parity test -- to the reference's text.
@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref is not built and /root/reference is absent")
    return oracle


def test_ref_logspace_addcpp_equals_oracle_bit_for_bit(ref):
    rng = np.random.default_rng(3)
    a = np.concatenate((rng.uniform(-800, 5, 4000), [-np.inf, 0.0, -27.631021115928547, -745.2, -1e-300]))
    b = np.concatenate((rng.uniform(-800, 5, 4000), [-3.0, 0.0, -27.631021115928547, -np.inf, -1e-300]))
    L = ref.ref_lib()
    for u, v in zip(a, b):
        r, o = L.ref_logspace_addcpp(float(u), float(v)), ref.logspace_add(float(u), float(v), ref.MATH_LIBM)
        assert (math.isnan(r) and math.isnan(o)) or r == o, (u, v, r, o)
    assert math.isnan(L.ref_logspace_addcpp(-math.inf, -math.inf))          # the Inf - Inf case of src/code.cpp:7


@pytest.mark.parametrize("S,T", [(2, 1), (3, 2), (12, 5), (12, 300), (7, 64), (32, 40)])
def test_ref_viterbi_equals_oracle(ref, S, T):
    """random emissions with -Inf entries and planted exact ties, structured and dense transitions, shuffled
    observation indices: the compiled reference routine and oracle.c's restatement return the same path"""
    rng = np.random.default_rng(100 * S + T)
    for trial in range(6):
        nrow = T + int(rng.integers(0, 4))
        em = np.round(np.log(rng.random((nrow, S))), int(rng.choice([1, 3, 15])))     # coarse rounding plants ties
        em[rng.random((nrow, S)) < 0.15] = -np.inf
        if trial % 2 == 0:
            stp = float(rng.choice([0.5, 0.95, 0.999]))
            tp = np.full((S, S), (1 - stp) / max(S - 1, 1)); np.fill_diagonal(tp, stp)
        else:
            tp = rng.dirichlet(np.ones(S), size=S)
        tr = np.log(tp) if S > 1 else np.zeros((1, 1))
        obs = rng.permutation(nrow)[:T] + 1 if trial == 5 else np.arange(1, T + 1)
        want = ref.ref_viterbi(em, tr, obs)
        got = ref.viterbi(em, tr, obs, init_mode=0, decode=ref.DECODE_REFERENCE, math=ref.MATH_LIBM)
        assert np.array_equal(want, got), (S, T, trial, np.nonzero(want != got)[0][:5])


def test_ref_viterbi_decode_rule_is_not_a_backtrace(ref):
    """the quirk DESIGN.md section 2 describes, shown on the reference's own code: path[t] is read from the column
    winner of t+1 (no chaining), the last entry repeats, and one observation yields state 0"""
    em = np.log(np.array([[0.9, 0.1], [0.1, 0.9], [0.9, 0.1], [0.1, 0.9]]))
    tr = np.log(np.array([[0.5, 0.5], [0.5, 0.5]]))
    p = ref.ref_viterbi(em, tr, [1, 2, 3, 4])
    assert p[-1] == p[-2]
    assert ref.ref_viterbi(em[:1], tr, [1]).tolist() == [0.0]
    assert np.array_equal(p, ref.viterbi(em, tr, [1, 2, 3, 4]))


@pytest.mark.parametrize("ps", cases.PARAM_SETS[:8], ids=[p[0] for p in cases.PARAM_SETS[:8]])
def test_ref_viterbi_on_restated_emissions_reproduces_the_sequence_decode(ref, ps):
    """HaplotypeHMM / alleleHMM end in `viterbi(l, log(tProbs), seq_len(T))` (R/callHSCN.R:99-104, R/callASCN.R:98-101):
    feeding the compiled reference routine the restated emission and transition matrices gives the oracle's paths"""
    P = ref.params(**cases.param_dict(ps))
    tr = ref.transition(P, ref.MATH_LIBM)
    for name, n, x, cn in cases.edge_sequences():
        l = ref.emission(n, x, cn, P, ref.MATH_LIBM)                    # bins x states, as R holds it
        want = ref.ref_viterbi(l, tr, np.arange(1, n.size + 1))
        got = ref.hmm_sequence(n, x, cn, P, ref.MATH_LIBM)
        assert np.array_equal(want, got), (ps[0], name)


def test_thin_decision_margins_occur_only_at_cn0_bins(oracle):
    """bins decoded on a margin below 1e-9 are where a different but equally accurate dbinom could flip
    the call.  tools/margin_report.py counts them; on the synthetic cohort they are rare and all sit at
    homozygous-deletion (CN 0) bins, whose two live states have mirror-image likelihoods."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("margin_report", os.path.join(os.path.dirname(__file__), "..", "tools", "margin_report.py"))
    mr = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mr)
    r = mr.report(n_cells=12, model="hscn", likelihood="binomial", seed=11)
    assert r["bins"] > 40000
    assert r["lt_1e-9"] <= 1e-3 * r["bins"]
    assert r["lt_1e-9"] == r["lt_1e-9_at_cn0"]
    r2 = mr.report(n_cells=6, model="ascn", likelihood="betabinomial", seed=11)
    assert r2["lt_1e-9"] == r2["lt_1e-9_at_cn0"] <= 1e-3 * r2["bins"]


@pytest.mark.parametrize("model,lik,rho", [("hscn", "binomial", 0.0), ("hscn", "betabinomial", 0.02),
                                           ("ascn", "betabinomial", 0.02)])
def test_tables_identical_across_math_flavours(oracle, model, lik, rho):
    """CPU half of the table-identity argument (the GPU half is tests/test_gpu_parity.py): the libm, deterministic and
    R-style (saddle-point dbinom) flavours may decode different raw paths at homozygous-deletion bins only, and the
    clamp block R/callHSCN.R:282-290 removes exactly those differences -- every post-clamp column is identical."""
    from signals_b200 import synth
    d = synth.make_cohort(120, seed=77, likelihood=lik, ragged_drop=0.03)
    n, x, cn, so = d["n"], d["x"], d["cn"].astype(np.int32), d["seg_off"]
    P = oracle.params(model=model, likelihood=lik, S=12, rho=rho, selftransitionprob=0.95)
    base_path = oracle.decode_batch(n, x, cn, so, P, oracle.MATH_DET, nthreads=4)
    base = oracle.postprocess(cn, base_path)
    for flavour in (oracle.MATH_LIBM, oracle.MATH_RMATH):
        path = oracle.decode_batch(n, x, cn, so, P, flavour, nthreads=4)
        assert (cn[path != base_path] == 0).all()
        tab = oracle.postprocess(cn, path)
        for k in tab:
            assert np.array_equal(tab[k], base[k], equal_nan=True), (flavour, k)
