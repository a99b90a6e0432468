"""GPU parity tests: every call goes through the C ABI (libsignals_b200.so) and is compared with the CPU
oracle on the same seeded inputs, with the committed golden fixtures, and -- at BASELINE.json's full
sizes -- through size-independent properties.  Integer outputs (state paths, table codes) must be
bit-exact; emission log-likelihoods are bit-exact against the oracle's deterministic flavour and within
1e-9 relative (1e-12 absolute floor near zero) of its libm flavour."""
import math
import os

import numpy as np
import pytest

import cases
from signals_b200 import _lib, hmm, synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "hmm_golden.npz")
IDS = [p[0] for p in cases.PARAM_SETS]


def _kw(ps, **extra):
    d = cases.param_dict(ps)
    out = dict(model=d["model"], likelihood=d["likelihood"], n_states=d["S"], loherror=d["loherror"], eps=d["eps"],
               rho=d["rho"], Abias=d["Abias"], selftransitionprob=d["selftransitionprob"])
    out.update(extra)
    return out


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


@pytest.fixture(params=["split", "split_staged", "fused"])
def sched(request, monkeypatch):
    """the schedules of the reference decode: "split" (k_emit + k_recur, the default; small batches store the decoded states
    directly), "split_staged" (k_recur's variant for large batches: four-bin blocks, states staged in shared memory) and
    "fused" (k_decode)"""
    monkeypatch.setenv("SGB_SCHED", request.param.split("_")[0])
    monkeypatch.setenv("SGB_RECUR_STAGED", "1" if request.param == "split_staged" else "0")
    return request.param


def test_device_math_bits_equal_host_bits(oracle):
    """the deterministic math gives the same bits on the device as on the host, function by function"""
    rng = np.random.default_rng(1)
    lib = _lib.load()
    sets = {
        0: np.concatenate((-rng.uniform(0, 750, 200000), [0.0, -0.0, -np.inf, np.nan, -708.0, -708.5, 1.5, 709.5])),
        1: np.concatenate((np.exp(rng.uniform(-740, 700, 200000)), rng.uniform(0.5, 2, 100000),
                           [0.0, 1.0, np.inf, -1.0, np.nan, 5e-324, 2.2e-308])),
        2: np.concatenate((rng.uniform(0, 1, 200000), np.exp(-rng.uniform(0, 740, 100000)), [0.0, 1.0, np.nan])),
        3: np.concatenate((rng.uniform(-1, 1, 200000), -rng.uniform(0, 1, 100000), [-1.0, 0.0, np.nan, 3.0, 1e10])),
        4: np.concatenate((rng.uniform(1e-9, 12, 200000), np.exp(rng.uniform(0, 14, 100000)),
                           [0.0, 1.0, 2.0, 10.0, np.inf, np.nan, -1.0, 1e-300])),
        6: np.concatenate((np.arange(0, 4100, dtype=np.float64), [1e5, 1e6, 2.0 ** 28])),
    }
    for fn, xs in sets.items():
        xs = np.ascontiguousarray(xs, dtype=np.float64)
        out = np.empty_like(xs)
        hmm.check(lib.sgb_math_eval(fn, xs.ctypes.data, None, xs.size, out.ctypes.data))
        ref = oracle.math_eval(fn, xs, math=oracle.MATH_DET)
        nan = np.isnan(ref)
        assert np.array_equal(np.isnan(out), nan), fn
        assert np.array_equal(_bits(out)[~nan], _bits(ref)[~nan]), (fn, int((_bits(out)[~nan] != _bits(ref)[~nan]).sum()))
    a = -rng.uniform(0, 700, 300000); b = -rng.uniform(0, 80, 300000)
    a[:5] = [-np.inf, -np.inf, np.nan, -3.0, 0.0]; b[:5] = [-27.6, -np.inf, -1.0, -3.0, 0.0]
    out = hmm.logspace_add(a, b)
    ref = oracle.math_eval(5, a, b, oracle.MATH_DET)
    nan = np.isnan(ref)
    assert np.array_equal(np.isnan(out), nan) and np.array_equal(_bits(out)[~nan], _bits(ref)[~nan])


def test_logspace_addcpp_scalar(oracle):
    assert abs(hmm.logspace_addcpp(0.0, 0.0) - math.log(2.0)) < 1e-15
    assert hmm.logspace_addcpp(-math.inf, -27.6) == -27.6
    assert math.isnan(hmm.logspace_addcpp(-math.inf, -math.inf))
    assert hmm.logspace_addcpp(-3.0, math.log(1e-12)) == oracle.logspace_add(-3.0, math.log(1e-12), oracle.MATH_DET)


@pytest.mark.parametrize("ps", cases.PARAM_SETS, ids=IDS)
def test_edge_sequences_bit_exact(oracle, ps, sched):
    g = np.load(GOLD)
    n, x, cn, seg = g["edge_n"], g["edge_x"], g["edge_cn"], g["edge_seg"]
    P = oracle.params(**cases.param_dict(ps))
    ref = oracle.decode_batch(n, x, cn, seg, P, oracle.MATH_DET)
    got, ll = hmm.decode_arrays(n, x, cn, seg, want_loglik=True, **_kw(ps))
    assert np.array_equal(got, ref), (ps[0], np.nonzero(got != ref)[0][:10])
    assert np.array_equal(got, g["edge_path_%s_det" % ps[0]])
    # emission matrix: bit-exact vs the deterministic oracle / golden, 1e-9 relative vs libm
    gold_l = g["edge_l_%s_det" % ps[0]]
    assert np.array_equal(np.isinf(ll), np.isinf(gold_l))
    fin = np.isfinite(gold_l)
    assert np.array_equal(_bits(ll)[fin], _bits(gold_l)[fin])
    S = cases.param_dict(ps)["S"]
    for s in range(seg.size - 1):
        a, b = int(seg[s]), int(seg[s + 1])
        lm = oracle.emission(n[a:b], x[a:b], cn[a:b], P, oracle.MATH_LIBM)
        blk = ll[a * S: b * S].reshape(S, b - a).T
        f = np.isfinite(lm)
        assert np.array_equal(np.isfinite(blk), f)
        assert np.allclose(blk[f], lm[f], rtol=1e-9, atol=1e-12)
    # chained (textbook) decode
    gotc = hmm.decode_arrays(n, x, cn, seg, **_kw(ps, decode="chained"))
    assert np.array_equal(gotc, g["edge_chained_%s_det" % ps[0]])


@pytest.mark.parametrize("ps", cases.PARAM_SETS[:5], ids=IDS[:5])
def test_small_cohort_bit_exact_vs_golden(oracle, ps, sched):
    g = np.load(GOLD)
    d = cases.small_cohort("betabinomial" if ps[2] == "betabinomial" else "binomial")
    got = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    ref = g["cohort_path_%s_det" % ps[0]]
    assert np.array_equal(got, ref), int((got != ref).sum())
    # uint8 and int32 copy-number columns give the same answer
    got32 = hmm.decode_arrays(d["n"], d["x"], d["cn"].astype(np.int32), d["seg_off"], **_kw(ps))
    assert np.array_equal(got32, ref)


def test_libm_oracle_differs_only_at_structural_ties(oracle):
    d = cases.small_cohort("binomial", n_cells=60, seed=5)
    ps = cases.PARAM_SETS[0]
    got = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    lm = oracle.decode_batch(d["n"], d["x"], d["cn"].astype(np.int32), d["seg_off"],
                             oracle.params(**cases.param_dict(ps)), oracle.MATH_LIBM, nthreads=4)
    diff = np.nonzero(got != lm)[0]
    assert diff.size < 1e-3 * got.size
    assert (d["cn"][diff] == 0).all()       # mirror-image likelihoods at CN 0: exact ties decided by rounding


def test_chunked_pipeline_is_invariant(oracle, monkeypatch, sched):
    d = cases.small_cohort("binomial", n_cells=30, seed=8)
    ps = cases.PARAM_SETS[0]
    ref = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    for chunk in ("1", "5000", "40000"):
        monkeypatch.setenv("SGB_CHUNK_BINS", chunk)
        got = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
        assert np.array_equal(got, ref), chunk
    monkeypatch.delenv("SGB_CHUNK_BINS")


def test_empty_and_degenerate_inputs(sched):
    ps = cases.PARAM_SETS[0]
    e = np.zeros(0, np.int32)
    out = hmm.decode_arrays(e, e, e, np.zeros(1, np.int64), **_kw(ps))
    assert out.size == 0
    # zero-length sequences in between are skipped
    n = np.array([10, 12, 9], np.int32); x = np.array([5, 6, 0], np.int32); cn = np.array([2, 2, 1], np.int32)
    seg = np.array([0, 0, 2, 2, 3, 3], np.int64)
    out = hmm.decode_arrays(n, x, cn, seg, **_kw(ps))
    assert out.tolist() == [1, 1, 0]        # T=2 sequence repeats its first call; T=1 sequence yields 0


def test_input_validation_errors(sched):
    ps = cases.PARAM_SETS[0]
    n = np.array([10, 12], np.int32); x = np.array([5, 6], np.int32); cn = np.array([2, 2], np.int32)
    seg = np.array([0, 2], np.int64)
    with pytest.raises(hmm.SignalsB200Error, match="negative"):
        hmm.decode_arrays(np.array([10, -1], np.int32), x, cn, seg, **_kw(ps))
    with pytest.raises(hmm.SignalsB200Error, match="2\\^28"):
        hmm.decode_arrays(np.array([10, 1 << 28], np.int32), x, cn, seg, **_kw(ps))
    with pytest.raises(hmm.SignalsB200Error, match="copy-number"):
        hmm.decode_arrays(n, x, np.array([2, 300], np.int32), seg, **_kw(ps))
    with pytest.raises(hmm.SignalsB200Error, match="n_states"):
        hmm.decode_arrays(n, x, cn, seg, **_kw(ps, n_states=33))
    with pytest.raises(hmm.SignalsB200Error, match="non-decreasing"):
        hmm.decode_arrays(n, x, cn, np.array([0, 2, 1], np.int64), **_kw(ps))
    with pytest.raises(ValueError, match="binomial or betabinomial"):
        hmm.decode_arrays(n, x, cn, seg, **_kw(ps, likelihood="poisson"))
    # the library still works after an error
    assert hmm.decode_arrays(n, x, cn, seg, **_kw(ps)).tolist() == [1, 1]


def test_standalone_viterbi(oracle):
    rng = np.random.default_rng(4)
    from oracle import oracle_np as onp
    for S, T, stp in ((2, 3, 0.7), (12, 200, 0.95), (12, 90, 0.0), (12, 60, 1.0), (5, 1, 0.9), (1, 10, 0.9),
                      (33, 70, 0.9), (64, 40, 0.99), (128, 25, 0.9)):
        e = np.log(rng.dirichlet(np.ones(S) * 0.5, size=T + 3))
        e[rng.random(e.shape) < 0.1] = -np.inf
        if S > 1:
            tr = np.log(rng.dirichlet(np.ones(S), size=S)) if stp in (0.7, 0.95, 0.99) else onp._transition(np.arange(S), stp)
        else:
            tr = np.zeros((1, 1))
        obs = rng.integers(1, T + 4, size=T)
        for init_mode, fn in ((0, hmm.viterbi), (1, hmm.viterbiR)):
            ref = oracle.viterbi(e, tr, obs, init_mode=init_mode, math=oracle.MATH_DET)
            assert np.array_equal(fn(e, tr, obs), ref), (S, T, stp, init_mode)
        refc = oracle.viterbi(e, tr, obs, decode=oracle.DECODE_CHAINED, math=oracle.MATH_DET)
        assert np.array_equal(hmm.viterbi(e, tr, obs, decode="chained"), refc), (S, T, stp)
    # ties -> first index; NaN treated as -Inf
    e = np.log(np.full((6, 4), 0.25)); tr = np.log(np.full((4, 4), 0.25))
    assert hmm.viterbi(e, tr, np.arange(1, 7)).tolist() == [0.0] * 6
    e2 = e.copy(); e2[2, 0] = np.nan
    assert np.array_equal(hmm.viterbi(e2, tr, np.arange(1, 7)), oracle.viterbi(e2, tr, np.arange(1, 7), math=oracle.MATH_DET))
    with pytest.raises(hmm.SignalsB200Error, match="observations"):
        hmm.viterbi(e, tr, [1, 2, 9])


def test_standalone_viterbi_against_compiled_reference(oracle):
    """sgb_viterbi vs oracle/_ref (the reference's src/code.cpp compiled verbatim against oracle/rcpp_mock): the
    routine is additions and comparisons only, so the paths must agree exactly wherever the device's log(S) for the
    initial column has the same bits as libm's"""
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref is not built")
    rng = np.random.default_rng(12)
    checked = 0
    for S in (2, 3, 4, 7, 8, 12, 16, 25):
        if float(oracle.math_eval(1, np.array([float(S)]), math=oracle.MATH_DET)[0]) != math.log(S):
            continue
        for T in (1, 2, 50, 400):
            em = np.round(np.log(rng.random((T + 2, S))), int(rng.choice([1, 2, 15])))
            em[rng.random(em.shape) < 0.1] = -np.inf
            stp = 0.95
            tp = np.full((S, S), (1 - stp) / (S - 1)); np.fill_diagonal(tp, stp)
            for tr in (np.log(tp), np.log(rng.dirichlet(np.ones(S), size=S))):
                obs = np.arange(1, T + 1)
                assert np.array_equal(hmm.viterbi(em, tr, obs), oracle.ref_viterbi(em, tr, obs)), (S, T)
                checked += 1
    assert checked >= 16


def test_postprocess_codes(oracle):
    rng = np.random.default_rng(9)
    cn = rng.integers(0, 12, 5000).astype(np.int32)
    path = rng.integers(0, 12, 5000).astype(np.uint8)
    got = hmm.postprocess_arrays(cn, path)
    ref = oracle.postprocess(cn, path)
    for k in ref:
        if k == "state_BAF":
            assert np.array_equal(_bits(got[k]), _bits(ref[k])), k
        else:
            assert np.array_equal(got[k], ref[k]), k


def test_between_pass_statistics(oracle):
    """infloherror = weighted.mean(BAF, totalcounts) over A-Hom bins (R/callHSCN.R:898-902) and the A|B histogram"""
    d = cases.small_cohort("binomial", n_cells=40, seed=21)
    ps = cases.PARAM_SETS[0]
    path = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    sums, hist = hmm.pass_stats_arrays(d["n"], d["x"], d["cn"], path, n_states=12)
    ref = oracle.postprocess(d["cn"].astype(np.int32), path)
    ahom = (ref["B"] == 0) & (ref["A"] > 0)
    assert sums.tolist() == [int(d["x"][ahom].sum()), int(d["n"][ahom].sum()), int(ahom.sum())]
    h = np.zeros((12, 12), np.int64)
    sel = ref["B"] > 0
    np.add.at(h, (ref["A"][sel], ref["B"][sel]), 1)
    assert np.array_equal(hist, h)
    # the R expression: weighted.mean(x = BAF, w = totalcounts)
    baf = d["x"][ahom] / d["n"][ahom]
    wm = float((baf * d["n"][ahom]).sum() / d["n"][ahom].sum())
    got = hmm.infer_loherror(d["n"], d["x"], d["cn"], path, maxloherror=1.0)
    assert abs(got - wm) < 1e-12 and 0.01 < got < 0.04          # the cohort was simulated with loherror = 0.02
    assert hmm.infer_loherror(d["n"], d["x"], d["cn"], path, maxloherror=0.015) == 0.015


def test_r_level_mirror_end_to_end(oracle):
    """.callHaplotypeSpecificCN_ / ASCN twin on a long table, against the oracle's table-level restatement"""
    d = synth.make_cohort(6, lens=synth.LENS_DLPBINS, seed=31, ragged_drop=0.05)
    df = synth.to_frame(d).sample(frac=1.0, random_state=1)        # the batch function sorts (R/callHSCN.R:241)
    for model, fn in (("hscn", hmm.callHaplotypeSpecificCN_), ("ascn", hmm.callAlleleSpecificCN_)):
        res = fn(df, loherror=0.02, maxCN=11, selftransitionprob=0.95)
        s = res.sort_values(["cell_id", "chr", "start"], kind="stable")
        assert len(res) == len(df) and list(res["cell_id"]) == list(s["cell_id"])
        # rebuild the oracle's view in the reference's row order
        seg = hmm._segments(res, by_cell=True)
        P = oracle.params(model=model, S=12, selftransitionprob=0.95)
        path = oracle.decode_batch(res["totalcounts"].to_numpy(), res["alleleB"].to_numpy(), res["state"].to_numpy(),
                                   seg, P, oracle.MATH_DET)
        ref = oracle.postprocess(res["state"].to_numpy(), path)
        assert np.array_equal(res["A"].to_numpy(), ref["A"]) and np.array_equal(res["B"].to_numpy(), ref["B"])
        assert np.array_equal(res["state_min"].to_numpy(), ref["state_min"])
        assert (res["state_AS_phased"] == res["A"].astype(str) + "|" + res["B"].astype(str)).all()
        assert set(res["LOH"]) <= {"LOH", "NO"} and set(res["phase"]) <= {"Balanced", "A", "B"}
        assert ((res["LOH"] == "LOH") == (res["state_min"] == 0)).all()
        hi = res[res["state"] > 4]
        assert (hi["state_AS"] == hi["state"].astype(str)).all()
    # single-sequence functions return the reference's list(minorcn, l)
    one = res[(res["cell_id"] == res["cell_id"].iloc[0]) & (res["chr"] == "1")]
    r = hmm.HaplotypeHMM(one["totalcounts"], one["alleleB"], one["state"], np.arange(12), selftransitionprob=0.95)
    P = oracle.params(S=12, selftransitionprob=0.95)
    pth, l = oracle.hmm_sequence(one["totalcounts"].to_numpy(), one["alleleB"].to_numpy(), one["state"].to_numpy(), P,
                                 oracle.MATH_DET, want_l=True)
    assert np.array_equal(r["minorcn"], pth) and r["l"].shape == l.shape
    assert np.array_equal(np.isinf(r["l"]), np.isinf(l)) and np.array_equal(_bits(r["l"])[np.isfinite(l)], _bits(l)[np.isfinite(l)])
    rr = hmm.HaplotypeHMM(one["totalcounts"], one["alleleB"], one["state"], np.arange(12), selftransitionprob=0.95,
                          viterbiver="R")
    tp = np.full((12, 12), 0.05 / 11); np.fill_diagonal(tp, 0.95)
    assert np.array_equal(rr["minorcn"], oracle.viterbi(l, np.log(tp), np.arange(1, len(pth) + 1), init_mode=1,
                                                        math=oracle.MATH_DET))
    ra = hmm.alleleHMM(one["totalcounts"], one["alleleB"], one["state"], np.arange(12), selftransitionprob=0.95)
    assert ra["l"].shape == l.shape
    cell = res[res["cell_id"] == res["cell_id"].iloc[0]].drop(columns=["state_min"])
    a1 = hmm.assignHaplotypeHMM(cell, np.arange(12), selftransitionprob=0.95)
    assert "state_min" in a1.columns and len(a1) == len(cell)


def test_device_resident_api_matches_host_api(oracle, sched):
    import torch
    d = cases.small_cohort("binomial", n_cells=50, seed=12)
    ps = cases.PARAM_SETS[0]
    ref = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    dev = torch.device("cuda:0")
    tn = torch.from_numpy(d["n"]).to(dev); tx = torch.from_numpy(d["x"]).to(dev); tc = torch.from_numpy(d["cn"]).to(dev)
    out = torch.zeros(d["n"].size, dtype=torch.uint8, device=dev)
    plan = hmm.DevicePlan(d["n"].size, d["seg_off"], n_states=12)
    P = hmm.make_params(**_kw(ps))
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        out.zero_()
        plan.decode(P, tn, tx, tc, out, stream=stream)
        plan.check(stream)
        assert np.array_equal(out.cpu().numpy(), ref)
    assert plan.scratch_bytes() > 0
    # the `l` matrix written on the device
    S = 12
    ll = torch.empty(d["n"].size * S, dtype=torch.float64, device=dev)
    plan.decode(P, tn, tx, tc, out, d_loglik=ll, stream=stream)
    plan.check(stream)
    _, ll_host = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], want_loglik=True, **_kw(ps))
    a, b = ll.cpu().numpy(), ll_host
    assert np.array_equal(np.isinf(a), np.isinf(b)) and np.array_equal(_bits(a)[np.isfinite(b)], _bits(b)[np.isfinite(b)])
    # validation errors surface through plan.check
    bad = tn.clone(); bad[3] = -5
    plan.decode(P, bad, tx, tc, out, stream=stream)
    with pytest.raises(hmm.SignalsB200Error, match="negative"):
        plan.check(stream)
    plan.close()


def test_in_place_and_packed_staging_agree(oracle, monkeypatch):
    """(fused schedule) k_decode reads 16-byte aligned uint8-cn columns in place; anything else goes through k_pack's packed
    records.  Same rows at aligned and misaligned addresses, ragged lengths, column end inside a 4-bin block."""
    import torch
    monkeypatch.setenv("SGB_SCHED", "fused")
    d = cases.small_cohort("binomial", n_cells=40, seed=21)
    keep = d["seg_off"][-2] + 3                      # cut the last sequence to 3 bins: N is not a multiple of 4
    seg = d["seg_off"].copy(); seg[-1] = keep
    n, x, cn = d["n"][:keep], d["x"][:keep], d["cn"][:keep]
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for ps in (cases.PARAM_SETS[0], [q for q in cases.PARAM_SETS if q[1] == "ascn" and q[2] == "binomial"][0]):
        P = hmm.make_params(**_kw(ps))
        ref = oracle.decode_batch(n, x, cn.astype(np.int32), seg, oracle.params(**cases.param_dict(ps)), oracle.MATH_DET)
        plan = hmm.DevicePlan(int(keep), seg, n_states=cases.param_dict(ps)["S"])
        for off in (0, 1, 4):                        # element offset of the views: 0 and 4 keep n/x 16-byte aligned
            bn = torch.zeros(keep + 8, dtype=torch.int32, device=dev); bx = torch.zeros_like(bn)
            bc = torch.full((keep + 8,), 7, dtype=torch.uint8, device=dev)
            tn, tx, tc = bn[off: off + keep], bx[off: off + keep], bc[off: off + keep]
            tn.copy_(torch.from_numpy(n)); tx.copy_(torch.from_numpy(x)); tc.copy_(torch.from_numpy(cn))
            out = torch.zeros(keep, dtype=torch.uint8, device=dev)
            plan.decode(P, tn, tx, tc, out, stream=stream)
            plan.check(stream)
            assert np.array_equal(out.cpu().numpy(), ref), (ps[0], off)
            assert plan.last_staging() == (1 if off == 1 else 2), (off, plan.last_staging())
        # an int32 copy-number column (what R passes) is narrowed on the device and then read in place
        out.zero_()
        plan.decode(P, torch.from_numpy(n).to(dev), torch.from_numpy(x).to(dev), torch.from_numpy(cn.astype(np.int32)).to(dev),
                    out, stream=stream)
        plan.check(stream)
        assert np.array_equal(out.cpu().numpy(), ref) and plan.last_staging() == 3
        badc = torch.from_numpy(cn.astype(np.int32)).to(dev); badc[keep - 1] = 300
        plan.decode(P, torch.from_numpy(n).to(dev), torch.from_numpy(x).to(dev), badc, out, stream=stream)
        with pytest.raises(hmm.SignalsB200Error, match="copy-number"):
            plan.check(stream)
        # validation on the in-place path
        bad = torch.from_numpy(n).to(dev); bad[5] = 1 << 28
        plan.decode(P, bad, torch.from_numpy(x).to(dev), torch.from_numpy(cn).to(dev), out, stream=stream)
        with pytest.raises(hmm.SignalsB200Error, match="2\\^28"):
            plan.check(stream)
        plan.close()


@pytest.mark.parametrize("seed", range(cases.N_RANDOM_CASES))
def test_randomised_models_and_inputs_bit_exact(oracle, seed, sched):
    """random state counts (all three kernel buckets), transition regimes, eps / Abias / loherror settings, total copy
    numbers beyond the tables and the state space, counts on both sides of every log-factorial branch, counts
    outside the support, 1-bin sequences; uint8 (in place) and int32 (narrowed) copy-number columns"""
    n, x, cn, seg, kw = cases.random_case(seed)
    P = oracle.params(model=kw["model"], likelihood=kw["likelihood"], S=kw["n_states"], loherror=kw["loherror"], eps=kw["eps"],
                      rho=kw["rho"], Abias=kw["Abias"], selftransitionprob=kw["selftransitionprob"])
    ref = oracle.decode_batch(n, x, cn.astype(np.int32), seg, P, oracle.MATH_DET)
    got = hmm.decode_arrays(n, x, cn, seg, **kw)
    assert np.array_equal(got, ref), (kw, np.nonzero(got != ref)[0][:8])
    if seed % 3 == 0:
        Pc = oracle.params(model=kw["model"], likelihood=kw["likelihood"], S=kw["n_states"], loherror=kw["loherror"],
                           eps=kw["eps"], rho=kw["rho"], Abias=kw["Abias"], selftransitionprob=kw["selftransitionprob"],
                           decode="chained")
        refc = oracle.decode_batch(n, x, cn.astype(np.int32), seg, Pc, oracle.MATH_DET)
        assert np.array_equal(hmm.decode_arrays(n, x, cn, seg, decode="chained", **kw), refc)


def test_shutdown_and_reinitialise():
    """sgb_shutdown releases every per-device resource; the next call rebuilds the context (sgb_init is optional)"""
    lib = _lib.load()
    d = cases.small_cohort("binomial", n_cells=6, seed=3)
    ps = cases.PARAM_SETS[0]
    ref = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    for _ in range(2):
        assert lib.sgb_shutdown() == 0
        assert np.array_equal(hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps)), ref)
    assert lib.sgb_shutdown() == 0 and lib.sgb_init(0) == 0


def test_multi_device_sharding_matches_single_device():
    lib = _lib.load()
    if lib.sgb_device_count() < 2:
        pytest.skip("needs >= 2 CUDA devices")
    d = cases.small_cohort("binomial", n_cells=60, seed=14)
    ps = cases.PARAM_SETS[0]
    ref = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **_kw(ps))
    got = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], n_devices=2, **_kw(ps))
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("model,lik,rho", [("hscn", "binomial", 0.0), ("ascn", "betabinomial", 0.02)])
def test_full_size_properties(oracle, model, lik, rho):
    """BASELINE.json configs 2/3 at full size (10,000 cells x 5,700 bins = 57 M cell-bins)"""
    n_cells = 10000
    d = synth.make_cohort(n_cells, seed=20260002 if model == "hscn" else 20260003, likelihood=lik)
    kw = dict(model=model, likelihood=lik, n_states=12, loherror=0.02, eps=1e-12, rho=rho, selftransitionprob=0.95)
    out = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **kw)
    so = d["seg_off"]
    # (1) deterministic / idempotent
    assert np.array_equal(out, hmm.decode_arrays(d["n"], d["x"], d["cn"], so, **kw))
    # (2) the reference's decode repeats the last call of every sequence (src/code.cpp:37-42)
    assert np.array_equal(out[so[1:] - 1], out[so[1:] - 2])
    # (3) masks: HSCN never calls a state above max(cn,1) (R/callHSCN.R:80), ASCN never above cn/2
    #     (R/callASCN.R:81) -- except in the LAST bin of a sequence, which repeats the previous call whatever
    #     its own copy number is (src/code.cpp:37); the table-level clamp B <= state handles that one
    cnv = d["cn"].astype(np.int64)
    inner = np.ones(out.size, dtype=bool)
    inner[so[1:] - 1] = False
    if model == "hscn":
        assert (out[inner] <= np.maximum(cnv, 1)[inner]).all()
    else:
        assert (out[inner] <= (cnv // 2)[inner]).all()
    # (4) cells are independent: decoding cells in another order permutes the output
    perm = np.random.default_rng(1).permutation(400)
    per_cell = 5700
    idx = (perm[:, None] * per_cell + np.arange(per_cell)[None, :]).ravel()
    seg_small = np.arange(0, 400 * 23 + 1)
    seg_small = np.concatenate(([0], np.cumsum(np.tile(np.array(synth.LENS_5700), 400))))
    sub = hmm.decode_arrays(d["n"][idx], d["x"][idx], d["cn"][idx], seg_small, **kw)
    assert np.array_equal(sub, out[idx])
    # (5) bit-exact against the oracle on ALL 57 M cell-bins (the deterministic flavour does ~1e7 cell-bins/s on 8 threads)
    ref = oracle.decode_batch(d["n"], d["x"], d["cn"].astype(np.int32), so,
                              oracle.params(model=model, likelihood=lik, S=12, rho=rho, selftransitionprob=0.95),
                              oracle.MATH_DET, nthreads=os.cpu_count() or 1)
    assert np.array_equal(out, ref), int((out != ref).sum())
    # (6) acceptance bar of the reference's tests (> 99 % recovery of the simulated states)
    truth = d["truth_b"] if model == "hscn" else np.minimum(d["truth_b"], d["cn"] - d["truth_b"])
    called = np.minimum(out, d["cn"]) if model == "hscn" else out
    assert (called == truth).mean() > 0.99
    # (7) a checksum of the whole result, stable across runs / chunk sizes
    assert int(out.astype(np.uint64).sum()) == int(hmm.decode_arrays(d["n"], d["x"], d["cn"], so, **kw).astype(np.uint64).sum())


@pytest.mark.parametrize("model,lik,rho", [("hscn", "binomial", 0.0), ("ascn", "betabinomial", 0.02),
                                           ("ascn", "binomial", 0.0), ("hscn", "betabinomial", 0.02)])
def test_long_sequences_30000_bin_grid(oracle, model, lik, rho, sched):
    """BASELINE.json configs[4] geometry (23 sequences of 476..2,463 bins per cell), 200 cells = 6 M cell-bins, against the
    oracle: reference and chained decode, uint8 and int32 copy-number columns through the host API, and 16-byte
    aligned / one-element-offset device columns through the device API (in-place ring vs packed records)."""
    import torch
    d = synth.make_cohort(200, lens=synth.LENS_30000, seed=20260005, likelihood=lik, ragged_drop=0.02)
    n, x, cn, so = d["n"], d["x"], d["cn"], d["seg_off"]
    N = n.size
    kw = dict(model=model, likelihood=lik, n_states=12, loherror=0.02, eps=1e-12, rho=rho, selftransitionprob=0.95)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for decode in ("reference", "chained"):
        P = oracle.params(model=model, likelihood=lik, S=12, rho=rho, selftransitionprob=0.95, decode=decode)
        ref = oracle.decode_batch(n, x, cn.astype(np.int32), so, P, oracle.MATH_DET, nthreads=os.cpu_count() or 1)
        for cnv in (cn, cn.astype(np.int32)):
            got = hmm.decode_arrays(n, x, cnv, so, decode=decode, **kw)
            assert np.array_equal(got, ref), (decode, cnv.dtype, int((got != ref).sum()))
        plan = hmm.DevicePlan(N, so, n_states=12, decode=decode)
        PP = hmm.make_params(decode=decode, **kw)
        for off in (0, 1):
            bn = torch.zeros(N + 8, dtype=torch.int32, device=dev); bx = torch.zeros_like(bn)
            bc = torch.full((N + 8,), 9, dtype=torch.uint8, device=dev)
            tn, tx, tc = bn[off: off + N], bx[off: off + N], bc[off: off + N]
            tn.copy_(torch.from_numpy(n)); tx.copy_(torch.from_numpy(x)); tc.copy_(torch.from_numpy(cn))
            out = torch.zeros(N, dtype=torch.uint8, device=dev)
            plan.decode(PP, tn, tx, tc, out, stream=stream)
            plan.check(stream)
            assert np.array_equal(out.cpu().numpy(), ref), (decode, off)
        plan.close()


@pytest.mark.parametrize("model,lik,rho", [("hscn", "binomial", 0.0), ("hscn", "betabinomial", 0.02),
                                           ("ascn", "binomial", 0.0), ("ascn", "betabinomial", 0.02)])
def test_tables_identical_across_math_flavours_and_gpu(oracle, model, lik, rho):
    """north_star asks for identical hscn / ascn TABLES.  State paths computed with different fp64 math (glibc libm, the
    deterministic functions the GPU uses, R's own saddle-point dbinom) differ only where the model has an exact tie --
    homozygous-deletion bins, whose two mirror-image states round differently -- and the reference's clamp block
    (R/callHSCN.R:282-290: B = min(state_min, state) with state = 0 there) erases exactly those.  So the post-clamp columns
    (A, B, state_min, LOH, phase, state_phase, state_BAF; R/util.R:1078-1097) must be IDENTICAL in all four."""
    d = synth.make_cohort(300, seed=77, likelihood=lik, ragged_drop=0.03)
    n, x, cn, so = d["n"], d["x"], d["cn"], d["seg_off"]
    assert (cn == 0).sum() > 1000                   # the cohort does contain hom-del bins
    kw = dict(model=model, likelihood=lik, n_states=12, loherror=0.02, eps=1e-12, rho=rho, selftransitionprob=0.95)
    P = oracle.params(model=model, likelihood=lik, S=12, rho=rho, selftransitionprob=0.95)
    gpu_path = hmm.decode_arrays(n, x, cn, so, **kw)
    gpu_tab = hmm.postprocess_arrays(cn, gpu_path)
    n_path_diff = 0
    for flavour in (oracle.MATH_DET, oracle.MATH_LIBM, oracle.MATH_RMATH):
        path = oracle.decode_batch(n, x, cn.astype(np.int32), so, P, flavour, nthreads=os.cpu_count() or 1)
        diff = path != gpu_path
        n_path_diff += int(diff.sum())
        assert (cn[diff] == 0).all()                # raw paths may differ at CN-0 bins only
        tab = oracle.postprocess(cn.astype(np.int32), path)
        for k in tab:
            if k == "state_BAF":
                assert np.array_equal(_bits(gpu_tab[k]), _bits(tab[k])), (flavour, k)
            else:
                assert np.array_equal(gpu_tab[k], tab[k]), (flavour, k, int((gpu_tab[k] != tab[k]).sum()))
    if model == "hscn":
        assert n_path_diff > 0                      # the test would be vacuous if rounding never decided a tie


def test_typed_columns_pageable_and_pinned(oracle, monkeypatch):
    """sgb_hmm_decode_typed: every column in its own element type (R hands over INTSXP = int32 or REALSXP = double
    vectors, pageable), states returned as uint8 or as doubles, staged through page-locked buffers by worker threads
    (narrowed to uint16 / uint8) or copied directly; all must give the oracle's paths."""
    import torch
    d = cases.small_cohort("binomial", n_cells=120, seed=33)
    n, x, cn, so = d["n"], d["x"], d["cn"], d["seg_off"]
    ps = cases.PARAM_SETS[0]
    ref = oracle.decode_batch(n, x, cn.astype(np.int32), so, oracle.params(**cases.param_dict(ps)), oracle.MATH_DET, nthreads=4)
    monkeypatch.setenv("SGB_CHUNK_BINS", "70000")          # several chunks, so the slot staging is reused
    combos = [(np.int32, np.int32, np.int32), (np.float64, np.float64, np.float64), (np.float64, np.int32, np.uint8),
              (np.uint16, np.uint16, np.uint8), (np.int32, np.float64, np.float64)]
    for tn, tx, tc in combos:
        for out_dtype in (np.uint8, np.float64, np.int32):
            got = hmm.decode_arrays(n.astype(tn), x.astype(tx), cn.astype(tc), so, out_dtype=out_dtype, **_kw(ps))
            assert got.dtype == np.dtype(out_dtype) and np.array_equal(got.astype(np.uint8), ref), (tn, tx, tc, out_dtype)
    # page-locked int32 columns: staged (default) and direct (SGB_HOST_STAGE=never) feeds agree
    hn = torch.from_numpy(n).pin_memory(); hx = torch.from_numpy(x).pin_memory()
    hc = torch.from_numpy(cn.astype(np.int32)).pin_memory(); ho = torch.zeros(n.size, dtype=torch.uint8).pin_memory()
    lib = _lib.load()
    P = hmm.make_params(**_kw(ps))
    import ctypes
    for mode in ("auto", "never", "always"):
        monkeypatch.setenv("SGB_HOST_STAGE", mode)
        ho.zero_()
        hmm.check(lib.sgb_hmm_decode(ctypes.byref(P), n.size, hn.data_ptr(), hx.data_ptr(), hc.data_ptr(), 4, so.size - 1,
                                     so.ctypes.data, ho.data_ptr(), None))
        assert np.array_equal(ho.numpy(), ref), mode
    monkeypatch.delenv("SGB_HOST_STAGE")
    # a chunk with a count beyond uint16 goes over as int32 (same answer as the oracle); the other chunks stay narrow
    n2 = n.copy(); x2 = x.copy()
    n2[100000] = 70000; x2[100000] = 35000
    ref2 = oracle.decode_batch(n2, x2, cn.astype(np.int32), so, oracle.params(**cases.param_dict(ps)), oracle.MATH_DET, nthreads=4)
    assert np.array_equal(hmm.decode_arrays(n2, x2, cn, so, **_kw(ps)), ref2)
    assert np.array_equal(hmm.decode_arrays(n2.astype(np.float64), x2.astype(np.float64), cn, so, **_kw(ps)), ref2)
    # what the reference would turn into NaN / -Inf rows is an error here, with a message that names the problem
    bad = n.astype(np.float64); bad[7] = np.nan
    with pytest.raises(hmm.SignalsB200Error, match="not an integer"):
        hmm.decode_arrays(bad, x, cn, so, **_kw(ps))
    bad = x.astype(np.float64); bad[7] = 2.5
    with pytest.raises(hmm.SignalsB200Error, match="not an integer"):
        hmm.decode_arrays(n, bad, cn, so, **_kw(ps))
    bad = cn.astype(np.float64); bad[7] = 300.0
    with pytest.raises(hmm.SignalsB200Error, match="copy-number"):
        hmm.decode_arrays(n, x, bad, so, **_kw(ps))
    bad = n.astype(np.float64); bad[7] = -3.0
    with pytest.raises(hmm.SignalsB200Error, match="negative"):
        hmm.decode_arrays(bad, x, cn, so, **_kw(ps))
    bad = n.astype(np.float64); bad[7] = 2.0 ** 40
    with pytest.raises(hmm.SignalsB200Error, match="not an integer|2\\^28"):
        hmm.decode_arrays(bad, x, cn, so, **_kw(ps))
    assert np.array_equal(hmm.decode_arrays(n, x, cn, so, **_kw(ps)), ref)       # and the library still works


@pytest.mark.gpu
def test_resident_columns_several_passes_one_upload(oracle, sched):
    """sgb_cols_*: the table goes to the device once, then every pass of the reference's workflow (R/callHSCN.R:883-935:
    binomial at the default loherror, then the inferred loherror / beta-binomial; R/callASCN.R:372-414) decodes the resident
    columns -- each pass must give the oracle's paths, for any element type of the columns and of the result."""
    d = cases.small_cohort("betabinomial", n_cells=90, seed=41)
    n, x, cn, so = d["n"], d["x"], d["cn"], d["seg_off"]
    passes = [cases.PARAM_SETS[0], cases.PARAM_SETS[2], cases.PARAM_SETS[4], cases.PARAM_SETS[11], cases.PARAM_SETS[5]]
    refs = [oracle.decode_batch(n, x, cn.astype(np.int32), so, oracle.params(**cases.param_dict(ps)), oracle.MATH_DET, nthreads=4)
            for ps in passes]
    for tn, tx, tc in ((np.int32, np.int32, np.int32), (np.float64, np.float64, np.float64), (np.int32, np.int32, np.uint8),
                       (np.uint16, np.float64, np.int32)):
        with hmm.ResidentColumns(n.astype(tn), x.astype(tx), cn.astype(tc), so) as cols:
            for ps, ref in zip(passes, refs):
                for out_dtype in (np.uint8, np.float64):
                    got = cols.decode(out_dtype=out_dtype, **_kw(ps))
                    assert got.dtype == np.dtype(out_dtype) and np.array_equal(got.astype(np.uint8), ref), (ps[0], tn, out_dtype)
            # chained decode on the same columns (the plan is rebuilt with backpointer scratch), then reference again
            ch = cols.decode(decode="chained", **_kw(passes[0]))
            ref_ch = oracle.decode_batch(n, x, cn.astype(np.int32), so, oracle.params(decode="chained", **cases.param_dict(passes[0])),
                                         oracle.MATH_DET, nthreads=4)
            assert np.array_equal(ch, ref_ch)
            assert np.array_equal(cols.decode(**_kw(passes[0])), refs[0])
    # validation: what the conversion can see fails the upload, what only the device sees fails the pass
    bad = n.astype(np.float64); bad[11] = np.nan
    with pytest.raises(hmm.SignalsB200Error, match="not an integer"):
        hmm.ResidentColumns(bad, x, cn, so)
    bad = cn.astype(np.int32); bad[11] = 700
    with pytest.raises(hmm.SignalsB200Error, match="copy-number"):
        hmm.ResidentColumns(n, x, bad, so)
    bad = n.copy(); bad[11] = -4
    cols = hmm.ResidentColumns(bad, x, cn, so)
    with pytest.raises(hmm.SignalsB200Error, match="negative"):
        cols.decode(**_kw(passes[0]))
    cols.close()
    with pytest.raises(hmm.SignalsB200Error, match="released"):
        cols.decode(**_kw(passes[0]))
    # empty table and a table whose sequences do not cover every row
    e = hmm.ResidentColumns(np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert e.decode(**_kw(passes[0])).size == 0
    e.close()
    so2 = so[: so.size // 2].copy()
    with hmm.ResidentColumns(n, x, cn, so2) as cols:
        got = cols.decode(**_kw(passes[0]))
        assert np.array_equal(got[: so2[-1]], refs[0][: so2[-1]]) and not got[so2[-1]:].any()
