"""The table passes either side of the HMM hot path: device results through the C ABI against the
NumPy / pandas / C restatements in oracle/ (which are themselves checked on hand-computed cases here, without a GPU)."""
import numpy as np
import pytest

import cases
from oracle import table_np as tnp
from signals_b200 import hmm, synth, table


def _joined(seed, n_cells=30, dup_rate=0.4, shuffle=True):
    """rows as CNbins[haplotypes] would give them: several haplotype-block rows per (cell, bin), unsorted"""
    rng = np.random.default_rng(seed)
    d = synth.make_cohort(n_cells, lens=(40, 35, 12, 9, 30), seed=seed, ragged_drop=0.1)
    seg_len = np.diff(d["seg_off"])
    cell = np.repeat(d["cell"], seg_len); chrom = np.repeat(d["chrom"], seg_len)
    pos = np.arange(d["n"].size) - np.repeat(d["seg_off"][:-1], seg_len)
    # drop most bins of a few cells / chromosomes so that the minbins / minbinschr filters bite
    keep = np.ones(cell.size, bool)
    keep &= ~((cell == 3) & (rng.random(cell.size) < 0.8))
    keep &= ~((cell == 7) & (chrom == 2) & (pos > 5))
    keep &= ~((cell == 9) & (chrom == 4) & (pos > 3))           # chromosome 4 plays "X": never filters
    cell, chrom, pos = cell[keep], chrom[keep], pos[keep]
    n, x, cn = d["n"][keep], d["x"][keep], d["cn"][keep].astype(np.int32)
    rep = 1 + (rng.random(cell.size) < dup_rate) + (rng.random(cell.size) < dup_rate / 3)
    idx = np.repeat(np.arange(cell.size), rep)
    part = rng.random(idx.size)
    tot = np.maximum(0, np.rint(n[idx] * part / np.maximum(1, rep[idx])).astype(np.int64))
    b = rng.binomial(tot, np.clip(x[idx] / np.maximum(n[idx], 1), 0, 1))
    rows = dict(cell=cell[idx], chr=chrom[idx], start=pos[idx] * 500000 + 1, end=(pos[idx] + 1) * 500000, state=cn[idx],
                copy=cn[idx] + 0.25, alleleA=tot - b, alleleB=b, totalcounts=tot)
    if shuffle:
        p = rng.permutation(idx.size)
        rows = {k: v[p] for k, v in rows.items()}
    return rows


def test_oracle_table_restatements_on_hand_cases():
    # order + segments
    perm, seg = tnp.table_order([1, 0, 1, 0, 0], [0, 1, 0, 0, 1], [5, 9, 2, 7, 3])
    assert perm.tolist() == [3, 4, 1, 2, 0] and seg.tolist() == [0, 1, 3, 5]
    # combine: two rows of one bin are summed, zero-count rows dropped, small cells dropped
    g = tnp.combine_bafcn(cell=[0, 0, 0, 1], chr_=[0, 0, 0, 0], start=[1, 1, 2, 1], end=[2, 2, 3, 2], state=[2, 2, 2, 3],
                          copy=[2.1, 2.1, 2.0, 3.0], alleleA=[1, 2, 0, 4], alleleB=[3, 1, 0, 1], totalcounts=[4, 3, 0, 5],
                          chrX_code=-1, filtern=0, minbins=0, minbinschr=0)
    assert g[["cell_id", "start", "alleleA", "alleleB", "totalcounts"]].values.tolist() == [[0, 1, 3, 4, 7], [1, 1, 4, 1, 5]]
    # Tarone's Z: the expanded-sums form equals the vector form
    rng = np.random.default_rng(0)
    n = rng.integers(5, 200, 5000); m = rng.binomial(n, 0.4)
    z = tnp.tarones_Z(m, n)
    z2 = tnp.tarones_Z_from_sums(m.sum(), n.sum(), (m.astype(float) ** 2).sum(), (m * n).astype(float).sum(), (n.astype(float) ** 2).sum())
    assert abs(z - z2) < 1e-9 * max(1.0, abs(z))
    # qc_per_cell: one cell, two states of 12 bins each with known medians
    A = np.array([1] * 12 + [2] * 12); B = np.array([1] * 12 + [0] * 12)
    x = np.array([5] * 12 + [1] * 12); nn = np.array([10] * 24)
    q = tnp.qc_per_cell(np.zeros(24, int), A, B, x, nn, 1)
    assert abs(q[0] - (0.5 * abs(0.5 - 0.5) + 0.5 * abs(0.0 - 0.1))) < 1e-15


def test_oracle_fix_assignments_as_written(oracle):
    # a singleton takes its successor's B whatever the likelihoods say (R/callHSCN.R:1114 makes LLprev == LLafter)
    state = np.array([2, 2, 2, 2, 2, 2]); A = np.array([1, 1, 2, 1, 1, 1]); B = state - A
    aA = np.array([10, 11, 20, 9, 10, 12]); aB = np.array([10, 9, 0, 11, 10, 8])
    Ao, Bo = oracle.fix_assignments(state, A, B, aA, aB, np.array([0, 6]), 0.02)
    assert Bo[2] == B[3]                                      # the 2|0 singleton became 1|1 like its successor
    # second block: a run whose allele totals clearly favour the other assignment is swapped
    state = np.full(8, 3); A = np.full(8, 1); B = np.full(8, 2)
    aA = np.full(8, 40); aB = np.full(8, 20)                  # alleleA dominates although B = 2 > A = 1 was called
    Ao, Bo = oracle.fix_assignments(state, A, B, aA, aB, np.array([0, 8]), 0.02)
    assert (Ao == 2).all() and (Bo == 1).all()


@pytest.mark.gpu
def test_table_order_and_combine_bit_exact(oracle):
    for seed in (1, 2, 3):
        r = _joined(seed)
        perm, seg = table.table_order(r["cell"], r["chr"], r["start"])
        operm, oseg = tnp.table_order(r["cell"], r["chr"], r["start"])
        assert np.array_equal(perm, operm) and np.array_equal(seg, oseg)
        for filtern, minbins, minbinschr in ((0, 100, 10), (0, 0, 0), (3, 60, 6)):
            got = table.combine_bafcn_arrays(r["cell"], r["chr"], r["start"], r["state"], r["alleleA"], r["alleleB"], r["totalcounts"],
                                             int(r["cell"].max()) + 1, 5, filtern, minbins, minbinschr, chrX_code=4)
            ref = tnp.combine_bafcn(r["cell"], r["chr"], r["start"], r["end"], r["state"], r["copy"], r["alleleA"], r["alleleB"],
                                    r["totalcounts"], 4, filtern, minbins, minbinschr)
            assert len(ref) == got["cell"].size, (seed, filtern, minbins, len(ref), got["cell"].size)
            for k, col in (("cell", "cell_id"), ("chr", "chr"), ("start", "start"), ("state", "state"), ("alleleA", "alleleA"),
                           ("alleleB", "alleleB"), ("totalcounts", "totalcounts")):
                assert np.array_equal(got[k], ref[col].to_numpy()), (seed, k)
            assert np.array_equal(r["end"][got["first_row"]], ref["end"].to_numpy())
            _, oseg2 = tnp.table_order(ref["cell_id"].to_numpy(), ref["chr"].to_numpy(), ref["start"].to_numpy())
            assert np.array_equal(got["seg_off"], oseg2 if len(ref) else np.array([0]))
    # conflicting states inside one key are refused; so are codes outside their declared range
    r = _joined(5, shuffle=False)
    bad = r["state"].copy()
    dup = np.nonzero((r["cell"][1:] == r["cell"][:-1]) & (r["start"][1:] == r["start"][:-1]) & (r["chr"][1:] == r["chr"][:-1]))[0][0]
    bad[dup] += 1
    with pytest.raises(hmm.SignalsB200Error, match="disagree"):
        table.combine_bafcn_arrays(r["cell"], r["chr"], r["start"], bad, r["alleleA"], r["alleleB"], r["totalcounts"], 30, 5)
    with pytest.raises(hmm.SignalsB200Error, match="code outside"):
        table.combine_bafcn_arrays(r["cell"], r["chr"], r["start"], r["state"], r["alleleA"], r["alleleB"], r["totalcounts"], 5, 5)
    # empty table
    e = np.zeros(0, np.int32)
    got = table.combine_bafcn_arrays(e, e, e, e, e, e, e, 1, 1)
    assert got["cell"].size == 0 and got["seg_off"].tolist() == [0]


@pytest.mark.gpu
def test_table_order_at_scale():
    """2,000 cells x 5,700 bins (11.4 M rows) shuffled: the permutation restores the sorted order and the boundaries are the 46,000 sequences"""
    d = synth.make_cohort(2000, seed=9)
    seg_len = np.diff(d["seg_off"])
    cell = np.repeat(d["cell"], seg_len); chrom = np.repeat(d["chrom"], seg_len)
    pos = (np.arange(d["n"].size) - np.repeat(d["seg_off"][:-1], seg_len)).astype(np.int32) * 500000
    p = np.random.default_rng(3).permutation(cell.size)
    perm, seg = table.table_order(cell[p], chrom[p], pos[p])
    assert np.array_equal(p[perm], np.arange(cell.size)) and np.array_equal(seg, d["seg_off"])


@pytest.mark.gpu
def test_between_pass_sums_tarone_and_qc(oracle):
    d = cases.small_cohort("betabinomial", n_cells=60, seed=41)
    ps = cases.PARAM_SETS[0]
    kw = dict(model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12, selftransitionprob=0.95)
    path = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **kw)
    tab = oracle.postprocess(d["cn"].astype(np.int32), path)
    A, B = tab["A"], tab["B"]
    cell_of_row = np.repeat(d["cell"], np.diff(d["seg_off"]))
    keep_cells = np.random.default_rng(1).random(60) < 0.8
    row_keep = keep_cells[cell_of_row]
    h = np.zeros((12, 12), np.int64); sel = (B > 0) & row_keep
    np.add.at(h, (A[sel], B[sel]), 1)
    modal = np.unravel_index(np.argmax(h), h.shape)
    got = table.pass_stats2_arrays(d["n"], d["x"], d["cn"], path, 12, sel=modal, row_keep=row_keep)
    assert np.array_equal(got["hist"], h)
    ah = (B == 0) & (A > 0) & row_keep
    assert got["ahom"].tolist() == [float(d["x"][ah].sum()), float(d["n"][ah].sum()), float(ah.sum())]
    ms = (A == modal[0]) & (B == modal[1]) & row_keep
    m, n = d["x"][ms].astype(np.int64), d["n"][ms].astype(np.int64)
    assert got["sel"].tolist() == [float(ms.sum()), float(m.sum()), float(n.sum()), float((m * m).sum()), float((m * n).sum()), float((n * n).sum())]
    z_dev, z_ref = table.tarones_Z_from_sums(got["sel"]), tnp.tarones_Z(m, n)
    assert abs(z_dev - z_ref) <= 1e-9 * max(1.0, abs(z_ref))
    assert abs(table.tarones_Z(m, n) - z_ref) <= 1e-9 * max(1.0, abs(z_ref))
    infl = min(got["ahom"][0] / got["ahom"][1], 0.035)
    assert abs(infl - tnp.infloherror(d["x"], d["n"], A, B, row_keep)) < 1e-12
    # qc_per_cell
    q = table.qc_per_cell_arrays(d["n"], d["x"], d["cn"], path, cell_of_row, 60)
    qr = tnp.qc_per_cell(cell_of_row, A, B, d["x"], d["n"], 60)
    assert np.array_equal(np.isnan(q), np.isnan(qr))
    ok = ~np.isnan(qr)
    assert np.allclose(q[ok], qr[ok], rtol=1e-12, atol=1e-15)


@pytest.mark.gpu
@pytest.mark.parametrize("lik,rho", [("binomial", 0.0), ("betabinomial", 0.02)])
def test_fix_assignments_bit_exact(oracle, lik, rho):
    d = cases.small_cohort("binomial", n_cells=40, seed=51)
    kw = dict(model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12, selftransitionprob=0.5)   # low stp: singletons
    path = hmm.decode_arrays(d["n"], d["x"], d["cn"], d["seg_off"], **kw)
    tab = oracle.postprocess(d["cn"].astype(np.int32), path)
    A, B, state = tab["A"], tab["B"], d["cn"].astype(np.int32)
    aB = d["x"]; aA = d["n"] - d["x"]
    single = np.zeros(A.size, bool)
    single[1:-1] = ((A[1:-1] != A[:-2]) | (B[1:-1] != B[:-2])) & ((A[1:-1] != A[2:]) | (B[1:-1] != B[2:]))
    assert single.sum() > 20
    Ao, Bo = table.fix_assignments_arrays(state, A, B, aA, aB, d["seg_off"], 0.03, lik, rho)
    Ar, Br = oracle.fix_assignments(state, A, B, aA, aB, d["seg_off"], 0.03, lik, rho, math=oracle.MATH_DET)
    assert np.array_equal(Ao, Ar) and np.array_equal(Bo, Br), int((Bo != Br).sum())
    assert ((Ao != A) | (Bo != B)).sum() > 0                  # the pass does change assignments on this input
    # the libm flavour (the natural restatement of R's arithmetic) may only differ where the two densities tie
    Al, Bl = oracle.fix_assignments(state, A, B, aA, aB, d["seg_off"], 0.03, lik, rho, math=oracle.MATH_LIBM)
    assert (Bl != Bo).mean() < 1e-3
