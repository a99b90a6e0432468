"""NumPy / pandas restatement of the table passes either side of the HMM hot path (TEST INFRASTRUCTURE ONLY, see
oracle/__init__.py).  Each function cites the reference lines it follows (/root/reference, signals 0.14.1).
Integer results must match the device bit for bit; the two floating-point ones (tarones_Z, qc_per_cell) are sums of
doubles taken in another order than the device's, compared within 1e-12 relative."""
import numpy as np
import pandas as pd


def table_order(cell, chr_, start):
    """.[order(cell_id, chr, start)]  (R/callHSCN.R:241-242) on integer codes; data.table's order is stable.
    Returns (perm, seg_off): the (cell, chr) runs replace dplyr::filter(CNBAF, cell_id == cell) (:247,266) and the
    per-chromosome loop of assignHaplotypeHMM (:150-154)."""
    perm = np.lexsort((start, chr_, cell)).astype(np.int32)          # lexsort is stable; last key is primary
    c, h = np.asarray(cell)[perm], np.asarray(chr_)[perm]
    change = np.ones(perm.size, bool)
    if perm.size > 1:
        change[1:] = (c[1:] != c[:-1]) | (h[1:] != h[:-1])
    seg_off = np.concatenate((np.nonzero(change)[0], [perm.size])).astype(np.int64)
    return perm, seg_off


def combine_bafcn(cell, chr_, start, end, state, copy, alleleA, alleleB, totalcounts, chrX_code, filtern=0, minbins=100,
                  minbinschr=10):
    """combineBAFCN after the join (R/combinedata.R:309-325).  Returns a DataFrame sorted by (cell, chr, start)."""
    df = pd.DataFrame(dict(chr=chr_, start=start, end=end, cell_id=cell, state=state, copy=copy, alleleA=alleleA,
                           alleleB=alleleB, totalcounts=totalcounts))
    df = df[df["totalcounts"] > filtern]                                              # :310
    g = df.groupby(["chr", "start", "end", "cell_id", "state", "copy"], sort=False, as_index=False)[
        ["alleleA", "alleleB", "totalcounts"]].sum()                                  # :311
    g["BAF"] = g["alleleB"] / g["totalcounts"]                                        # :312
    g["n"] = g.groupby("cell_id")["cell_id"].transform("size")                        # :315
    g = g[g["n"] > minbins].drop(columns="n")                                         # :316-317
    g["n"] = g.groupby(["chr", "cell_id"])["chr"].transform("size")                   # :318
    g.loc[g["chr"] == chrX_code, "n"] = 100                                           # :319
    g["ncell"] = g.groupby("cell_id")["n"].transform("min")                           # :320
    g = g[g["ncell"] > minbinschr].drop(columns=["n", "ncell"])                       # :322-324
    return g.sort_values(["cell_id", "chr", "start"], kind="stable").reset_index(drop=True)


def tarones_Z(m, n):
    """R/callHSCN.R:609-616, evaluated the way R does (double-precision vectors)"""
    m = np.asarray(m, np.float64); n = np.asarray(n, np.float64)
    p_hat = m.sum() / n.sum()
    S = ((m - n * p_hat) ** 2 / (p_hat * (1 - p_hat))).sum()
    return (S - n.sum()) / np.sqrt(2 * (n * (n - 1)).sum())


def tarones_Z_from_sums(sm, sn, smm, smn, snn):
    """the same statistic from the five sums the device returns (expanding the square)"""
    p = sm / sn
    S = (smm - 2.0 * p * smn + p * p * snn) / (p * (1.0 - p))
    return (S - sn) / np.sqrt(2.0 * (snn - sn))


def infloherror(alleleB, totalcounts, A, B, keep_rows=None, maxloherror=0.035):
    """weighted.mean(BAF, w = totalcounts) over state_phase == "A-Hom" rows of ascn_filt, capped (R/callHSCN.R:895-900)"""
    sel = (np.asarray(B) == 0) & (np.asarray(A) > 0)
    if keep_rows is not None:
        sel &= np.asarray(keep_rows, bool)
    x = np.asarray(alleleB, np.float64)[sel]; n = np.asarray(totalcounts, np.float64)[sel]
    return min(float(((x / n) * n).sum() / n.sum()), maxloherror)


def qc_per_cell(cell, A, B, alleleB, totalcounts, n_cells):
    """average_distance of qc_per_cell (R/util.R:675-691); NaN for cells the reference would have no row for"""
    df = pd.DataFrame(dict(cell_id=cell, A=A, B=B, BAF=np.asarray(alleleB, np.float64) / np.asarray(totalcounts, np.float64)))
    df = df[~((df["A"] == 0) & (df["B"] == 0))]                                       # state_AS_phased != "0|0"
    df["n"] = df.groupby(["B", "A", "cell_id"])["A"].transform("size")
    ntot = len(df)
    df["frac"] = df["n"] / ntot
    df = df[df["n"] > 10]
    s = df.groupby(["B", "A", "n", "frac", "cell_id"], as_index=False)["BAF"].median()
    s["expBAF"] = s["B"] / (s["B"] + s["A"])
    s["distance"] = np.sqrt((s["expBAF"] - s["BAF"]) ** 2)
    out = np.full(n_cells, np.nan)
    for c, grp in s.groupby("cell_id"):
        out[int(c)] = float((grp["distance"] * grp["frac"]).sum() / grp["frac"].sum())
    return out
