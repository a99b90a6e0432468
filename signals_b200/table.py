"""Host-side mirror of the reference's table passes either side of the HMM hot path, on the GPU:

    table_order(cell, chr, start)                      .[order(cell_id, chr, start)] + sequence boundaries, R/callHSCN.R:241-247
    combine_bafcn_arrays(...) / combineBAFCN(joined)   the aggregation and filters of combineBAFCN, R/combinedata.R:309-325
    pass_stats2_arrays(...)                            infloherror / tarones_Z / modal-state sums, R/callHSCN.R:609-616, 883-900
    tarones_Z(m, n)                                    R/callHSCN.R:609-616
    qc_per_cell_arrays(...)                            average_distance of qc_per_cell, R/util.R:667-691
    fix_assignments_arrays(...)                        fix_assignments, R/callHSCN.R:1072-1286

Everything numeric runs in libsignals_b200.so; there is no CPU implementation here.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np

from . import _lib
from ._lib import check
from .hmm import _as_counts, _ptr


def _i32(a, name):
    return _as_counts(a, name)


def table_order(cell, chr_, start, n_cells=None, n_chr=None):
    cell = _i32(cell, "cell"); chr_ = _i32(chr_, "chr"); start = _i32(start, "start")
    N = cell.size
    n_cells = int(cell.max()) + 1 if n_cells is None and N else (n_cells or 1)
    n_chr = int(chr_.max()) + 1 if n_chr is None and N else (n_chr or 1)
    perm = np.zeros(N, np.int32); seg = np.zeros(N + 1, np.int64); nseg = ctypes.c_int64(0)
    check(_lib.load().sgb_table_order(N, _ptr(cell), _ptr(chr_), _ptr(start), n_cells, n_chr, _ptr(perm), ctypes.byref(nseg), _ptr(seg)))
    return perm, seg[: nseg.value + 1].copy()


def combine_bafcn_arrays(cell, chr_, start, state, alleleA, alleleB, totalcounts, n_cells, n_chr, filtern=0.0, minbins=100,
                         minbinschr=10, chrX_code=-1):
    cols = [_i32(v, k) for v, k in ((cell, "cell"), (chr_, "chr"), (start, "start"), (state, "state"), (alleleA, "alleleA"),
                                    (alleleB, "alleleB"), (totalcounts, "totalcounts"))]
    N = cols[0].size
    outs = [np.zeros(N, np.int32) for _ in range(8)]
    seg = np.zeros(N + 1, np.int64); nrow = ctypes.c_int64(0); nseg = ctypes.c_int64(0)
    check(_lib.load().sgb_combine_bafcn(N, *[_ptr(c) for c in cols], int(n_cells), int(n_chr), float(filtern), int(minbins),
                                        int(minbinschr), int(chrX_code), ctypes.byref(nrow), *[_ptr(o) for o in outs],
                                        ctypes.byref(nseg), _ptr(seg)))
    k = nrow.value
    names = ("first_row", "cell", "chr", "start", "state", "alleleA", "alleleB", "totalcounts")
    res = {nm: o[:k].copy() for nm, o in zip(names, outs)}
    res["seg_off"] = seg[: nseg.value + 1].copy()
    return res


def combineBAFCN(joined, filtern=0, minbins=100, minbinschr=10):
    """The part of combineBAFCN after the CNbins x haplotypes join (R/combinedata.R:309-325) on a DataFrame with columns
    chr, start, end, cell_id, state, copy, alleleA, alleleB, totalcounts.  Returns the CNBAF table ordered by
    (cell_id, chr, start) (the reference keeps first-appearance order; .callHaplotypeSpecificCN_ sorts it this way)."""
    import pandas as pd
    cells, cell_code = np.unique(joined["cell_id"].to_numpy().astype(str), return_inverse=True)
    chrs, chr_code = np.unique(joined["chr"].to_numpy().astype(str), return_inverse=True)
    chrx = int(np.nonzero(chrs == "X")[0][0]) if (chrs == "X").any() else -1
    r = combine_bafcn_arrays(cell_code, chr_code, joined["start"].to_numpy(), joined["state"].to_numpy(), joined["alleleA"].to_numpy(),
                             joined["alleleB"].to_numpy(), joined["totalcounts"].to_numpy(), len(cells), len(chrs), filtern, minbins,
                             minbinschr, chrx)
    f = r["first_row"]
    out = pd.DataFrame({"chr": chrs[r["chr"]], "start": r["start"], "end": joined["end"].to_numpy()[f], "cell_id": cells[r["cell"]],
                        "state": r["state"], "copy": joined["copy"].to_numpy()[f], "alleleA": r["alleleA"], "alleleB": r["alleleB"],
                        "totalcounts": r["totalcounts"]})
    out["BAF"] = out["alleleB"] / out["totalcounts"]
    return out


def pass_stats2_arrays(n, x, cn, path, n_states=12, sel=(1, 1), row_keep=None):
    """dict(ahom=(sum alleleB, sum totalcounts, rows), sel=(rows, sum m, sum n, sum m^2, sum mn, sum n^2), hist=A x B)"""
    n = _i32(n, "totalcounts"); x = _i32(x, "alleleB")
    cn = np.asarray(cn)
    if cn.dtype != np.uint8:
        cn = _as_counts(cn, "state")
    cn = np.ascontiguousarray(cn)
    path = np.ascontiguousarray(path, dtype=np.uint8)
    keep = None if row_keep is None else np.ascontiguousarray(row_keep, dtype=np.uint8)
    out = np.zeros(9, np.float64); hist = np.zeros(int(n_states) ** 2, np.int64)
    check(_lib.load().sgb_pass_stats2(n.size, _ptr(n), _ptr(x), _ptr(cn), cn.dtype.itemsize, _ptr(path),
                                      _ptr(keep) if keep is not None else None, int(n_states), int(sel[0]), int(sel[1]), _ptr(out), _ptr(hist)))
    return dict(ahom=out[:3].copy(), sel=out[3:].copy(), hist=hist.reshape(int(n_states), int(n_states)))


def tarones_Z_from_sums(sel):
    rows, sm, sn, smm, smn, snn = (float(v) for v in sel)
    p = sm / sn
    S = (smm - 2.0 * p * smn + p * p * snn) / (p * (1.0 - p))
    return (S - sn) / math.sqrt(2.0 * (snn - sn))


def tarones_Z(alleleA, totalcounts):
    """tarones_Z(alleleA, totalcounts) of the reference (R/callHSCN.R:609-616) for two vectors"""
    m = _i32(alleleA, "alleleA"); n = _i32(totalcounts, "totalcounts")
    ones = np.ones(m.size, np.uint8)
    r = pass_stats2_arrays(n, m, ones, np.zeros(m.size, np.uint8), n_states=2, sel=(1, 0))
    return tarones_Z_from_sums(r["sel"])


def qc_per_cell_arrays(n, x, cn, path, cell_of_row, n_cells):
    n = _i32(n, "totalcounts"); x = _i32(x, "alleleB"); cell_of_row = _i32(cell_of_row, "cell")
    cn = np.asarray(cn)
    if cn.dtype != np.uint8:
        cn = _as_counts(cn, "state")
    cn = np.ascontiguousarray(cn)
    path = np.ascontiguousarray(path, dtype=np.uint8)
    out = np.zeros(int(n_cells), np.float64)
    check(_lib.load().sgb_qc_per_cell(n.size, _ptr(n), _ptr(x), _ptr(cn), cn.dtype.itemsize, _ptr(path), _ptr(cell_of_row), int(n_cells), _ptr(out)))
    return out


def fix_assignments_arrays(state, A, B, alleleA, alleleB, seg_off, loherror, likelihood="binomial", rho=0.0):
    cols = [_i32(v, k) for v, k in ((state, "state"), (A, "A"), (B, "B"), (alleleA, "alleleA"), (alleleB, "alleleB"))]
    seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
    N = cols[0].size
    Ao = np.zeros(N, np.int32); Bo = np.zeros(N, np.int32)
    check(_lib.load().sgb_fix_assignments(N, *[_ptr(c) for c in cols], seg_off.size - 1, _ptr(seg_off), float(loherror),
                                          _lib.LIK_BETABINOMIAL if likelihood == "betabinomial" else _lib.LIK_BINOMIAL, float(rho),
                                          _ptr(Ao), _ptr(Bo)))
    return Ao, Bo
