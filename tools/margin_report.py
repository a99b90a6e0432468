#!/usr/bin/env python
"""Decision-margin report (last row): how many bins of a cohort are decoded on a margin so thin that a
different but equally accurate math library (R's own dbinom vs any lgamma formula: ~1e-14 relative) could flip them.

For every bin t >= 1 of every sequence the reference's decode reads two argmaxes (src/code.cpp:29-31, 39-41):
the column winner i1 = which_max(T1[, t]) and its best predecessor which_max(T1[, t-1] + trans[, i1] + e[t, i1]).
The margin of a bin is the smaller of the two gaps between the best and the second-best candidate.  CPU only
(uses the oracle's emission matrix); prints counts below 1e-9 / 1e-12 / exact ties, split by total copy number 0.

    python tools/margin_report.py [n_cells] [hscn|ascn] [binomial|betabinomial]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def margins(l, tr, init):
    """l: T x S emission (bins x states), tr: S x S log transition [from, to]; returns per-bin margin for t = 1..T-1"""
    T, S = l.shape
    out = np.full(T, np.inf)
    t1 = init + l[0]
    for t in range(1, T):
        cand = (t1[:, None] + tr) + l[t][None, :]            # [from, to], the reference's operation order
        new = cand.max(axis=0)
        i1 = int(np.argmax(new))
        col = np.sort(new)[::-1]
        pre = np.sort(cand[:, i1])[::-1]
        g_col = col[0] - col[1] if S > 1 and np.isfinite(col[0]) else np.inf
        g_pre = pre[0] - pre[1] if S > 1 and np.isfinite(pre[0]) else np.inf
        out[t] = min(g_col if not np.isnan(g_col) else 0.0, g_pre if not np.isnan(g_pre) else 0.0)
        t1 = new
    return out


def report(n_cells=40, model="hscn", likelihood="binomial", seed=11):
    from oracle import cpu_oracle as co
    from signals_b200 import synth
    d = synth.make_cohort(n_cells, lens=synth.LENS_DLPBINS, seed=seed, likelihood=likelihood)
    P = co.params(model=model, likelihood=likelihood, S=12, rho=0.02 if likelihood == "betabinomial" else 0.0,
                  selftransitionprob=0.95)
    tr = co.transition(P, co.MATH_LIBM)
    cn = d["cn"].astype(np.int32)
    so = d["seg_off"]
    m = np.full(d["n"].size, np.inf)
    with np.errstate(invalid="ignore"):
        for s in range(so.size - 1):
            a, b = int(so[s]), int(so[s + 1])
            l = co.emission(d["n"][a:b], d["x"][a:b], cn[a:b], P, co.MATH_LIBM)
            m[a:b] = margins(np.asarray(l), np.asarray(tr), -np.log(12.0))
    dec = np.ones(m.size, bool)
    dec[so[:-1]] = False                                      # bin 0 of a sequence takes no decision of its own
    res = {"bins": int(dec.sum())}
    for name, thr in (("lt_1e-9", 1e-9), ("lt_1e-12", 1e-12), ("exact_tie", 0.0)):
        hit = dec & ((m <= thr) if thr == 0.0 else (m < thr))
        res[name] = int(hit.sum())
        res[name + "_at_cn0"] = int((hit & (cn == 0)).sum())
        res[name + "_next_to_cn0"] = int((hit & ((cn == 0) | (np.roll(cn, 1) == 0))).sum())
    return res


if __name__ == "__main__":
    a = sys.argv[1:]
    print(report(int(a[0]) if a else 40, a[1] if len(a) > 1 else "hscn", a[2] if len(a) > 2 else "binomial"))
