#!/usr/bin/env python
"""Regenerate tests/golden/hmm_golden.npz.

The reference has no golden vectors for this path and cannot be run here (no R, Rcpp, VGAM):
these fixtures are produced by the CPU oracle (oracle/oracle.c) on the committed inputs of
tests/cases.py, so they pin the oracle against drift and give the GPU tests fixed expected
outputs.  Run from the repository root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
from oracle import cpu_oracle as co  # noqa: E402

out = {}
seqs = cases.edge_sequences()
n, x, cn, seg = cases.concat_sequences(seqs)
out["edge_n"], out["edge_x"], out["edge_cn"], out["edge_seg"] = n, x, cn, seg
for ps in cases.PARAM_SETS:
    P = co.params(**cases.param_dict(ps))
    for math, tag in ((co.MATH_DET, "det"), (co.MATH_LIBM, "libm")):
        out["edge_path_%s_%s" % (ps[0], tag)] = co.decode_batch(n, x, cn, seg, P, math=math)
    # chained decode too
    Pc = co.params(**cases.param_dict(ps), decode="chained")
    out["edge_chained_%s_det" % ps[0]] = co.decode_batch(n, x, cn, seg, Pc, math=co.MATH_DET)
    # emission matrices, sequence by sequence, column-major blocks like the C ABI's loglik_out
    blocks = []
    for s in range(len(seg) - 1):
        a, b = seg[s], seg[s + 1]
        l = co.emission(n[a:b], x[a:b], cn[a:b], P, math=co.MATH_DET)       # T x S
        blocks.append(np.asfortranarray(l).ravel(order="F"))
    out["edge_l_%s_det" % ps[0]] = np.concatenate(blocks)

for lik in ("binomial", "betabinomial"):
    d = cases.small_cohort(lik)
    cn32 = d["cn"].astype(np.int32)
    for ps in cases.PARAM_SETS[:5]:
        if (ps[2] == "betabinomial") != (lik == "betabinomial"):
            continue
        P = co.params(**cases.param_dict(ps))
        out["cohort_path_%s_det" % ps[0]] = co.decode_batch(d["n"], d["x"], cn32, d["seg_off"], P, math=co.MATH_DET)

dst = os.path.join(ROOT, "tests", "golden", "hmm_golden.npz")
np.savez_compressed(dst, **out)
print("wrote", dst, os.path.getsize(dst), "bytes,", len(out), "arrays")
