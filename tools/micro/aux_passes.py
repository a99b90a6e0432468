"""Runs the post-decode / between-pass device passes once on a 4,000-cell table (for an ncu launch list):
pass_stats2, qc_per_cell, fix_assignments, postprocess."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
from signals_b200 import hmm, table, synth
n_cells = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
d = synth.make_cohort(n_cells, seed=9)
n, x, cn, so = d["n"], d["x"], d["cn"], d["seg_off"]
path = hmm.decode_arrays(n, x, cn, so, selftransitionprob=0.95)
seg_len = np.diff(so)
cell_of_row = np.repeat(d["cell"], seg_len).astype(np.int32)
print("rows", n.size)
table.pass_stats2_arrays(n, x, cn, path)
table.qc_per_cell_arrays(n, x, cn, path, cell_of_row, n_cells)
pp = hmm.postprocess_arrays(cn, path)
A, B = pp["A"], pp["B"]
table.fix_assignments_arrays(cn.astype(np.int32), A, B, (n - x).astype(np.int32), x, so, 0.02)
print("done")
