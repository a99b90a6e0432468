"""HBM-roofline lines for the input-side kernels: device time of sgb_table_order / sgb_combine_bafcn on a
shuffled long table, against the bytes the radix sort has to move.
    python tools/gpu_table_bench.py [n_cells]"""
import ctypes, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from signals_b200 import _lib, synth, table

n_cells = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
d = synth.make_cohort(n_cells, seed=9)
seg_len = np.diff(d["seg_off"])
cell = np.repeat(d["cell"], seg_len).astype(np.int32); chrom = np.repeat(d["chrom"], seg_len).astype(np.int32)
pos = ((np.arange(d["n"].size) - np.repeat(d["seg_off"][:-1], seg_len)) * 500000).astype(np.int32)
p = np.random.default_rng(3).permutation(cell.size)
N = cell.size
lib = _lib.load()
ms = ctypes.c_double(0.0)
try:
    hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    hbm = 6650.0
def bits(v):
    b = 1
    while (1 << b) <= v: b += 1
    return b
key_bits = bits(int(pos.max())) + bits(22) + bits(n_cells - 1) + 1       # position field sized from the largest start
passes = (key_bits + 7) // 8
# per pass: histogram read (8 B key) + scatter read (8 + 4) + scatter write (8 + 4) = 32 B per row; + keys written once (12 B) and inputs read (12 B)
alg = N * (passes * 32 + 24)
for rep in range(3):
    perm, seg = table.table_order(cell[p], chrom[p], pos[p], n_cells, 23)
    lib.sgb_table_last_ms(ctypes.byref(ms))
assert np.array_equal(p[perm], np.arange(N)) and np.array_equal(seg, d["seg_off"])
print(json.dumps(dict(op="sgb_table_order", rows=N, key_bits=key_bits, radix_passes=passes, device_ms=ms.value, rows_per_s=N / (ms.value * 1e-3),
                      algorithmic_bytes=alg, achieved_gbs=alg / (ms.value * 1e-3) / 1e9, hbm_peak_gbs=hbm, frac=alg / (ms.value * 1e-3) / 1e9 / hbm)))
st = d["cn"][p].astype(np.int32); tot = d["n"][p]; b = d["x"][p]; a = tot - b
for rep in range(3):
    r = table.combine_bafcn_arrays(cell[p], chrom[p], pos[p], st, a, b, tot, n_cells, 23, 0, 100, 10, 22)
    lib.sgb_table_last_ms(ctypes.byref(ms))
assert r["cell"].size == N and np.array_equal(r["totalcounts"], d["n"]) and np.array_equal(r["seg_off"], d["seg_off"])
alg2 = N * (passes * 32 + 24 + 7 * 4 + 8 * 4 + 8 * 4 + 16)        # + reduce reads 7 columns, writes / compacts 8 columns twice, flags
print(json.dumps(dict(op="sgb_combine_bafcn", rows=N, device_ms=ms.value, rows_per_s=N / (ms.value * 1e-3), algorithmic_bytes=alg2,
                      achieved_gbs=alg2 / (ms.value * 1e-3) / 1e9, hbm_peak_gbs=hbm, frac=alg2 / (ms.value * 1e-3) / 1e9 / hbm)))
