"""Per-group durations of k_recur by plane class, and which groups shared an SM.
    python -m signals_b200.build --variant dbg --define SGB_RECUR_TIMING ; python tools/micro/recur_timing.py cfg2 dbg"""
import os, sys, ctypes, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ["SGB_LIB"] = os.path.join(ROOT, "signals_b200", "libsignals_b200_%s.so" % (sys.argv[2] if len(sys.argv) > 2 else "dbg"))
os.environ["SGB_SCHED"] = "split"
import numpy as np
import tools.gpu_sched_compare as gc
from signals_b200 import _lib
name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
print(json.dumps(gc.run(name, "split", 3)))
lib = _lib.load()
f = lib.sgb_debug_recur_read
f.argtypes = [ctypes.c_void_p, ctypes.c_int]
buf = np.zeros(4 * 8192, dtype=np.uint64)
print("rc", f(buf.ctypes.data, buf.size))
b = buf.reshape(-1, 4)
b = b[b[:, 1] > 0]
npl = (b[:, 0] & 255).astype(int); tmax = ((b[:, 0] >> 8) & 0xffff).astype(int); smid = ((b[:, 0] >> 24) & 0xff).astype(int)
t0 = b[:, 1].astype(np.int64); t1 = b[:, 2].astype(np.int64); cyc = b[:, 3].astype(np.int64)
base = t0.min()
print("groups", len(b), "kernel span us", (t1.max() - base) / 1e3)
for p in sorted(set(npl)):
    m = npl == p
    print("npl", p, "groups", m.sum(), "Tmax max", tmax[m].max(), "dur us max %.1f mean %.1f" % ((t1 - t0)[m].max() / 1e3, (t1 - t0)[m].mean() / 1e3),
          "cycles/step mean %.0f" % (cyc[m] / np.maximum(tmax[m], 1)).mean(), 
          "end us max %.1f" % ((t1[m] - base).max() / 1e3))
order = np.argsort(-(t1 - base))[:8]
for i in order:
    print("late group: npl", npl[i], "Tmax", tmax[i], "start %.1f end %.1f cyc/step %.0f" % ((t0[i] - base) / 1e3, (t1[i] - base) / 1e3, cyc[i] / max(tmax[i], 1)))

# heavy groups: cycles/step against how many groups of which class shared their SM at the same time
import collections
per_sm = collections.defaultdict(list)
for i in range(len(b)):
    per_sm[smid[i]].append(i)
for p in (3, 2, 1):
    rows = []
    for i in np.where(npl == p)[0]:
        if tmax[i] < 300: continue
        mates = [j for j in per_sm[smid[i]] if j != i and t0[j] < t1[i] and t1[j] > t0[i]]
        rows.append((cyc[i] / tmax[i], len(mates), sum(npl[j] for j in mates)))
    rows.sort()
    print("npl", p, "cyc/step, co-resident groups, sum of their planes:", [(int(a), b_, c_) for a, b_, c_ in rows[:6]], "...", [(int(a), b_, c_) for a, b_, c_ in rows[-6:]])
