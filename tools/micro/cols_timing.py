"""Breakdown of the sgb_cols two-pass workflow on the target workload (pageable int32 in, double out)."""
import os, sys, time, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from signals_b200 import hmm, _lib
lib = _lib.load()
d = bench.make_data("target")
N = d["n"].size
pn, px, pc = d["n"].astype(np.int32), d["x"].astype(np.int32), d["cn"].astype(np.int32)
so = d["seg_off"]
po = np.zeros(N, dtype=np.float64)
pu = np.zeros(N, dtype=np.uint8)
P = hmm.make_params(model="hscn", likelihood="binomial", rho=0.0, **bench.PARAMS)
for rep in range(3):
    t = [time.perf_counter()]
    h = ctypes.c_void_p()
    hmm.check(lib.sgb_cols_upload(N, pn.ctypes.data, _lib.T_I32, px.ctypes.data, _lib.T_I32, pc.ctypes.data, _lib.T_I32, so.size - 1, so.ctypes.data, ctypes.byref(h)))
    t.append(time.perf_counter())
    hmm.check(lib.sgb_cols_decode(h, ctypes.byref(P), po.ctypes.data, _lib.T_F64)); t.append(time.perf_counter())
    hmm.check(lib.sgb_cols_decode(h, ctypes.byref(P), po.ctypes.data, _lib.T_F64)); t.append(time.perf_counter())
    hmm.check(lib.sgb_cols_decode(h, ctypes.byref(P), pu.ctypes.data, _lib.T_U8)); t.append(time.perf_counter())
    lib.sgb_cols_free(h); t.append(time.perf_counter())
    print("upload %.1f  decode#1(double) %.1f  decode#2(double) %.1f  decode#3(uint8) %.1f  free %.1f ms" % tuple((b - a) * 1e3 for a, b in zip(t, t[1:])))
import torch
for rep in range(3):
    t0 = time.perf_counter(); pl = hmm.DevicePlan(N, so, n_states=12); torch.cuda.synchronize(); t1 = time.perf_counter()
    pl.close(); t2 = time.perf_counter()
    print("DevicePlan create %.1f ms  close %.1f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3))
t0 = time.perf_counter(); a = torch.empty(72 << 20, dtype=torch.uint8).pin_memory(); t1 = time.perf_counter()
print("pin 72 MB: %.1f ms" % ((t1 - t0) * 1e3))
