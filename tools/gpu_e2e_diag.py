"""e2e diagnostics on the GPU box: raw PCIe bandwidth and sgb_hmm_decode time vs chunk size."""
import ctypes, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from signals_b200 import _lib, hmm, synth
n_cells = int(os.environ.get("CELLS", "10000"))
d = synth.make_cohort(n_cells, seed=20260002)
N = d["n"].size
hn = torch.from_numpy(d["n"]).pin_memory(); hx = torch.from_numpy(d["x"]).pin_memory(); hc = torch.from_numpy(d["cn"]).pin_memory()
hout = torch.zeros(N, dtype=torch.uint8).pin_memory()
dev = torch.device("cuda:0")
dn = torch.empty_like(hn, device=dev)
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); dn.copy_(hn, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("H2D pinned %.1f GB/s" % (hn.numel() * 4 / dt / 1e9))
do = torch.zeros(N, dtype=torch.uint8, device=dev)
torch.cuda.synchronize(); t0 = time.perf_counter(); hout.copy_(do, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("D2H pinned %.1f GB/s" % (N / dt / 1e9))
lib = _lib.load()
P = hmm.make_params(model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12, selftransitionprob=0.95)
args = (ctypes.byref(P), N, hn.data_ptr(), hx.data_ptr(), hc.data_ptr(), 1, d["seg_off"].size - 1, d["seg_off"].ctypes.data, hout.data_ptr(), None)
for chunk in ("1000000", "2000000", "4000000", "8000000", "16000000", "32000000", "100000000"):
    os.environ["SGB_CHUNK_BINS"] = chunk
    for _ in range(2):
        hmm.check(lib.sgb_hmm_decode(*args))
    t0 = time.perf_counter()
    for _ in range(5):
        hmm.check(lib.sgb_hmm_decode(*args))
    dt = (time.perf_counter() - t0) / 5
    print("chunk %10s: %.2f ms  %.3g cell-bins/s" % (chunk, dt * 1e3, N / dt))
# host-side plan cost alone
so = d["seg_off"]
t0 = time.perf_counter()
for _ in range(5):
    pl = hmm.DevicePlan(N, so, n_states=12); pl.close()
print("plan create+destroy (230k sequences): %.2f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
