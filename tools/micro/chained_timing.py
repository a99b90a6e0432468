"""Device-resident chained decode (k_decode<CHAINED> + k_backtrace) on the target workload, against the reference rule."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from signals_b200 import hmm
name = sys.argv[1] if len(sys.argv) > 1 else "target"
n_cells, lens, model, lik, rho, seed, desc = bench.WORKLOADS[name]
d = bench.make_data(name)
dev = torch.device("cuda:0")
tn, tx, tc = (torch.from_numpy(d[k]).to(dev) for k in ("n", "x", "cn"))
N = d["n"].size
out = torch.zeros(N, dtype=torch.uint8, device=dev)
for dec in ("reference", "chained"):
    P = hmm.make_params(model=model, likelihood=lik, rho=rho, decode=dec, **bench.PARAMS)
    plan = hmm.DevicePlan(N, d["seg_off"], n_states=12, decode=dec)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        plan.decode(P, tn, tx, tc, out, stream=st)
    plan.check(st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        plan.decode(P, tn, tx, tc, out, stream=st)
    e1.record(); torch.cuda.synchronize()
    print(name, dec, "ms_per_step %.3f" % (e0.elapsed_time(e1) / 5), "scratch MB %.0f" % (plan.scratch_bytes() / 1e6), "staging", plan.last_staging())
    plan.close()
