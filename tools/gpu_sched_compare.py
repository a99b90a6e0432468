"""Times the device-resident decode under both schedules (SGB_SCHED=split|fused) on the bench workloads (GPU box)."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from signals_b200 import hmm

def run(name, sched, steps=10):
    os.environ["SGB_SCHED"] = sched
    n_cells, lens, model, lik, rho, seed, desc = bench.WORKLOADS[name]
    d = bench.make_data(name)
    dev = torch.device("cuda:0")
    tn, tx, tc = (torch.from_numpy(d[k]).to(dev) for k in ("n", "x", "cn"))
    N = d["n"].size
    out = torch.zeros(N, dtype=torch.uint8, device=dev)
    P = hmm.make_params(model=model, likelihood=lik, rho=rho, **bench.PARAMS)
    plan = hmm.DevicePlan(N, d["seg_off"], n_states=12)
    timing = not os.environ.get("SGB_NO_TIMING")      # (the per-kernel events sit between the launches)
    plan.enable_timing(timing)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        plan.decode(P, tn, tx, tc, out, stream=st)
    plan.check(st)
    plan.enable_timing(timing)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot = 0.0
    for _ in range(steps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); plan.decode(P, tn, tx, tc, out, stream=st); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    a, b, c = plan.read_timing3(64) if timing else (np.zeros(1), np.zeros(1), np.zeros(1))
    res = dict(workload=name, sched=sched, ms_per_step=tot / steps, prep_ms=float(a.mean()), emit_ms=float(b.mean()),
               decode_ms=float(c.mean()), staging=plan.last_staging(), scratch_MB=plan.scratch_bytes() / 1e6,
               checksum=int(out.to(torch.int64).sum().item()))
    plan.close()
    return res

if __name__ == "__main__":
    names = sys.argv[1:] or ["cfg2", "target", "cfg3"]
    for nm in names:
        rs = [run(nm, s) for s in ("fused", "split")]
        assert rs[0]["checksum"] == rs[1]["checksum"], rs
        for r in rs:
            print(json.dumps(r))
