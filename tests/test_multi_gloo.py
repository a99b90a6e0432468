"""world_size-2 gloo run of the cell-axis sharding + path gather used by bench.py (CPU; the oracle stands in
for the GPU decode, which is exactly what the N>1 host logic wraps)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import cpu_oracle as co
    from signals_b200 import dist as sd, synth
    d = synth.make_cohort(9, lens=synth.LENS_DLPBINS, seed=77, ragged_drop=0.05)
    so = d["seg_off"]
    cn = d["cn"].astype(np.int32)
    P = co.params(S=12, selftransitionprob=0.95)
    cuts = [sd.shard_sequences(so, r, world) for r in range(world)]
    a, b = cuts[rank]
    r0, r1 = so[a], so[b]
    local = co.decode_batch(d["n"][r0:r1], d["x"][r0:r1], cn[r0:r1], so[a:b + 1] - r0, P, co.MATH_DET)
    sizes = [int(so[e] - so[s]) for s, e in cuts]
    full = sd.gather_paths(torch.from_numpy(local), sizes)
    # between-pass sums of the sharded cohort: every rank reduces its own rows (the oracle's table restatement stands in for
    # sgb_pass_stats2), one all-reduce gives every rank the cohort's sums
    tab = co.postprocess(cn[r0:r1], local)
    A, B = tab["A"], tab["B"]
    ah = (B == 0) & (A > 0); ms = (A == 1) & (B == 1)
    x_, n_ = d["x"][r0:r1].astype(np.int64), d["n"][r0:r1].astype(np.int64)
    out9 = np.array([x_[ah].sum(), n_[ah].sum(), ah.sum(), ms.sum(), x_[ms].sum(), n_[ms].sum(), (x_[ms] ** 2).sum(),
                     (x_[ms] * n_[ms]).sum(), (n_[ms] ** 2).sum()], np.float64)
    hist = np.zeros((12, 12), np.int64)
    np.add.at(hist, (A[B > 0], B[B > 0]), 1)
    o_all, h_all = sd.allreduce_pass_stats(out9, hist)
    if rank == 0:
        ref = co.decode_batch(d["n"], d["x"], cn, so, P, co.MATH_DET)
        ok = bool(np.array_equal(full.numpy(), ref))
        tabr = co.postprocess(cn, ref)
        Ar, Br = tabr["A"], tabr["B"]
        ahr = (Br == 0) & (Ar > 0); msr = (Ar == 1) & (Br == 1)
        xr, nr = d["x"].astype(np.int64), d["n"].astype(np.int64)
        ref9 = [xr[ahr].sum(), nr[ahr].sum(), ahr.sum(), msr.sum(), xr[msr].sum(), nr[msr].sum(), (xr[msr] ** 2).sum(),
                (xr[msr] * nr[msr]).sum(), (nr[msr] ** 2).sum()]
        hr = np.zeros((12, 12), np.int64)
        np.add.at(hr, (Ar[Br > 0], Br[Br > 0]), 1)
        ok = ok and o_all.tolist() == [float(v) for v in ref9] and bool(np.array_equal(h_all.numpy(), hr))
        q.put(ok)
    dist.barrier()
    dist.destroy_process_group()


def test_shard_and_gather_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
