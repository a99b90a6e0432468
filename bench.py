#!/usr/bin/env python
"""bench.py -- headline benchmark of the HSCN/ASCN HMM hot path (emission + Viterbi-style decode).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload target|cfg2|cfg3] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic DLP-shaped cell-bins.
  value     cell-bins/s with inputs already resident in HBM (k_setup + k_decode, which reads the caller's columns
            in place; k_pack only on the packed-record path [+ all-gather of the uint8 state paths when N > 1,
            pipelined with the next step's decode]); CUDA events on the launch stream, max over ranks
  e2e       the same metric through the host-buffer C-ABI call sgb_hmm_decode: pinned host inputs are
            copied to the device, decoded, and the state paths copied back, every step
  roofline  k_decode against the FP64 pipe (the path is FP64-bound, not HBM- or tensor-bound):
            algorithmic flop-eq per launch / mean k_decode duration (CUDA events over the timed region) vs a
            DFMA peak measured in this process; the HBM view (10 B per cell-bin) is reported beside it
  cpu_baseline  the oracle's libm flavour (CPU restatement of the reference) on this box's host cores
Under torchrun (N > 1) every rank owns one GPU and its own shard of the cohort's cells (weak scaling: rank 0 holds
exactly the 1-GPU data, the other ranks further cells of the same clones).
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (n_cells, lens-key, model, likelihood, rho, seed, description)
    "target": (10000, "5700", "hscn", "binomial", 0.0, 20260002,
               "synthetic 10,000 cells x 5,700 500kb bins HSCN, maxCN=11, binomial (north_star target)"),
    "cfg2": (1000, "5700", "hscn", "binomial", 0.0, 20260001,
             "synthetic 1,000 cells x 5,700 500kb bins HSCN, maxCN=11, binomial (BASELINE configs[1])"),
    "cfg3": (10000, "5700", "ascn", "betabinomial", 0.02, 20260003,
             "synthetic 10,000 cells x 5,700 bins ASCN beta-binomial rho=0.02 (BASELINE configs[2])"),
    # the two large configs are generated as 2,000 distinct cells tiled up to size (same value distributions)
    "cfg4": (50000, "5700", "hscn", "binomial", 0.0, 20260004,
             "synthetic 50,000 cells x 5,700 bins HSCN binomial, cells split over the ranks (BASELINE configs[3])"),
    "cfg5": (20000, "30000", "hscn", "binomial", 0.0, 20260005,
             "synthetic 20,000 cells x 30,000 100kb bins HSCN binomial, cells split over the ranks (BASELINE configs[4])"),
}
BIG = {"cfg4", "cfg5"}
PARAMS = dict(n_states=12, loherror=0.02, eps=1e-12, Abias=0.0, selftransitionprob=0.95)
# kernels of one device-resident step, by what sgb_plan_last_staging reports
KERNELS = {
    4: ["k_setup", "k_split_class", "k_split_scan", "k_split_scatter", "k_split_groups", "k_split_gscan", "k_emit", "k_recur"],
    2: ["k_setup", "k_decode"],
    3: ["k_setup", "k_narrow_cn", "k_decode"],
    1: ["k_setup", "k_pack", "k_decode"],
    0: ["k_setup", "k_pack", "k_decode"],
}
F_BIN = {("hscn", "binomial"): 303, ("hscn", "betabinomial"): 404, ("ascn", "binomial"): 303, ("ascn", "betabinomial"): 404}
F_STATE = {("hscn", "binomial"): 99, ("hscn", "betabinomial"): 300, ("ascn", "binomial"): 198, ("ascn", "betabinomial"): 600}


def make_data(name, rank=0, strong=None):
    """strong=None: the workload's default (cfg4 / cfg5 fix the TOTAL number of cells, the others the cells per GPU)"""
    from signals_b200 import synth
    n_cells, lens, model, lik, rho, seed, desc = WORKLOADS[name]
    lens = synth.LENS_5700 if lens == "5700" else synth.LENS_30000
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if strong is None:
        strong = name in BIG
    if not strong:
        # weak scaling = one cohort (one set of clones) whose cells are spread over the ranks: rank 0 holds exactly
        # the 1-GPU data, the other ranks further cells of the same clones
        return synth.make_cohort(n_cells, lens=lens, seed=seed, likelihood=lik, shard=rank if rank else None)
    mine = n_cells // world                       # strong scaling: the TOTAL number of cells is fixed
    base = min(2000, mine)                        # generated as <= 2,000 distinct cells tiled up to size
    reps = max(1, mine // base)
    b = synth.make_cohort(base, lens=lens, seed=seed, likelihood=lik, shard=rank if rank else None)
    if reps == 1:
        return b
    d = dict(b)
    for k in ("n", "x", "cn", "truth_b"):
        d[k] = np.tile(b[k], reps)
    per = int(b["seg_off"][-1])
    d["seg_off"] = np.concatenate([b["seg_off"][:-1] + r * per for r in range(reps)] + [np.array([reps * per])]).astype(np.int64)
    d["n_cells"] = base * reps
    return d


def flop_eq(d, model, lik, S=12):
    """algorithmic flop-eq of one pass, SURVEY 8(d) convention: F = F_bin + L*F_state + 3*L_prev*L + L per cell-bin"""
    c = d["cn"].astype(np.int64)
    if model == "hscn":
        L = np.minimum(np.maximum(c, 1), S - 1) + 1
    else:
        L = np.where(c == 0, 0, np.minimum(c // 2, S - 1) + 1)
    Lp = np.empty_like(L)
    Lp[1:] = L[:-1]
    Lp[d["seg_off"][:-1]] = 0
    key = (model, lik)
    return float((F_BIN[key] + L * F_STATE[key] + 3 * Lp * L + L).sum())


class ClockSampler(threading.Thread):
    """polls NVML while the timed regions run (clocks + throttle reasons)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.ok = index, [], False, False
        self.active = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                if self.active:
                    self.samples.append((sm, rs, pw))
            except Exception:
                pass
            time.sleep(0.01)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": getattr(self, "max_sm", None), "reasons": ["unavailable"]}
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        seen = set()
        for _, rs, _ in self.samples:
            for bit, nm in names.items():
                if rs & bit:
                    seen.add(nm)
        sms = sorted(s for s, _, _ in self.samples)
        return {"sm_mhz": float(sms[len(sms) // 2]), "sm_max_mhz": float(self.max_sm), "reasons": sorted(seen),
                "samples": len(sms), "power_w_max": max(p for _, _, p in self.samples)}


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


ORACLE_FLAGS = "gcc -O2 -mfma -ffp-contract=off (oracle/Makefile; -O2 is R's default level)"


def cpu_reference_run(d, model, lik, rho, seconds_target, steps, warmup, one_thread_seconds=0.0):
    """times the oracle's libm flavour (the CPU restatement of the reference path) on all host threads, and -- when
    one_thread_seconds > 0 -- on ONE thread as well (the reference's default ncores = 1, R/callHSCN.R:736)"""
    from oracle import cpu_oracle as co
    co.build()
    cores = os.cpu_count() or 1
    P = co.params(model=model, likelihood=lik, S=PARAMS["n_states"], loherror=PARAMS["loherror"], eps=PARAMS["eps"],
                  rho=rho, Abias=PARAMS["Abias"], selftransitionprob=PARAMS["selftransitionprob"])
    cn32 = d["cn"].astype(np.int32)
    so = d["seg_off"]
    nseq_per_cell = len(d["lens"])
    # calibrate on a few cells, then size the sample for ~seconds_target of work per step
    cal_cells = min(d["n_cells"], max(2 * cores, 8))
    m = so[cal_cells * nseq_per_cell]
    t0 = time.perf_counter()
    co.decode_batch(d["n"][:m], d["x"][:m], cn32[:m], so[: cal_cells * nseq_per_cell + 1], P, co.MATH_LIBM, nthreads=cores)
    rate = m / (time.perf_counter() - t0)
    cells = int(min(d["n_cells"], max(cal_cells, rate * seconds_target / (m / cal_cells))))
    m = so[cells * nseq_per_cell]
    args = (d["n"][:m], d["x"][:m], cn32[:m], so[: cells * nseq_per_cell + 1], P, co.MATH_LIBM)
    for _ in range(warmup):
        co.decode_batch(*args, nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(steps):
        co.decode_batch(*args, nthreads=cores)
    dt = (time.perf_counter() - t0) / steps
    out = dict(value=float(m / dt), unit="cell-bins/s", cores=cores, kind="port", cpu=cpu_model(), flags=ORACLE_FLAGS,
               sample="first %d of %d cells (%d cell-bins) of the workload per step, oracle/oracle.c libm flavour, %d threads over "
                      "sequences (mirrors mclapply(mc.cores))" % (cells, d["n_cells"], m, cores))
    if one_thread_seconds > 0:
        c1 = int(min(cells, max(4, rate / cores * one_thread_seconds / (so[nseq_per_cell]))))
        m1 = so[c1 * nseq_per_cell]
        t0 = time.perf_counter()
        co.decode_batch(d["n"][:m1], d["x"][:m1], cn32[:m1], so[: c1 * nseq_per_cell + 1], P, co.MATH_LIBM, nthreads=1)
        out["one_thread"] = dict(value=float(m1 / (time.perf_counter() - t0)), cores=1,
                                 sample="first %d cells (%d cell-bins), 1 thread (the reference's default ncores = 1)" % (c1, m1))
    return out, dt, int(m)


def gpu_measure(name, d, steps, warmup, rank, world, dist, sampler, want_e2e=True):
    import torch
    from signals_b200 import _lib, hmm
    n_cells, _, model, lik, rho, seed, desc = WORKLOADS[name]
    dev = torch.device("cuda", torch.cuda.current_device())
    P = hmm.make_params(model=model, likelihood=lik, rho=rho, **PARAMS)
    N = d["n"].size
    # pinned host buffers (e2e inputs / outputs) and resident device copies (value)
    hn = torch.from_numpy(d["n"]).pin_memory(); hx = torch.from_numpy(d["x"]).pin_memory()
    # the state column as uint8 (comparable with round 1; what a caller that keeps narrow arrays passes) unless BENCH_CN32=1;
    # the int32 column R's INTSXP vectors give is measured beside it (value_int32_state, e2e.variants)
    cn_host = d["cn"].astype(np.int32) if os.environ.get("BENCH_CN32") == "1" else d["cn"]
    hc = torch.from_numpy(cn_host).pin_memory()
    hout = torch.zeros(N, dtype=torch.uint8).pin_memory()
    tn, tx, tc = hn.to(dev), hx.to(dev), hc.to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    so = d["seg_off"]
    gather_mode = "none"
    if world == 1:
        out = torch.zeros(N, dtype=torch.uint8, device=dev)
        plan = hmm.DevicePlan(N, so, n_states=PARAMS["n_states"])
        plan.enable_timing(True)
        plans = [plan]

        def step():
            plan.decode(P, tn, tx, tc, out, stream=stream)

        def drain():
            pass
    else:
        # Cells are independent: each rank decodes its own shard; the only exchange is the gather of the uint8
        # paths.  Preferred: every rank PUSHES its slice into all peers' gather buffers with copy-engine (DMA)
        # copies over NVLink (symmetric memory, no SMs involved) and one barrier closes the exchange.  The push of
        # step k runs on a side stream into the gather buffer of parity k&1, so it overlaps the decode of step
        # k+1 (which writes the other buffer); the timed region ends only after the last push and its barrier.
        # (Splitting ONE step's shard in halves to overlap inside a step was measured and lost more to the second
        # kernel tail than it hid.)  BENCH_GATHER=p2p_serial keeps push and decode back to back on one stream;
        # BENCH_GATHER=nccl, or symmetric memory being unavailable, uses NCCL all-gather.
        plan = hmm.DevicePlan(N, so, n_states=PARAMS["n_states"])
        plan.enable_timing(True)
        plans = [plan]
        hdl = None
        # BENCH_GATHER: "root" (default: what north_star asks for -- every rank pushes its paths to rank 0 only),
        # "p2p" (all-gather by pushes to every peer), "p2p_serial" (the same without overlap), "nccl" (NCCL all-gather)
        want = os.environ.get("BENCH_GATHER", "root")
        to_root = want == "root"
        # gather buffers in rotation: a rank may run up to nbuf steps ahead of the slowest one; the ranks meet at one
        # symmetric-memory barrier per rotation (before a buffer is reused), not per step, plus one at the very end
        nbuf = 1 if want == "p2p_serial" else max(2, int(os.environ.get("BENCH_GATHER_BUFS", "4")))
        if want in ("root", "p2p", "p2p_serial"):
            try:
                import torch.distributed._symmetric_memory as symm_mem
                if hasattr(symm_mem, "enable_symm_mem_for_group"):
                    try:
                        symm_mem.enable_symm_mem_for_group(dist.group.WORLD.group_name)
                    except Exception:
                        pass
                gathered_all = symm_mem.empty(nbuf * N * world, dtype=torch.uint8, device=dev)
                hdl = symm_mem.rendezvous(gathered_all, dist.group.WORLD)
                peers = [hdl.get_buffer(r, (nbuf * N * world,), torch.uint8) for r in range(world)]
                gather_mode = "p2p_dma_push_to_root_pipelined" if to_root else ("p2p_dma_push" if nbuf == 1 else "p2p_dma_push_pipelined")
            except Exception as ex:          # symmetric memory unavailable: NCCL path below
                hdl = None
                sys.stderr.write("symmetric memory unavailable (%s); using NCCL all_gather\n" % ex)
        if hdl is None:
            nbuf = 1
            gathered_all = torch.empty(N * world, dtype=torch.uint8, device=dev)
            gather_mode = "nccl_all_gather"
        W = N * world
        outs = [gathered_all[b * W + rank * N: b * W + (rank + 1) * N] for b in range(nbuf)] if hdl is not None \
            else [torch.zeros(N, dtype=torch.uint8, device=dev)]
        side = torch.cuda.Stream(device=dev) if nbuf >= 2 else None
        trace = [] if os.environ.get("BENCH_TRACE") == "1" else None       # per-step (decode start, decode end, push end) events
        ev_dec = [torch.cuda.Event() for _ in range(nbuf)]
        ev_push = [None] * nbuf
        step_no = [0]

        # the copies to the peers are spread over a few streams so that several copy engines run at once
        n_cp = max(1, min(int(os.environ.get("BENCH_PUSH_STREAMS", "3")), world - 1))
        cp_streams = [torch.cuda.Stream(device=dev) for _ in range(n_cp - 1)]

        def push(b, barrier=True):
            here = torch.cuda.current_stream()
            if cp_streams:
                start = torch.cuda.Event()
                start.record(here)
            for k in range(1, world):
                r = (rank + k) % world
                if to_root and r != 0:
                    continue
                st = ([here] + cp_streams)[(k - 1) % n_cp]
                if st is not here:
                    st.wait_event(start)
                with torch.cuda.stream(st):
                    peers[r][b * W + rank * N: b * W + (rank + 1) * N].copy_(outs[b], non_blocking=True)
            for st in cp_streams:
                here.wait_stream(st)
            if barrier:
                hdl.barrier()

        def step():
            b = step_no[0] % nbuf
            step_no[0] += 1
            if hdl is None:
                plan.decode(P, tn, tx, tc, outs[0], stream=stream)
                dist.all_gather_into_tensor(gathered_all, outs[0])
            elif side is None:
                plan.decode(P, tn, tx, tc, outs[0], stream=stream)
                push(0)
            else:
                if ev_push[b] is not None:
                    torch.cuda.current_stream().wait_event(ev_push[b])      # the push that last read outs[b] is done
                if trace is not None:
                    t0 = torch.cuda.Event(enable_timing=True); t0.record()
                plan.decode(P, tn, tx, tc, outs[b], stream=stream)
                ev_dec[b] = torch.cuda.Event(enable_timing=trace is not None)
                ev_dec[b].record()
                with torch.cuda.stream(side):
                    side.wait_event(ev_dec[b])
                    push(b, barrier=(b == nbuf - 1))
                    ev_push[b] = torch.cuda.Event(enable_timing=trace is not None)
                    ev_push[b].record()
                if trace is not None:
                    trace.append((t0, ev_dec[b], ev_push[b]))

        def drain():
            # joins the side stream: every gather issued so far has landed (one closing barrier over all ranks)
            if side is not None:
                with torch.cuda.stream(side):
                    hdl.barrier()
                torch.cuda.current_stream().wait_stream(side)

    for _ in range(max(warmup, 3)):
        step()
    drain()
    for pl in plans:
        pl.check(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    for pl in plans:
        pl.enable_timing(True)          # restart the rings so they only hold timed steps
    small = N * (8 + cn_host.dtype.itemsize) < 200e6               # inputs fit in the 126 MB L2: flush it between (individually timed) steps
    sampler.active = True
    if not small:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        drain()
        e1.record()
        torch.cuda.synchronize()
        elapsed = e0.elapsed_time(e1)
    else:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        elapsed = 0.0
        for _ in range(steps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); drain(); e1.record()
            torch.cuda.synchronize()
            elapsed += e0.elapsed_time(e1)
        del flush
    sampler.active = False
    if world > 1 and trace:
        # diagnostic only: where a step's time goes on this rank (ms relative to the first traced decode start)
        tr_ = trace[-steps:]
        base = tr_[0][0]
        rows = ["%2d start %.3f dec_end %.3f push_end %.3f" % (i, base.elapsed_time(a_), base.elapsed_time(b_), base.elapsed_time(c_))
                for i, (a_, b_, c_) in enumerate(tr_)]
        sys.stderr.write("rank %d trace\n  " % rank + "\n  ".join(rows) + "\n")
    if world > 1:
        dist.barrier()
    ms = torch.tensor([elapsed], dtype=torch.float64, device=dev)
    by_rank = None
    if world > 1:
        allms = [torch.zeros_like(ms) for _ in range(world)]
        dist.all_gather(allms, ms)
        by_rank = [float(v.item()) / steps for v in allms]
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    for pl in plans:
        pl.check(stream)
    if world > 1:
        last = (step_no[0] - 1) % nbuf          # the buffer the last step decoded into and gathered
        out = outs[last]
        gathered = gathered_all[last * W: (last + 1) * W]
    t3 = plans[0].read_timing3(64)
    prep_ms, emit_ms, decode_ms = float(t3[0].mean()), float(t3[1].mean()), float(t3[2].mean())
    stg = plans[0].last_staging()
    ngroups, e_bytes = plans[0].split_info(stream)
    res = dict(ms_per_step=total_ms / steps, value=N * world * steps / (total_ms * 1e-3),
               decode_ms=decode_ms, emit_ms=emit_ms, pack_ms=prep_ms, N=N, gather=gather_mode, staging=stg,
               kernels=KERNELS[stg], launches_per_step=len(KERNELS[stg]), e_bytes=e_bytes, ngroups=ngroups,
               cn_dtype=str(cn_host.dtype), ms_per_step_by_rank=by_rank,
               scratch_bytes=sum(pl.scratch_bytes() for pl in plans), checksum=int(out.to(torch.int64).sum().item()))
    if world > 1:
        # the gathering rank(s) must hold every rank's paths: compare checksums of the gathered slices with their owners'
        sums = torch.stack([gathered[r * N: (r + 1) * N].to(torch.int64).sum() for r in range(world)])
        ref = torch.zeros(world, dtype=torch.int64, device=dev)
        ref[rank] = res["checksum"]
        dist.all_reduce(ref)
        if gather_mode != "p2p_dma_push_to_root_pipelined" or rank == 0:
            assert torch.equal(sums, ref), "gathered paths differ from the owners' results"

    if want_e2e:
        lib = _lib.load()
        args = (ctypes.byref(P), N, hn.data_ptr(), hx.data_ptr(), hc.data_ptr(), cn_host.dtype.itemsize, d["seg_off"].size - 1,
                d["seg_off"].ctypes.data, hout.data_ptr(), None)
        for _ in range(max(2, min(warmup, 3))):
            hmm.check(lib.sgb_hmm_decode(*args))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e2e_steps = max(3, min(steps, 10))
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.active = True
        t0 = time.perf_counter()
        f0.record()
        for _ in range(e2e_steps):
            hmm.check(lib.sgb_hmm_decode(*args))       # synchronous: returns when the paths are back on the host
        f1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        sampler.active = False
        ems = torch.tensor([max(f0.elapsed_time(f1), wall * 1e3)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        e2e_ms = float(ems.item()) / e2e_steps
        nseq = d["seg_off"].size - 1
        res["e2e"] = dict(value=N * world / (e2e_ms * 1e-3), unit="cell-bins/s", ms_per_step=e2e_ms,
                          h2d_bytes_per_step=int(N * (8 + cn_host.dtype.itemsize) + nseq * 20), d2h_bytes_per_step=int(N), steps=e2e_steps,
                          api="sgb_hmm_decode (host buffers: page-locked int32 counts + %s state in, uint8 states out; chunked "
                              "copy / compute overlap, copied directly)" % cn_host.dtype,
                          bytes_note="h2d/d2h = bytes of the caller's tensors consumed / produced per step")
        assert int(hout.to(torch.int64).sum().item()) == res["checksum"], "host-API and device-API results differ"
        if want_e2e == "all" and world == 1:
            # the same call as an R session makes it: PAGEABLE vectors, INTSXP (int32) or REALSXP (double) columns, states
            # returned as doubles (state_min is numeric in the reference, R/callHSCN.R:169) -- sgb_hmm_decode_typed
            variants = {}
            so_ = d["seg_off"]
            for label, dt_in, dt_out in (("pageable_int32_in_uint8_out", np.int32, np.uint8),
                                         ("pageable_int32_in_double_out", np.int32, np.float64),
                                         ("pageable_double_in_double_out", np.float64, np.float64)):
                pn, px, pc = d["n"].astype(dt_in), d["x"].astype(dt_in), d["cn"].astype(dt_in)
                po = np.zeros(N, dtype=dt_out)
                code = {np.int32: _lib.T_I32, np.float64: _lib.T_F64, np.uint8: _lib.T_U8}
                targs = (ctypes.byref(P), N, pn.ctypes.data, code[dt_in], px.ctypes.data, code[dt_in], pc.ctypes.data, code[dt_in],
                         so_.size - 1, so_.ctypes.data, po.ctypes.data, code[dt_out], 1)
                for _ in range(2):
                    hmm.check(lib.sgb_hmm_decode_typed(*targs))
                t0 = time.perf_counter()
                for _ in range(3):
                    hmm.check(lib.sgb_hmm_decode_typed(*targs))
                ms = (time.perf_counter() - t0) / 3 * 1e3
                assert int(po.astype(np.int64).sum()) == res["checksum"], label
                variants[label] = dict(value=N / (ms * 1e-3), ms_per_step=ms,
                                       h2d_bytes_per_step=int(N * 3 * np.dtype(dt_in).itemsize), d2h_bytes_per_step=int(N * np.dtype(dt_out).itemsize))
                del pn, px, pc, po
            hc32 = torch.from_numpy(d["cn"].astype(np.int32)).pin_memory()
            a32 = (ctypes.byref(P), N, hn.data_ptr(), hx.data_ptr(), hc32.data_ptr(), 4, d["seg_off"].size - 1,
                   d["seg_off"].ctypes.data, hout.data_ptr(), None)
            for label, env in (("pinned_int32_state_direct", None), ("pinned_int32_state_staged_narrowed", "1")):
                if env:
                    os.environ["SGB_HOST_NARROW"] = env
                for _ in range(2):
                    hmm.check(lib.sgb_hmm_decode(*a32))
                t0 = time.perf_counter()
                for _ in range(3):
                    hmm.check(lib.sgb_hmm_decode(*a32))
                ms = (time.perf_counter() - t0) / 3 * 1e3
                os.environ.pop("SGB_HOST_NARROW", None)
                variants[label] = dict(value=N / (ms * 1e-3), ms_per_step=ms, h2d_bytes_per_step=int(N * 12), d2h_bytes_per_step=int(N))
            del hc32
            # through the .Call shim itself (rpkg/src/r_shim.c compiled against the mock of R's C API, tests/r_mock): the call
            # an R session makes, on pageable INTSXP / REALSXP vectors, numeric state_min back
            shim_so = os.path.join(ROOT, "tests", "r_mock", "libshimtest.so")
            if os.path.exists(shim_so):
                try:
                    sh = ctypes.CDLL(shim_so)
                    sh.drv_decode_batch_timed.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long,
                                                          ctypes.c_void_p, ctypes.c_long, ctypes.c_char_p, ctypes.c_int] + \
                        [ctypes.c_double] * 5 + [ctypes.c_int] * 3 + [ctypes.POINTER(ctypes.c_double)] * 2 + [ctypes.c_char_p]
                    cn32 = d["cn"].astype(np.int32)
                    segd = so_.astype(np.float64)
                    for label, asd in (("r_shim_call_pageable_INTSXP", 0), ("r_shim_call_pageable_REALSXP", 1)):
                        ms_ = ctypes.c_double(0.0); cs_ = ctypes.c_double(0.0); eb = ctypes.create_string_buffer(1024)
                        rc = sh.drv_decode_batch_timed(1 if model == "ascn" else 0, d["n"].ctypes.data, d["x"].ctypes.data, cn32.ctypes.data,
                                                       N, segd.ctypes.data, segd.size - 1, lik.encode(), PARAMS["n_states"] - 1,
                                                       PARAMS["loherror"], PARAMS["eps"], rho, PARAMS["Abias"], PARAMS["selftransitionprob"],
                                                       asd, 1, 3, ctypes.byref(ms_), ctypes.byref(cs_), eb)
                        if rc == 0 and int(cs_.value) == res["checksum"]:
                            variants[label] = dict(value=N / (ms_.value * 1e-3), ms_per_step=ms_.value,
                                                   h2d_bytes_per_step=int(N * 3 * (8 if asd else 4)), d2h_bytes_per_step=int(N * 8))
                        else:
                            variants[label] = dict(error=eb.value.decode() or "checksum mismatch")
                except Exception as ex:
                    variants["r_shim_call"] = dict(error=str(ex))
            # the two HMM passes of callHaplotypeSpecificCN (R/callHSCN.R:883-935: pass 1 at the default loherror, pass 2 at the
            # inferred one) on ONE upload of the table: columns copied to the device once, hmm.DevicePlan.decode twice, both state
            # vectors copied back -- what a caller who keeps the table on the device between the passes pays for the pair
            try:
                P2 = hmm.make_params(model=model, likelihood=lik, rho=rho, **dict(PARAMS, loherror=0.031))
                pl2 = hmm.DevicePlan(N, d["seg_off"], n_states=PARAMS["n_states"])
                dn_, dx_, dc_ = (torch.empty(N, dtype=t_.dtype, device=dev) for t_ in (hn, hx, hc))
                o1_, o2_ = torch.empty(N, dtype=torch.uint8, device=dev), torch.empty(N, dtype=torch.uint8, device=dev)
                h1_, h2_ = torch.empty(N, dtype=torch.uint8).pin_memory(), torch.empty(N, dtype=torch.uint8).pin_memory()
                st_ = torch.cuda.current_stream().cuda_stream
                def two_pass():
                    dn_.copy_(hn, non_blocking=True); dx_.copy_(hx, non_blocking=True); dc_.copy_(hc, non_blocking=True)
                    pl2.decode(P, dn_, dx_, dc_, o1_, stream=st_)
                    h1_.copy_(o1_, non_blocking=True)
                    pl2.decode(P2, dn_, dx_, dc_, o2_, stream=st_)
                    h2_.copy_(o2_, non_blocking=True)
                    torch.cuda.synchronize()
                for _ in range(2):
                    two_pass()
                pl2.check(st_)
                t0 = time.perf_counter()
                for _ in range(3):
                    two_pass()
                ms = (time.perf_counter() - t0) / 3 * 1e3
                assert int(h1_.to(torch.int64).sum().item()) == res["checksum"], "two-pass variant: pass 1 differs"
                variants["two_passes_one_upload_pinned"] = dict(value=2 * N / (ms * 1e-3), ms_per_step=ms, ms_per_pass=ms / 2,
                                                                h2d_bytes_per_step=int(N * (8 + cn_host.dtype.itemsize)), d2h_bytes_per_step=int(2 * N),
                                                                note="value counts the cell-bins of both passes")
                pl2.close()
                del dn_, dx_, dc_, o1_, o2_, h1_, h2_
            except Exception as ex:
                variants["two_passes_one_upload_pinned"] = dict(error=str(ex))
            # the same pair of passes for an R-style caller: PAGEABLE int32 columns through sgb_cols_upload (one conversion +
            # upload), two sgb_cols_decode calls with numeric (double) results -- against 2 x pageable_int32_in_double_out
            try:
                pn, px, pc = d["n"].astype(np.int32), d["x"].astype(np.int32), d["cn"].astype(np.int32)
                po1, po2 = np.zeros(N, dtype=np.float64), np.zeros(N, dtype=np.float64)
                P2 = hmm.make_params(model=model, likelihood=lik, rho=rho, **dict(PARAMS, loherror=0.031))
                def two_pass_cols():
                    h = ctypes.c_void_p()
                    hmm.check(lib.sgb_cols_upload(N, pn.ctypes.data, _lib.T_I32, px.ctypes.data, _lib.T_I32, pc.ctypes.data, _lib.T_I32,
                                                  so_.size - 1, so_.ctypes.data, ctypes.byref(h)))
                    hmm.check(lib.sgb_cols_decode(h, ctypes.byref(P), po1.ctypes.data, _lib.T_F64))
                    hmm.check(lib.sgb_cols_decode(h, ctypes.byref(P2), po2.ctypes.data, _lib.T_F64))
                    lib.sgb_cols_free(h)
                two_pass_cols()
                t0 = time.perf_counter()
                for _ in range(3):
                    two_pass_cols()
                ms = (time.perf_counter() - t0) / 3 * 1e3
                assert int(po1.sum()) == res["checksum"], "sgb_cols two-pass variant: pass 1 differs"
                variants["two_passes_one_upload_pageable_int32_double_out"] = dict(
                    value=2 * N / (ms * 1e-3), ms_per_step=ms, ms_per_pass=ms / 2, h2d_bytes_per_step=int(N * 9),
                    d2h_bytes_per_step=int(2 * N), note="sgb_cols_upload + 2 x sgb_cols_decode + sgb_cols_free; value counts both passes")
                del pn, px, pc, po1, po2
            except Exception as ex:
                variants["two_passes_one_upload_pageable_int32_double_out"] = dict(error=str(ex))
            res["e2e"]["variants"] = variants
    for pl in plans:
        pl.close()
    del tn, tx, tc, out, hn, hx, hc, hout
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="target", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"],
                    help="weak: the workload's cells PER GPU (default for target / cfg2 / cfg3); strong: its cells in TOTAL, "
                         "split over the ranks (default for cfg4 / cfg5, and what BASELINE configs[2] and [3] describe)")
    ap.add_argument("--no-extras", action="store_true", help="skip the other workloads and the CPU baseline")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n_cells, lens, model, lik, rho, seed, desc = WORKLOADS[a.workload]
    strong = (a.workload in BIG) if a.scaling is None else (a.scaling == "strong")
    if strong:
        n_cells = n_cells // max(world, 1)
    bins_per_cell = 5700 if lens == "5700" else 30000
    config = {"workload": desc, "cells_per_gpu": n_cells, "bins_per_cell": bins_per_cell, "sequences_per_cell": 23,
              "maxCN": 11, "n_states": 12, "likelihood": lik, "model": model.upper(), "selftransitionprob": 0.95,
              "loherror": 0.02, "eps": 1e-12, "rho": rho, "sharding": "cells, %d rank(s), 1 GPU each" % max(world, 1),
              "l2": "inputs (%d MB per GPU) %s" % (n_cells * bins_per_cell * 12 // 1000000,
                                                   "larger than L2; no explicit flush" if n_cells * bins_per_cell * 12 > 200e6
                                                   else "fit in L2: 256 MB flush write between individually timed steps"),
              "columns": "totalcounts, alleleB int32; state uint8 (int32-state and pageable / double variants are reported beside the headline)",
              # best-effort NUMA binding of the ranks' pinned buffers is applied when world > 1 (both arms carry the key)
              "host_numa_binding": bool(world > 1 and os.environ.get("BENCH_NUMA", "1") == "1")}

    if a.impl == "reference":
        # The reference's own CPU implementation cannot run (no R / Rcpp / VGAM in this image; src/code.cpp needs
        # <Rcpp.h>), so this arm times the oracle: the C restatement of the same path, all host threads.
        if rank != 0:
            return 0
        d = make_data(a.workload, 0, strong)
        per_step = max(0.5, min(8.0, 100.0 / max(a.steps + a.warmup, 1)))      # whole run stays within a few minutes
        cb, dt, m = cpu_reference_run(d, model, lik, rho, per_step, max(a.steps, 1), max(a.warmup, 0), one_thread_seconds=5.0)
        line = {"impl": "reference", "metric": "cell_bins_per_sec", "value": cb["value"], "unit": "cell-bins/s",
                "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3,
                "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "cell-bins/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "reference = oracle port (R toolchain absent); each step decodes a BOUNDED SAMPLE of the workload "
                        "(see cpu_baseline.sample) and the rate is per cell-bin; cells are independent"}
        print(json.dumps(line))
        return 0

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        # several ranks stream pinned host buffers at once: keep this rank's threads (and so its first-touch
        # pinned pages) on the CPUs NVML reports as local to its GPU.  Best effort; N=1 keeps every core.
        if config["host_numa_binding"]:
            try:
                import pynvml
                pynvml.nvmlInit()
                pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
            except Exception as ex:
                sys.stderr.write("rank %d: no NUMA binding (%s)\n" % (rank, ex))
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from signals_b200 import _lib, hmm
    lib = _lib.load()
    hmm.check(lib.sgb_init(local_rank))
    sampler = ClockSampler(local_rank)
    sampler.start()

    d = make_data(a.workload, rank, strong)
    main_res = gpu_measure(a.workload, d, a.steps, a.warmup, rank, world, dist, sampler, want_e2e="all")
    fe = flop_eq(d, model, lik)
    clocks = sampler.summary()
    peak = ctypes.c_double(0.0)
    hmm.check(lib.sgb_fp64_peak(20, ctypes.byref(peak)))
    fp64_peak = max(peak.value, 37.0)            # max(measured DFMA rate, 37 TFLOP/s nominal)
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm_peak, hbm_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_kernel_traffic.json")))
        traffic = tj.get(a.workload, {}).get("split" if main_res["staging"] == 4 else "fused")
    except Exception:
        traffic = None
    N = main_res["N"]
    kern_ms = main_res["pack_ms"] + main_res["emit_ms"] + main_res["decode_ms"]
    achieved_tf = fe / (kern_ms * 1e-3) / 1e12
    split = main_res["staging"] == 4
    roofline = {"bound": "fp64", "kernel": "k_emit + k_recur (split schedule)" if split else "k_decode (fused schedule)",
                "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved_tf / fp64_peak,
                "traffic": traffic,
                "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture "
                                "(profiles/r02_kernel_traffic.json); algorithmic bytes = 13 B x cell-bins with int32 columns (10 with uint8 state)",
                "peak_source": "max(DFMA micro-kernel measured in this process = %.2f, 37.0 nominal)" % peak.value,
                "flop_eq_per_cell_bin": fe / N, "flop_eq_convention": "SURVEY 8(d): FMA=2, exp=40, log=log1p=50, lgamma=100, live states only",
                "kernel_ms": kern_ms, "prep_ms": main_res["pack_ms"], "emit_ms": main_res["emit_ms"], "decode_ms": main_res["decode_ms"],
                "hbm": {"achieved": 13.0 * N / (kern_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": 13.0 * N / (kern_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": hbm_src,
                        "algorithmic_bytes_per_cell_bin": 13}}
    if split and main_res["decode_ms"] > 0:
        # the two kernels of the split schedule against their own bounds: k_emit is FP64 / issue bound (all of the
        # emission flop-eq), k_recur is HBM-bound (reads the emission tiles, writes 1 B per cell-bin)
        rb = main_res["e_bytes"] + N
        roofline["per_kernel"] = {
            "k_emit": {"bound": "fp64", "ms": main_res["emit_ms"], "frac": fe / (main_res["emit_ms"] * 1e-3) / 1e12 / fp64_peak,
                       "note": "flop-eq of the whole step credited to k_emit (the recursion is < 8 % of it)"},
            "k_recur": {"bound": "hbm", "ms": main_res["decode_ms"], "bytes": rb,
                        "achieved_gbs": rb / (main_res["decode_ms"] * 1e-3) / 1e9,
                        "frac": rb / (main_res["decode_ms"] * 1e-3) / 1e9 / hbm_peak,
                        "bytes_per_cell_bin": rb / N}}
    line = {"metric": "cell_bins_per_sec", "value": main_res["value"], "unit": "cell-bins/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": main_res["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config, "clocks": clocks, "e2e": main_res.get("e2e"), "gpu_launches": main_res["launches_per_step"] * a.steps,
            "gather": main_res["gather"], "gpu_kernels_per_step": main_res["kernels"], "ms_per_step_by_rank": main_res["ms_per_step_by_rank"],
            "schedule": "split" if split else "fused", "roofline": roofline,
            "scratch_bytes": main_res["scratch_bytes"], "torch_peak_device_bytes": int(torch.cuda.max_memory_allocated())}

    if rank == 0 and world == 1 and not a.no_extras:
        others = {}
        for nm in sorted(WORKLOADS):
            if nm == a.workload or nm in BIG:
                continue
            dd = make_data(nm)
            r = gpu_measure(nm, dd, max(5, a.steps // 2), 3, 0, 1, None, sampler, want_e2e=True)
            f2 = flop_eq(dd, WORKLOADS[nm][2], WORKLOADS[nm][3])
            km = r["pack_ms"] + r["emit_ms"] + r["decode_ms"]
            others[nm] = {"workload": WORKLOADS[nm][6], "value": r["value"], "ms_per_step": r["ms_per_step"],
                          "kernel_ms": km, "schedule": "split" if r["staging"] == 4 else "fused", "e2e_value": r["e2e"]["value"],
                          "fp64_frac": f2 / (km * 1e-3) / 1e12 / fp64_peak, "flop_eq_per_cell_bin": f2 / r["N"]}
            del dd
        # the headline workload with the state column as the int32 vector R holds (k_narrow_cn joins the step)
        os.environ["BENCH_CN32"] = "1"
        r = gpu_measure(a.workload, d, max(5, a.steps // 2), 3, 0, 1, None, sampler, want_e2e=False)
        os.environ.pop("BENCH_CN32", None)
        others["%s_int32_state" % a.workload] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "kernels": r["kernels"],
                                                 "kernel_ms": r["pack_ms"] + r["emit_ms"] + r["decode_ms"]}
        line["other_workloads"] = others
        cb, _, _ = cpu_reference_run(d, model, lik, rho, 12.0, 1, 0, one_thread_seconds=6.0)
        line["cpu_baseline"] = cb
    if world > 1:
        # the in-library multi-device entry point (one host thread per device inside ONE process, what the R drop-in
        # uses) on the GPUs of this job: rank 0 decodes a small cohort with n_devices = world and with one device and
        # the two results must be equal.  Outside the timed region; the other ranks wait at the barrier.
        dist.barrier()
        if rank == 0:
            dd = make_data("cfg2", 0, False)
            kw = dict(model="hscn", likelihood="binomial", rho=0.0, **PARAMS)
            nd = min(world, lib.sgb_device_count())
            one = hmm.decode_arrays(dd["n"], dd["x"], dd["cn"], dd["seg_off"], **kw)
            t0 = time.perf_counter()
            many = hmm.decode_arrays(dd["n"], dd["x"], dd["cn"], dd["seg_off"], n_devices=nd, **kw)
            dt = time.perf_counter() - t0
            assert np.array_equal(one, many), "sgb_hmm_decode_multi differs from sgb_hmm_decode"
            line["multi_device_check"] = {"api": "sgb_hmm_decode_multi", "n_devices": nd, "cell_bins": int(one.size),
                                          "equal_to_single_device": True, "wall_ms_first_call": dt * 1e3}
            torch.cuda.set_device(local_rank)
        dist.barrier()
    sampler.stop_flag = True
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
