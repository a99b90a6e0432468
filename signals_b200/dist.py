"""Cell-axis sharding helpers for one-process-per-GPU runs (torch.distributed).

Cells -- indeed (cell, chromosome) sequences -- are independent (R/callHSCN.R:244-280 is a plain map over
cells, parallelised in the reference only by parallel::mclapply fork fan-out), so the path shards with no
data-path collective; the only communication is gathering the uint8 state paths.
"""
from __future__ import annotations

import numpy as np


def shard_sequences(seg_off, rank, world):
    """contiguous block of sequences for `rank`, balanced by bin count.  Returns (s_begin, s_end)."""
    seg_off = np.asarray(seg_off, dtype=np.int64)
    nseg = seg_off.size - 1
    if world <= 1 or nseg == 0:
        return 0, nseg
    total = seg_off[-1] - seg_off[0]
    cuts = [0]
    for d in range(1, world):
        want = seg_off[0] + total * d // world
        cuts.append(int(np.searchsorted(seg_off[:-1], want, side="left")))
    cuts.append(nseg)
    return cuts[rank], cuts[rank + 1]


def gather_paths(local_paths, sizes, group=None, dst=0):
    """Gather per-rank uint8 path tensors of (possibly different) lengths `sizes` onto rank `dst`.
    Uses all_gather on padded tensors (NCCL over NVLink on GPUs, gloo on CPU).  Returns the concatenated
    tensor on every rank (all_gather) -- callers that only need it on `dst` ignore it elsewhere."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return local_paths
    maxlen = int(max(sizes))
    buf = local_paths
    if local_paths.numel() < maxlen:
        buf = torch.zeros(maxlen, dtype=local_paths.dtype, device=local_paths.device)
        buf[: local_paths.numel()] = local_paths
    out = torch.empty(world * maxlen, dtype=local_paths.dtype, device=local_paths.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    parts = [out[r * maxlen: r * maxlen + int(sizes[r])] for r in range(world)]
    return torch.cat(parts)


def allreduce_pass_stats(out9, hist, group=None):
    """Between-pass sums of a cohort sharded over ranks: every rank holds the sums of its own cells
    (`table.pass_stats2_arrays`: nine doubles that are exact integers below 2^53, and the A x B histogram); one
    all-reduce makes them the cohort's -- infloherror (R/callHSCN.R:895-899), tarones_Z (:609-616) and fitBB's modal
    state (:619) then follow on every rank without gathering a single row.  float64 / int64 sums of exact integers:
    the result does not depend on the reduction order.  Tensors may live on the GPU (NCCL) or the CPU (gloo)."""
    import torch
    import torch.distributed as dist
    if dist.get_world_size(group) == 1:
        return out9, hist
    o = out9 if torch.is_tensor(out9) else torch.as_tensor(out9, dtype=torch.float64)
    h = hist if torch.is_tensor(hist) else torch.as_tensor(hist, dtype=torch.int64)
    o = o.clone(); h = h.clone()
    dist.all_reduce(o, group=group)
    dist.all_reduce(h, group=group)
    return o, h
