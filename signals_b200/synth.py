"""Seeded synthetic DLP-shaped input for the HSCN/ASCN HMM path.

Shapes and value distributions follow SURVEY.md 8(d), which in turn restates the
reference's simulator (R/sim-data.R:13-115: per-chromosome copy number, gamma-distributed
depth scaled by copy number, binomial / beta-binomial B-allele counts, `loherror` in LOH
regions).  The reference draws from R's RNG streams, which cannot be reproduced outside R,
so this generator uses NumPy's PCG64 and is the source of every test / bench input here.

Rows are emitted already in the order the reference's batch loop establishes
(order(cell_id, chr, start), R/callHSCN.R:241-242): cell-major, then chromosome, then bin;
`seg_off` delimits the (cell, chromosome) sequences the per-chromosome HMM runs on.
"""
from __future__ import annotations

import numpy as np

# 23 sequences per cell (chr1-22, X); largest-remainder split of the bin count by hg19 chr length
LENS_5700 = (468, 457, 372, 359, 340, 321, 299, 275, 265, 254, 253, 251, 216, 202, 192, 170, 152, 147,
             111, 118, 90, 96, 292)
LENS_30000 = (2463, 2403, 1957, 1889, 1787, 1691, 1572, 1446, 1395, 1339, 1334, 1322, 1138, 1061, 1013,
              893, 802, 771, 584, 623, 476, 507, 1534)
# the reference's real 500 kb grid (data/dlpbins.rda) without chrY
LENS_DLPBINS = (345, 362, 308, 282, 271, 265, 241, 224, 182, 208, 203, 207, 166, 144, 117, 115, 122, 123,
                94, 103, 60, 51, 170)
assert sum(LENS_5700) == 5700 and sum(LENS_30000) == 30000 and sum(LENS_DLPBINS) == 4363

CHROMS = tuple([str(i) for i in range(1, 23)] + ["X"])

# P(total copy number = 0..11) for a whole chromosome / piece
CN_PROBS = np.array([.01, .09, .50, .18, .12, .04, .03, .01, .008, .005, .004, .003])
CN_PROBS = CN_PROBS / CN_PROBS.sum()


def make_cohort(n_cells: int, lens=LENS_5700, seed: int = 20260001, likelihood: str = "binomial",
                rho: float = 0.02, loherror: float = 0.02, n_clones: int = 4, ragged_drop: float = 0.0,
                extra_piece_rate: float = 0.3, max_cn: int = 11, shard=None):
    """Returns a dict of flat arrays:
         n, x      int32 [N]   total / B-allele read counts per cell-bin  (totalcounts, alleleB)
         cn        uint8 [N]   total copy number state per cell-bin       (state)
         seg_off   int64 [n_cells*len(lens)+1]  sequence boundaries
         truth_b   uint8 [N]   simulated B-allele copy number
         cell, chrom  int32 [nseg]  cell index / chromosome index of every sequence
       shard: None, or the index of a further shard of the same cohort (see below)
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    lens = np.asarray(lens, dtype=np.int64)
    nchr = len(lens)
    probs = CN_PROBS[: max_cn + 1] / CN_PROBS[: max_cn + 1].sum()

    # clone-level whole-chromosome states
    clone_c = rng.choice(len(probs), size=(n_clones, nchr), p=probs)
    clone_b = (rng.random((n_clones, nchr)) * (clone_c + 1)).astype(np.int64)
    if shard is not None:
        # another shard of the SAME cohort (same clones, other cells): the clone-level draws above come from
        # `seed`, everything per cell from a stream of its own -- what sharding one tumour's cells over ranks gives
        rng = np.random.Generator(np.random.PCG64([seed, int(shard)]))
    clone_of_cell = rng.integers(0, n_clones, size=n_cells)

    nseg = n_cells * nchr
    seq_len = np.tile(lens, n_cells)
    seq_cell = np.repeat(np.arange(n_cells), nchr)
    seq_chr = np.tile(np.arange(nchr), n_cells)
    n_pieces = 1 + rng.poisson(extra_piece_rate, size=nseg)
    n_pieces = np.minimum(n_pieces, np.maximum(seq_len, 1))

    # piece table: first piece of a sequence carries the clone state, the others are re-drawn
    tot_pieces = int(n_pieces.sum())
    piece_seq = np.repeat(np.arange(nseg), n_pieces)
    first = np.ones(tot_pieces, dtype=bool)
    first[1:] = piece_seq[1:] != piece_seq[:-1]
    pc = rng.choice(len(probs), size=tot_pieces, p=probs)
    pb = (rng.random(tot_pieces) * (pc + 1)).astype(np.int64)
    cl = clone_of_cell[seq_cell[piece_seq]]
    pc = np.where(first, clone_c[cl, seq_chr[piece_seq]], pc)
    pb = np.where(first, clone_b[cl, seq_chr[piece_seq]], pb)

    # piece lengths: k-1 distinct cut points inside the sequence
    piece_len = np.empty(tot_pieces, dtype=np.int64)
    start = np.concatenate(([0], np.cumsum(n_pieces)))
    single = n_pieces == 1
    piece_len[start[:-1][single]] = seq_len[single]
    for s in np.nonzero(~single)[0]:
        k, T = int(n_pieces[s]), int(seq_len[s])
        cuts = np.sort(rng.choice(np.arange(1, T), size=k - 1, replace=False))
        piece_len[start[s]: start[s] + k] = np.diff(np.concatenate(([0], cuts, [T])))

    c = np.repeat(pc, piece_len)
    b = np.repeat(pb, piece_len)
    N = c.size
    assert N == int(seq_len.sum())

    # read depth: n = round(c * Gamma(1.76, rate 0.09)), redrawn until >= 1; hom-del: stray reads
    n = np.zeros(N, dtype=np.int64)
    todo = np.nonzero(c > 0)[0]
    while todo.size:
        g = rng.gamma(1.76, 1.0 / 0.09, size=todo.size)
        v = np.rint(c[todo] * g).astype(np.int64)
        n[todo] = v
        todo = todo[v < 1]
    hd = c == 0
    n[hd] = 1 + rng.poisson(1.0, size=int(hd.sum()))

    with np.errstate(divide="ignore", invalid="ignore"):
        p = np.where(c > 0, b / np.maximum(c, 1), 0.5)
    p = np.where((c > 0) & (b == 0), loherror, p)
    if likelihood == "betabinomial":
        s = (1.0 - rho) / rho
        pp = rng.beta(np.maximum(p * s, 1e-12), np.maximum((1.0 - p) * s, 1e-12))
        pp = np.where(p >= 1.0, 1.0, pp)
        pp = np.where(hd, 0.5, pp)
    else:
        pp = p
    x = rng.binomial(n, np.clip(pp, 0.0, 1.0))

    seg_off = np.concatenate(([0], np.cumsum(seq_len))).astype(np.int64)
    out = dict(n=n.astype(np.int32), x=x.astype(np.int32), cn=c.astype(np.uint8), truth_b=b.astype(np.uint8),
               seg_off=seg_off, cell=seq_cell.astype(np.int32), chrom=seq_chr.astype(np.int32),
               lens=tuple(int(v) for v in lens), n_cells=n_cells)
    if ragged_drop > 0.0:
        out = drop_bins(out, ragged_drop, rng)
    return out


def drop_bins(d, frac, rng):
    """ragged variant: remove a fraction of cell-bins (keeps at least one bin per sequence)"""
    N = d["n"].size
    keep = rng.random(N) >= frac
    keep[d["seg_off"][:-1]] = True
    csum = np.concatenate(([0], np.cumsum(keep)))
    out = dict(d)
    for k in ("n", "x", "cn", "truth_b"):
        out[k] = d[k][keep]
    out["seg_off"] = csum[d["seg_off"]].astype(np.int64)
    return out


def to_frame(d, start_step=500_000):
    """long CNBAF-style table (cell_id, chr, start, end, state, totalcounts, alleleB, ...) as the
    reference's combineBAFCN would hand to the HMM batch loop (R/combinedata.R:303-329)"""
    import pandas as pd
    seg_len = np.diff(d["seg_off"])
    cell = np.repeat(d["cell"], seg_len)
    chrom = np.repeat(d["chrom"], seg_len)
    pos = np.arange(d["n"].size) - np.repeat(d["seg_off"][:-1], seg_len)
    width = len(str(max(int(d["n_cells"]) - 1, 1)))
    df = pd.DataFrame({
        "cell_id": np.char.add("cell_", np.char.zfill(cell.astype(str), width)),
        "chr": np.asarray(CHROMS)[chrom],
        "start": pos * start_step + 1,
        "end": (pos + 1) * start_step,
        "state": d["cn"].astype(np.int32),
        "totalcounts": d["n"],
        "alleleB": d["x"],
    })
    df["alleleA"] = df["totalcounts"] - df["alleleB"]
    df["BAF"] = df["alleleB"] / df["totalcounts"]
    df["copy"] = df["state"].astype(float)
    return df
