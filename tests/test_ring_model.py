"""Model check of k_decode's in-place staging arithmetic (signals_b200/csrc/hmm_kernels.cu, STAGE RAW).

The kernel streams every lane's own sequence out of the caller's columns in aligned 4-bin blocks through a 3-slot
ring (issue two blocks ahead, wait for all but the newest copy group, read bin t from block (sh + t) >> 2).
This file restates those index formulas in Python and checks, for many (first row, length, column length)
combinations, the invariants the CUDA code relies on:
  * every copy reads only rows inside the column, from a 4-row aligned start, clamped at the column end;
  * a block is issued before the step that first reads it and at least one wait has passed since;
  * a ring slot is never overwritten while a later step still reads the block it holds;
  * the value read for bin t is row ostart + t.
It is a model of the arithmetic, kept next to the GPU parity tests that exercise the real kernel."""
import numpy as np
import pytest


def simulate(ostart, length, rows, tmax):
    """one lane of a warp whose longest sequence has tmax bins; returns the rows read per step"""
    sh = ostart & 3
    a0 = ostart - sh
    span = sh + length
    avail = min(rows - a0, 0x7fffffff)
    ring = [None, None, None]           # slot -> (block index, landed?)
    pending = []                        # copy groups in flight, oldest first: list of (slot, block, row, nrows) or None
    reads = []

    def issue(B, slot):
        if 4 * B < span:
            left = avail - 4 * B
            assert left > 0
            row = a0 + 4 * B
            nrows = 4 if left >= 4 else left
            assert row % 4 == 0 and 0 <= row and row + nrows <= rows          # aligned, inside the column
            pending.append((slot, B, row, nrows))
        else:
            pending.append(None)                                              # empty group, still committed

    def wait_all_but_newest():
        while len(pending) > 1:
            g = pending.pop(0)
            if g is not None:
                slot, B, row, nrows = g
                ring[slot] = (B, row, nrows)

    slot_m = 0
    if tmax > 0:
        issue(0, 0)
        issue(1, 1)
    for t in range(tmax):
        if (t & 3) == 0:
            if t:
                slot_m = 0 if slot_m == 2 else slot_m + 1
            target = 2 if slot_m == 0 else slot_m - 1
            # the slot being refilled must not hold a block any later step of this lane still reads
            held = ring[target]
            if held is not None:
                assert held[0] < (sh + t) >> 2, (t, held)        # older than every block read from this step on
            assert all(g is None or g[0] != target for g in pending)          # and no copy into it is still in flight
            issue((t >> 2) + 2, target)
            wait_all_but_newest()
        if t < length:
            u = sh + (t & 3)
            slot = slot_m + (u >> 2)
            slot = slot - 3 if slot >= 3 else slot
            blk = ring[slot]
            assert blk is not None and blk[0] == (sh + t) >> 2, (t, slot, blk)          # the right block has landed
            k = u & 3
            assert k < blk[2]                                                            # inside the copied part
            reads.append(blk[1] + k)
    return reads


@pytest.mark.parametrize("seed", range(6))
def test_ring_reads_the_lane_rows_in_order(seed):
    rng = np.random.default_rng(seed)
    for _ in range(400):
        length = int(rng.choice([0, 1, 2, 3, 4, 5, 7, 8, 9, 31, 100, int(rng.integers(1, 300))]))
        ostart = int(rng.integers(0, 50))
        tail = int(rng.choice([0, 0, 1, 2, 3, 5, 40]))               # rows of the column after this sequence
        rows = ostart + length + tail
        tmax = length + int(rng.choice([0, 0, 1, 3, 17]))            # another lane of the warp may be longer
        assert simulate(ostart, length, rows, tmax) == list(range(ostart, ostart + length))


def test_ring_column_end_is_clamped():
    # the last sequence ends with the column, mid-block: the final copy must be shorter than 4 rows
    assert simulate(ostart=5, length=6, rows=11, tmax=6) == list(range(5, 11))
    assert simulate(ostart=0, length=1, rows=1, tmax=1) == [0]
    assert simulate(ostart=3, length=2, rows=5, tmax=9) == [3, 4]
