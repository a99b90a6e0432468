"""Host-side logic that needs no GPU: synthetic generator, segment detection / ordering, sharding."""
import numpy as np
import pandas as pd

from signals_b200 import dist, hmm, synth


def test_generator_is_deterministic_and_dlp_shaped():
    a = synth.make_cohort(30, seed=5)
    b = synth.make_cohort(30, seed=5)
    for k in ("n", "x", "cn", "seg_off", "truth_b"):
        assert np.array_equal(a[k], b[k])
    assert a["n"].size == 30 * 5700 and a["seg_off"].size == 30 * 23 + 1
    assert np.array_equal(np.diff(a["seg_off"])[:23], np.array(synth.LENS_5700))
    assert a["n"].min() >= 1 and (a["x"] <= a["n"]).all() and (a["x"] >= 0).all()
    assert a["cn"].max() <= 11 and (a["truth_b"] <= a["cn"]).all()
    c = synth.make_cohort(30, seed=6)
    assert not np.array_equal(a["x"], c["x"])


def test_ragged_variant_keeps_every_sequence():
    d = synth.make_cohort(10, seed=2, ragged_drop=0.05)
    L = np.diff(d["seg_off"])
    assert (L >= 1).all() and d["seg_off"][-1] == d["n"].size
    assert 0.93 * 57000 < d["n"].size < 0.97 * 57000


def test_segments_follow_row_runs_in_r_sort_order():
    d = synth.make_cohort(3, lens=(4, 3, 5), seed=1)
    d["chrom"] = d["chrom"]            # chromosome indices 0..2 -> "1","2","3"
    df = synth.to_frame(d)
    df = df.sample(frac=1.0, random_state=0)
    s = df.copy(); s["chr"] = s["chr"].astype(str)
    s = s.sort_values(["cell_id", "chr", "start"], kind="stable").reset_index(drop=True)
    seg = hmm._segments(s, by_cell=True)
    assert np.array_equal(np.diff(seg), np.tile([4, 3, 5], 3))
    # character ordering of chromosomes, as data.table's order() on a character column: "1" < "10" < "2" < "X"
    t = pd.DataFrame({"cell_id": "c", "chr": ["2", "10", "X", "1"], "start": 1})
    assert t.sort_values(["cell_id", "chr", "start"])["chr"].tolist() == ["1", "10", "2", "X"]


def test_minor_cn_must_be_0_to_maxcn():
    import pytest
    assert hmm._check_minor_cn(range(12)) == 12
    with pytest.raises(NotImplementedError):
        hmm._check_minor_cn([0, 2, 4])
    with pytest.raises(ValueError):
        hmm.make_params(likelihood="poisson")


def test_shard_sequences_partitions_everything():
    d = synth.make_cohort(17, seed=3, ragged_drop=0.05)
    so = d["seg_off"]
    for world in (1, 2, 3, 4, 8):
        cuts = [dist.shard_sequences(so, r, world) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == so.size - 1
        for (a, b), (c, e) in zip(cuts[:-1], cuts[1:]):
            assert b == c
        bins = [so[b] - so[a] for a, b in cuts]
        assert max(bins) - min(bins) <= 2 * 468 + 1


def test_bench_shards_are_cells_of_one_cohort():
    """bench.py's weak scaling: rank 0 holds exactly the 1-GPU data; the other ranks hold other cells of the same
    clones, so the algorithmic work per rank is equal to well under a percent"""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(__file__), "..", "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    base = synth.make_cohort(200, seed=77)
    same = synth.make_cohort(200, seed=77, shard=None)
    assert all(np.array_equal(base[k], same[k]) for k in ("n", "x", "cn", "seg_off"))
    work = [bench.flop_eq(base, "hscn", "binomial") / base["n"].size]
    for r in (1, 2, 3):
        sh = synth.make_cohort(200, seed=77, shard=r)
        assert not np.array_equal(sh["n"], base["n"]) and np.array_equal(sh["seg_off"], base["seg_off"])
        # the same clone profiles: the set of whole-chromosome copy numbers seen at sequence starts is the cohort's
        assert set(np.unique(sh["cn"][sh["seg_off"][:-1]])) <= set(np.unique(base["cn"])) | set(np.unique(sh["cn"]))
        work.append(bench.flop_eq(sh, "hscn", "binomial") / sh["n"].size)
    assert max(work) / min(work) < 1.02, work
    # the deterministic part of the JSON contract that needs no GPU
    d = bench.make_data("cfg2", rank=0)
    assert d["n"].size == 1000 * 5700 and d["seg_off"].size == 1000 * 23 + 1
