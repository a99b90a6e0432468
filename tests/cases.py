"""Shared test inputs: hand-written edge cases and seeded cohorts (used by the golden generator too)."""
import numpy as np

from signals_b200 import synth

# (name, model, likelihood, S, loherror, eps, rho, Abias, selftransitionprob)
PARAM_SETS = [
    ("hscn_binom_default", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("hscn_binom_stp999", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.999),
    ("hscn_bb", "hscn", "betabinomial", 12, 0.03, 1e-12, 0.02, 0.0, 0.95),
    ("ascn_binom", "ascn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("ascn_bb", "ascn", "betabinomial", 12, 0.025, 1e-12, 0.02, 0.0, 0.95),
    ("hscn_iid", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.0),
    ("hscn_stp1", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 1.0),
    ("hscn_noeps_noloh", "hscn", "binomial", 12, 0.0, 0.0, 0.0, 0.0, 0.95),
    ("hscn_abias", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.1, 0.95),
    ("hscn_maxcn2", "hscn", "binomial", 3, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("hscn_maxcn10", "hscn", "binomial", 11, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("hscn_maxcn19", "hscn", "binomial", 20, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("ascn_bb_rho0", "ascn", "betabinomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("ascn_noeps", "ascn", "binomial", 12, 0.02, 0.0, 0.0, 0.0, 0.9),
    ("hscn_bb_bigloh", "hscn", "betabinomial", 12, 0.6, 1e-12, 0.3, 0.0, 0.5),
    ("hscn_stp_below_uniform", "hscn", "binomial", 12, 0.02, 1e-12, 0.0, 0.0, 0.01),
    ("ascn_stp_below_uniform", "ascn", "betabinomial", 12, 0.02, 1e-12, 0.05, 0.0, 0.03),
    ("hscn_maxcn14", "hscn", "binomial", 15, 0.02, 1e-12, 0.0, 0.0, 0.95),
    ("ascn_bb_maxcn15", "ascn", "betabinomial", 16, 0.02, 1e-12, 0.02, 0.0, 0.95),
]


def param_dict(ps):
    name, model, lik, S, loh, eps, rho, Ab, stp = ps
    return dict(model=model, likelihood=lik, S=S, loherror=loh, eps=eps, rho=rho, Abias=Ab, selftransitionprob=stp)


def edge_sequences():
    """list of (name, n, x, cn) single sequences covering the reference's corner cases"""
    out = []
    out.append(("T1", [20], [7], [2]))
    out.append(("T2", [20, 30], [7, 29], [2, 3]))
    out.append(("x0_xn_n1", [1, 1, 50, 50, 1, 3], [0, 1, 0, 50, 1, 0], [2, 2, 3, 3, 1, 4]))
    out.append(("cn0_homdel", [1, 2, 4, 4, 3, 2, 2, 5, 1], [0, 1, 2, 3, 1, 0, 2, 2, 1], [0] * 9))
    out.append(("cn0_inside", [30, 28, 2, 3, 33, 31], [10, 9, 1, 2, 11, 12], [3, 3, 0, 0, 3, 3]))
    out.append(("cn_gt_maxstate", [200, 210, 190, 220], [100, 110, 95, 20], [11, 11, 14, 20]))
    out.append(("n0_bins", [0, 10, 0, 12], [0, 5, 0, 6], [2, 2, 2, 2]))
    out.append(("x_gt_n", [10, 10, 10], [5, 12, 5], [2, 2, 2]))
    out.append(("ties_symmetric", [2, 2, 2, 2, 2, 2], [1, 1, 1, 1, 1, 1], [2, 2, 2, 2, 2, 2]))
    out.append(("big_counts", [3000, 2500, 5000, 2047, 2048, 2049], [1500, 100, 4999, 1000, 1024, 7],
                [2, 3, 4, 2, 2, 5]))
    out.append(("cn1_only", [15, 12, 18, 11], [0, 12, 1, 11], [1, 1, 1, 1]))
    rng = np.random.default_rng(7)
    T = 300
    cn = np.repeat([2, 3, 1, 4, 6, 2], 50)
    b = np.repeat([1, 1, 0, 2, 5, 0], 50)
    n = rng.poisson(12 * np.maximum(cn, 1)) + 1
    x = rng.binomial(n, np.where(b == 0, 0.02, b / cn))
    out.append(("piecewise300", n.tolist(), x.tolist(), cn.tolist()))
    return [(nm, np.array(a, np.int32), np.array(b_, np.int32), np.array(c, np.int32)) for nm, a, b_, c in out]


def concat_sequences(seqs):
    n = np.concatenate([s[1] for s in seqs]); x = np.concatenate([s[2] for s in seqs])
    cn = np.concatenate([s[3] for s in seqs])
    seg = np.concatenate(([0], np.cumsum([len(s[1]) for s in seqs]))).astype(np.int64)
    return n, x, cn, seg


def small_cohort(likelihood="binomial", n_cells=40, seed=11, ragged=0.05):
    return synth.make_cohort(n_cells, lens=synth.LENS_DLPBINS, seed=seed, likelihood=likelihood, ragged_drop=ragged)


def random_case(seed):
    """(n, x, cn, seg_off, kwargs of hmm.decode_arrays) of one randomised model + input; see the tests using it.
    Seeds >= 24 are the long draws: a third of the sequences has 500..3,000 bins (the 100 kb grid of BASELINE
    configs[4] reaches 2,463), and every fourth of them uses all 32 states with total copy numbers 31 / 40 / 255."""
    rng = np.random.default_rng(1000 + seed)
    long_mode = seed >= 24
    S = int(rng.choice([2, 3, 7, 12, 12, 12, 13, 16, 20, 32]))
    if long_mode and seed % 4 == 0:
        S = 32
    nseq = int(rng.integers(20, 90))
    lens = np.where(rng.random(nseq) < 0.2, rng.integers(1, 5, nseq), rng.integers(5, 160, nseq))
    if long_mode:
        lens = np.where(rng.random(nseq) < 0.33, rng.integers(500, 3001, nseq), lens)
    cn_menu = [0, 1, 2, 2, 3, 4, 6, 9, 11, 15, 16, 19, 25]
    if S == 32:
        cn_menu = cn_menu + [31, 31, 40, 255]
    n_l, x_l, c_l = [], [], []
    for T in lens:
        pieces = 1 + rng.poisson(1.0)
        cuts = np.sort(rng.integers(0, T + 1, pieces - 1))
        bounds = np.concatenate(([0], cuts, [T]))
        cn = np.zeros(T, np.int64)
        for a, b in zip(bounds[:-1], bounds[1:]):
            cn[a:b] = rng.choice(cn_menu) if rng.random() < 0.8 else rng.integers(0, 30)
        depth = rng.choice([3.0, 20.0, 200.0, 900.0])
        n = rng.poisson(depth * np.maximum(cn, 1) * rng.gamma(2.0, 0.5, T))
        n[rng.random(T) < 0.03] = 0
        b = (rng.random(T) * (cn + 1)).astype(np.int64)
        pb = np.where(cn > 0, b / np.maximum(cn, 1), 0.5)
        x = rng.binomial(n, np.clip(pb, 0.01, 0.99))
        bad = rng.random(T) < 0.01
        x[bad] = n[bad] + rng.integers(1, 4, int(bad.sum()))          # outside the support
        n_l.append(n); x_l.append(x); c_l.append(cn)
    n = np.concatenate(n_l).astype(np.int32); x = np.concatenate(x_l).astype(np.int32); cn = np.concatenate(c_l)
    seg = np.concatenate(([0], np.cumsum(lens))).astype(np.int64)
    model = str(rng.choice(["hscn", "ascn"]))
    lik = str(rng.choice(["binomial", "binomial", "betabinomial"]))
    kw = dict(model=model, likelihood=lik, n_states=S, loherror=float(rng.choice([0.0, 0.01, 0.02, 0.3])),
              eps=float(rng.choice([0.0, 1e-12, 1e-12, 1e-6])), rho=float(rng.choice([0.0, 0.005, 0.02, 0.1])) if lik == "betabinomial" else 0.0,
              Abias=float(rng.choice([0.0, 0.0, 0.0, 0.05])),
              selftransitionprob=float(rng.choice([0.0, 0.5 / S, 0.5, 0.95, 0.95, 0.9999, 1.0])))
    cn = cn.astype(np.uint8) if rng.random() < 0.5 else cn.astype(np.int32)
    return n, x, cn, seg, kw


N_RANDOM_CASES = 40
