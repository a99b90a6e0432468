"""The allocation-free viterbi entry point, sgb_viterbi_chained_noalloc.

Most of this library grows device buffers when it needs them.  A few callers cannot tolerate that -- services
that gave the whole GPU to one arena, loops that cannot take the device-wide stall of cudaMalloc -- so one entry
point keeps its scratch (the uint8 backpointers of the chained viterbi) in a fixed 4 MiB array inside the library
and takes device pointers for everything else.  (It still makes non-allocating CUDA calls and launches a kernel,
so it is not meant for cudaLaunchHostFunc callbacks.)

The test drives it through the C ABI exactly like such a caller would (here torch owns the device buffers) and
checks its paths against the ordinary, allocating entry point sgb_viterbi(decode = chained) AND against the CPU
oracle's chained viterbi: the integer paths must be identical to both.  It makes two calls in a row with different inputs, because
every call reuses the same array and must not be disturbed by what the previous one left behind, and it checks
that a problem larger than the fixed array is refused with a clear message.  The file name sorts last so a
`pytest -x` run reaches it only after every other test; it also passes when run on its own in a fresh process.
"""
import ctypes

import numpy as np
import pytest

from signals_b200 import _lib, hmm
from oracle import cpu_oracle as oracle

pytestmark = pytest.mark.gpu


def _problem(seed, T, S):
    rng = np.random.default_rng(seed)
    e = np.log(rng.dirichlet(np.ones(S) * 0.5, size=T + 2))
    e[rng.random(e.shape) < 0.08] = -np.inf
    tr = np.log(rng.dirichlet(np.ones(S), size=S))
    obs = rng.integers(1, T + 3, size=T).astype(np.int32)
    return np.asfortranarray(e), np.asfortranarray(tr), obs


def _noalloc(lib, torch, e, tr, obs, init_mode=0):
    dev = torch.device("cuda:0")
    d_e = torch.from_numpy(np.ascontiguousarray(e.ravel(order="F"))).to(dev)
    d_tr = torch.from_numpy(np.ascontiguousarray(tr.ravel(order="F"))).to(dev)
    d_obs = torch.from_numpy(obs).to(dev)
    d_path = torch.full((obs.size,), -7, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.sgb_viterbi_chained_noalloc(d_e.data_ptr(), e.shape[0], d_tr.data_ptr(), d_obs.data_ptr(), obs.size,
                                         tr.shape[0], init_mode, d_path.data_ptr(), ctypes.c_void_p(stream))
    hmm.check(rc)
    torch.cuda.synchronize()
    return d_path.cpu().numpy()


def test_noalloc_viterbi_matches_allocating_entry_point_twice_in_a_row():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    # two different problems back to back: the second one is smaller, so stale backpointers of the first
    # remain in the shared array beyond its extent and must not matter
    for seed, T, S in ((101, 4000, 12), (202, 700, 33)):
        e, tr, obs = _problem(seed, T, S)
        got = _noalloc(lib, torch, e, tr, obs)
        ref = hmm.viterbi(e, tr, obs, decode="chained")          # existing entry point (allocates its scratch)
        assert np.array_equal(got.astype(np.float64), ref), (seed, int((got != ref).sum()))
        # the independent comparand: the oracle's textbook backtrace (additions and comparisons only, so the
        # deterministic flavour differs from libm only in the -log(S) of the first column)
        orc = oracle.viterbi(e, tr, obs, decode=oracle.DECODE_CHAINED, math=oracle.MATH_DET)
        assert np.array_equal(got.astype(np.float64), orc), (seed, int((got != orc).sum()))
    # and once more on the first problem: identical answer again
    e, tr, obs = _problem(101, 4000, 12)
    assert np.array_equal(_noalloc(lib, torch, e, tr, obs).astype(np.float64), hmm.viterbi(e, tr, obs, decode="chained"))


def test_noalloc_viterbi_refuses_problems_beyond_its_fixed_scratch():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    S, T = 128, 40000                     # 5,120,000 backpointers > 4,194,304
    dev = torch.device("cuda:0")
    d_e = torch.zeros(8, dtype=torch.float64, device=dev)
    d_i = torch.ones(8, dtype=torch.int32, device=dev)
    rc = lib.sgb_viterbi_chained_noalloc(d_e.data_ptr(), 1, d_e.data_ptr(), d_i.data_ptr(), T, S, 0, d_i.data_ptr(), None)
    assert rc != 0
    msg = lib.sgb_last_error().decode()
    assert "4194304" in msg and "5120000" in msg
    # a T so large that T * S wraps around 2^64 must be refused too, not slip under the limit
    rc = lib.sgb_viterbi_chained_noalloc(d_e.data_ptr(), 1, d_e.data_ptr(), d_i.data_ptr(), 1 << 57, 128, 0, d_i.data_ptr(), None)
    assert rc != 0 and "4194304" in lib.sgb_last_error().decode()
    # an emission matrix without rows cannot hold any observation
    rc = lib.sgb_viterbi_chained_noalloc(d_e.data_ptr(), 0, d_e.data_ptr(), d_i.data_ptr(), 4, 2, 0, d_i.data_ptr(), None)
    assert rc != 0 and "no rows" in lib.sgb_last_error().decode()
    # exactly at the limit is accepted as a size (T * S == 4 MiB): build a real problem of that size
    e, tr, obs = _problem(7, 32768, 128)
    got = _noalloc(lib, torch, e, tr, obs)
    assert np.array_equal(got.astype(np.float64), hmm.viterbi(e, tr, obs, decode="chained"))
    assert np.array_equal(got.astype(np.float64), oracle.viterbi(e, tr, obs, decode=oracle.DECODE_CHAINED, math=oracle.MATH_DET))
