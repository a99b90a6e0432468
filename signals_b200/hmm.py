"""Host-side mirror of the reference's R interface for the HSCN / ASCN HMM path.

The reference is an R package; R is not available where this repository is built and tested,
so the host layer above the C ABI is written in Python with the SAME function names, argument
names, defaults and return shapes as the R functions it stands in for (the R shim a maintainer
drops into the package is in rpkg/):

    logspace_addcpp(logx, logy)                         src/code.cpp:5-8, R/RcppExports.R:4-6
    logspace_add(logx, logy)                            R/callASCN.R:9-11
    viterbi(emission, transition, observations)         src/code.cpp:11-45, R/RcppExports.R:8-10
    viterbiR(emission, transition, observations)        R/viterbi.R:1-34
    HaplotypeHMM(...)                                   R/callHSCN.R:36-108
    alleleHMM(...)                                      R/callASCN.R:41-103
    assignHaplotypeHMM(CNBAF, minor_cn, ...)            R/callHSCN.R:135-172
    assignalleleHMM(CNBAF, minor_cn, ...)               R/callASCN.R:128-161
    callHaplotypeSpecificCN_(CNBAF, ...)                R/callHSCN.R:217-295  (.callHaplotypeSpecificCN_)
    callAlleleSpecificCN_(CNBAF, ...)                   R/callASCN.R:372-447  (pass-2 loop + post block)
    add_states(df)                                      R/util.R:1078-1097

All numerics run on the GPU through libsignals_b200.so; nothing here computes a likelihood or a
state path on the CPU, and there is no fallback when the library or a device is missing.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from ._lib import SignalsB200Error, check, make_params  # noqa: F401


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _as_counts(a, name):
    a = np.asarray(a)
    if a.dtype.kind == "f":
        if a.size and (np.isnan(a).any() or (np.rint(a) != a).any()):
            raise ValueError("%s must be integer valued" % name)
    if a.size and (a.max() > 2 ** 31 - 1 or a.min() < -2 ** 31):
        raise ValueError("%s does not fit in a 32-bit integer" % name)
    return np.ascontiguousarray(a, dtype=np.int32)


def _check_minor_cn(minor_cn):
    m = np.asarray(minor_cn, dtype=np.float64)
    if m.ndim != 1 or m.size < 1 or not np.array_equal(m, np.arange(m.size)):
        raise NotImplementedError("the GPU path requires minor_cn = 0:maxCN (as every reference driver passes, "
                                  "R/callHSCN.R:232)")
    if m.size > 32:
        raise NotImplementedError("maxCN > 31 is not supported by the fused kernels")
    return int(m.size)


# ---------------------------------------------------------------------------------- Rcpp exports
def logspace_addcpp(logx, logy):
    out = ctypes.c_double(0.0)
    check(_lib.load().sgb_logspace_add(float(logx), float(logy), ctypes.byref(out)))
    return out.value


def logspace_add(logx, logy):
    a, b = np.broadcast_arrays(np.asarray(logx, dtype=np.float64), np.asarray(logy, dtype=np.float64))
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    out = np.empty_like(a)
    check(_lib.load().sgb_logspace_add_vec(_ptr(a), _ptr(b), a.size, _ptr(out)))
    return out


def _viterbi(emission, transition, observations, init_mode, decode):
    e = np.asfortranarray(emission, dtype=np.float64)
    if e.ndim != 2:
        raise ValueError("emission must be a bins x states matrix")
    tr = np.asfortranarray(transition, dtype=np.float64)
    if tr.ndim != 2 or tr.shape[0] != tr.shape[1]:
        raise ValueError("transition must be a square matrix")
    if e.shape[1] != tr.shape[0]:
        raise ValueError("emission and transition disagree on the number of states")
    obs = _as_counts(observations, "observations")
    path = np.zeros(obs.size, dtype=np.int32)
    check(_lib.load().sgb_viterbi(_ptr(e), e.shape[0], _ptr(tr), _ptr(obs), obs.size, tr.shape[0], int(init_mode),
                                  _lib.DECODE_CHAINED if decode == "chained" else _lib.DECODE_REFERENCE, _ptr(path)))
    return path.astype(np.float64)        # the R function returns a numeric vector of 0-based states


def viterbi(emission, transition, observations, decode="reference"):
    return _viterbi(emission, transition, observations, 0, decode)


def viterbiR(emission, transition, observations):
    return _viterbi(emission, transition, observations, 1, "reference")


# ---------------------------------------------------------------------------------- array-level batch
_TYPE_CODE = {np.dtype(np.uint8): _lib.T_U8, np.dtype(np.uint16): _lib.T_U16, np.dtype(np.int32): _lib.T_I32,
              np.dtype(np.float64): _lib.T_F64}


def _as_column(a, name):
    """a contiguous array in one of the element types the library takes as is (uint8, uint16, int32, float64); other
    integer types are converted to int32 (range-checked), other floats to float64"""
    a = np.asarray(a)
    if a.dtype in _TYPE_CODE:
        return np.ascontiguousarray(a)
    if a.dtype.kind == "f":
        return np.ascontiguousarray(a, dtype=np.float64)
    return _as_counts(a, name)


def decode_arrays(n, x, cn, seg_off, model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12,
                  rho=0.0, Abias=0.0, selftransitionprob=0.999, decode="reference", want_loglik=False,
                  n_devices=1, out_dtype=np.uint8):
    """Decode every sequence of a sorted long table given as flat arrays.  Returns the states (uint8 unless out_dtype
    says otherwise; and the per-sequence column-major `l` blocks when want_loglik).  Columns may be int32, float64
    (integer valued: R numeric vectors), uint16 or uint8, each on its own."""
    lib = _lib.load()
    seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
    P = make_params(model, likelihood, n_states, loherror, eps, rho, Abias, selftransitionprob, decode)
    if want_loglik:
        # the `l` matrix path takes int32 counts
        n = _as_counts(n, "totalcounts"); x = _as_counts(x, "alleleB")
        cn = np.asarray(cn)
        if cn.dtype != np.uint8:
            cn = _as_counts(cn, "state")
        cn = np.ascontiguousarray(cn)
        if not (n.size == x.size == cn.size):
            raise ValueError("n, x and cn must have the same length")
        if n_devices > 1:
            raise ValueError("want_loglik is single-device only")
        out = np.zeros(n.size, dtype=np.uint8)
        ll = np.empty(n.size * int(n_states), dtype=np.float64)
        check(lib.sgb_hmm_decode(ctypes.byref(P), n.size, _ptr(n), _ptr(x), _ptr(cn), cn.dtype.itemsize,
                                 seg_off.size - 1, _ptr(seg_off), _ptr(out), _ptr(ll)))
        return out, ll
    n = _as_column(n, "totalcounts"); x = _as_column(x, "alleleB"); cn = _as_column(cn, "state")
    if not (n.size == x.size == cn.size):
        raise ValueError("n, x and cn must have the same length")
    out = np.zeros(n.size, dtype=np.dtype(out_dtype))
    check(lib.sgb_hmm_decode_typed(ctypes.byref(P), n.size, _ptr(n), _TYPE_CODE[n.dtype], _ptr(x), _TYPE_CODE[x.dtype],
                                   _ptr(cn), _TYPE_CODE[cn.dtype], seg_off.size - 1, _ptr(seg_off), _ptr(out),
                                   _TYPE_CODE[out.dtype], int(n_devices)))
    return out


class ResidentColumns:
    """The three columns of a sorted long table kept on the device for several decode passes (sgb_cols_*): what the two HMM
    passes of callHaplotypeSpecificCN (R/callHSCN.R:883-935) and callAlleleSpecificCN's third one share.

        cols = ResidentColumns(n, x, cn, seg_off)          # one conversion + upload
        p1 = cols.decode(loherror=0.02)                     # pass 1
        p2 = cols.decode(loherror=0.031, likelihood="betabinomial", rho=0.02)
        cols.close()
    """

    def __init__(self, n, x, cn, seg_off):
        self._lib = _lib.load()
        n = _as_column(n, "totalcounts"); x = _as_column(x, "alleleB"); cn = _as_column(cn, "state")
        if not (n.size == x.size == cn.size):
            raise ValueError("n, x and cn must have the same length")
        seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        self.N = int(n.size)
        self._h = ctypes.c_void_p()
        check(self._lib.sgb_cols_upload(self.N, _ptr(n), _TYPE_CODE[n.dtype], _ptr(x), _TYPE_CODE[x.dtype], _ptr(cn),
                                        _TYPE_CODE[cn.dtype], seg_off.size - 1, _ptr(seg_off), ctypes.byref(self._h)))

    def decode(self, model="hscn", likelihood="binomial", n_states=12, loherror=0.02, eps=1e-12, rho=0.0, Abias=0.0,
               selftransitionprob=0.999, decode="reference", out_dtype=np.uint8):
        if not self._h:
            raise SignalsB200Error("the columns have been released")
        P = make_params(model, likelihood, n_states, loherror, eps, rho, Abias, selftransitionprob, decode)
        out = np.zeros(self.N, dtype=np.dtype(out_dtype))
        check(self._lib.sgb_cols_decode(self._h, ctypes.byref(P), _ptr(out), _TYPE_CODE[out.dtype]))
        return out

    def close(self):
        if self._h:
            self._lib.sgb_cols_free(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def postprocess_arrays(cn, path):
    """device version of the clamp block + numeric add_states columns (integer codes)"""
    cn = np.asarray(cn)
    if cn.dtype != np.uint8:
        cn = _as_counts(cn, "state")
    cn = np.ascontiguousarray(cn)
    path = np.ascontiguousarray(path, dtype=np.uint8)
    N = cn.size
    A = np.empty(N, np.int32); B = np.empty(N, np.int32); smin = np.empty(N, np.int32)
    loh = np.empty(N, np.uint8); phase = np.empty(N, np.uint8); sp = np.empty(N, np.uint8)
    baf = np.empty(N, np.float64)
    check(_lib.load().sgb_postprocess(N, _ptr(cn), cn.dtype.itemsize, _ptr(path), _ptr(A), _ptr(B), _ptr(smin),
                                      _ptr(loh), _ptr(phase), _ptr(sp), _ptr(baf)))
    return dict(A=A, B=B, state_min=smin, LOH=loh, phase=phase, state_phase=sp, state_BAF=baf)


def pass_stats_arrays(n, x, cn, path, n_states=12):
    """exact whole-table sums the drivers need between the two HMM passes (computed on the device):
    returns (sum_alleleB, sum_totalcounts, n_rows) over the A-Hom rows and the A x B histogram of rows with B > 0"""
    n = _as_counts(n, "totalcounts"); x = _as_counts(x, "alleleB")
    cn = np.asarray(cn)
    if cn.dtype != np.uint8:
        cn = _as_counts(cn, "state")
    cn = np.ascontiguousarray(cn)
    path = np.ascontiguousarray(path, dtype=np.uint8)
    sums = np.zeros(3, np.int64); hist = np.zeros(int(n_states) * int(n_states), np.int64)
    check(_lib.load().sgb_pass_stats(n.size, _ptr(n), _ptr(x), _ptr(cn), cn.dtype.itemsize, _ptr(path), int(n_states),
                                     _ptr(sums), _ptr(hist)))
    return sums, hist.reshape(int(n_states), int(n_states))


def infer_loherror(n, x, cn, path, maxloherror=0.035, n_states=12):
    """infloherror of callHaplotypeSpecificCN (R/callHSCN.R:898-902): weighted.mean(BAF, w = totalcounts) over the
    A-Hom bins, capped at maxloherror"""
    sums, _ = pass_stats_arrays(n, x, cn, path, n_states)
    if sums[1] == 0:
        return float("nan")
    return min(float(sums[0]) / float(sums[1]), float(maxloherror))


class DevicePlan:
    """Device-resident decode: inputs/outputs are torch CUDA tensors (or anything with data_ptr())."""

    def __init__(self, N, seg_off, n_states=12, decode="reference"):
        self._lib = _lib.load()
        seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        self.N, self.nseg, self.n_states = int(N), int(seg_off.size - 1), int(n_states)
        h = ctypes.c_void_p()
        check(self._lib.sgb_plan_create(self.N, self.nseg, _ptr(seg_off), self.n_states,
                                        _lib.DECODE_CHAINED if decode == "chained" else _lib.DECODE_REFERENCE,
                                        ctypes.byref(h)))
        self._h = h

    def decode(self, params, d_n, d_x, d_cn, d_out, d_loglik=None, stream=0):
        itemsize = d_cn.element_size()
        check(self._lib.sgb_hmm_decode_device(ctypes.byref(params), self._h, d_n.data_ptr(), d_x.data_ptr(),
                                              d_cn.data_ptr(), itemsize, d_out.data_ptr(),
                                              d_loglik.data_ptr() if d_loglik is not None else None,
                                              ctypes.c_void_p(stream)))

    def check(self, stream=0):
        check(self._lib.sgb_plan_check(self._h, ctypes.c_void_p(stream)))

    def enable_timing(self, on=True):
        check(self._lib.sgb_plan_enable_timing(self._h, 1 if on else 0))

    def read_timing(self, max_records=64):
        """(pack_ms[], decode_ms[]) of the most recent calls, measured with CUDA events on the launch stream"""
        pk = np.zeros(max_records, np.float32); dc = np.zeros(max_records, np.float32)
        k = ctypes.c_int32(0)
        check(self._lib.sgb_plan_read_timing(self._h, max_records, _ptr(pk), _ptr(dc), ctypes.byref(k)))
        return pk[: k.value].copy(), dc[: k.value].copy()

    def read_timing3(self, max_records=64):
        """(prep_ms[], emit_ms[], decode_ms[]): preparation (k_pack or the split schedule's partition kernels), k_emit
        (0 on the fused schedule), k_decode / k_recur -- CUDA events on the launch stream"""
        a = np.zeros(max_records, np.float32); b = np.zeros(max_records, np.float32); c = np.zeros(max_records, np.float32)
        k = ctypes.c_int32(0)
        check(self._lib.sgb_plan_read_timing3(self._h, max_records, _ptr(a), _ptr(b), _ptr(c), ctypes.byref(k)))
        return a[: k.value].copy(), b[: k.value].copy(), c[: k.value].copy()

    def split_info(self, stream=0):
        """(groups, emission-tile bytes) of the last split-schedule call; (0, 0) after a fused one"""
        a = ctypes.c_int64(0); b = ctypes.c_int64(0)
        check(self._lib.sgb_plan_split_info(self._h, ctypes.c_void_p(stream), ctypes.byref(a), ctypes.byref(b)))
        return int(a.value), int(b.value)

    def reserve_split(self):
        check(self._lib.sgb_plan_reserve_split(self._h))

    def scratch_bytes(self):
        return int(self._lib.sgb_plan_scratch_bytes(self._h))

    def last_staging(self):
        """4 = split schedule (k_emit + k_recur), 2 / 3 = fused kernel reading the caller's columns in place, 1 / 0 = fused
        kernel via k_pack's records, -1 = no call yet"""
        return int(self._lib.sgb_plan_last_staging(self._h))

    def close(self):
        if self._h:
            self._lib.sgb_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------- single-sequence HMMs
def _single(model, n, x, binstates, minor_cn, loherror, selftransitionprob, eps, likelihood, rho, Abias):
    if likelihood not in ("binomial", "betabinomial"):
        raise ValueError("Likelihood model must be one of binomial or betabinomial")
    S = _check_minor_cn(minor_cn)
    n = _as_counts(n, "n"); x = _as_counts(x, "x"); cn = _as_counts(binstates, "binstates")
    T = n.size
    seg = np.array([0, T], dtype=np.int64)
    path, ll = decode_arrays(n, x, cn, seg, model, likelihood, S, loherror, eps, rho, Abias, selftransitionprob,
                             want_loglik=True)
    l = ll.reshape(S, T).T.copy() if T else np.zeros((0, S))
    return {"minorcn": path.astype(np.float64), "l": l}


def HaplotypeHMM(n, x, binstates, minor_cn, loherror=0.02, selftransitionprob=0.999, eps=1e-12,
                 likelihood="binomial", rho=0.0, Abias=0.0, viterbiver="cpp"):
    res = _single("hscn", n, x, binstates, minor_cn, loherror, selftransitionprob, eps, likelihood, rho, Abias)
    if viterbiver == "R":
        # R/callHSCN.R:101-105: the pure-R twin is run on the same l / transition matrix
        S = len(minor_cn)
        if selftransitionprob == 0.0:
            tp = np.full((S, S), 1.0 / S)
        else:
            tp = np.full((S, S), (1.0 - selftransitionprob) / (S - 1))
            np.fill_diagonal(tp, selftransitionprob)
        with np.errstate(divide="ignore"):
            res["minorcn"] = viterbiR(res["l"], np.log(tp), np.arange(1, len(res["minorcn"]) + 1))
    return res


def alleleHMM(n, x, binstates, minor_cn, loherror=0.02, selftransitionprob=0.999, eps=1e-12, rho=0.0,
              likelihood="binomial"):
    return _single("ascn", n, x, binstates, minor_cn, loherror, selftransitionprob, eps, likelihood, rho, 0.0)


# ---------------------------------------------------------------------------------- data.frame level
def _segments(df, by_cell=True):
    """row-run boundaries of (cell_id, chr) (or chr only) in the CURRENT row order, like unique(CNBAF$chr)
    followed by dplyr::filter does for a table sorted by chromosome (R/callHSCN.R:150-154)"""
    chr_codes = df["chr"].astype("category").cat.codes.to_numpy()
    change = np.ones(len(df), dtype=bool)
    if len(df) > 1:
        change[1:] = chr_codes[1:] != chr_codes[:-1]
        if by_cell and "cell_id" in df.columns:
            cell_codes = df["cell_id"].astype("category").cat.codes.to_numpy()
            change[1:] |= cell_codes[1:] != cell_codes[:-1]
    starts = np.nonzero(change)[0]
    return np.concatenate((starts, [len(df)])).astype(np.int64)


def _assign(model, CNBAF, minor_cn, eps, loherror, selftransitionprob, likelihood, rho, Abias):
    S = _check_minor_cn(minor_cn)
    df = CNBAF.copy()
    # the reference gathers rows per unique(chr) and writes the concatenated result back in row order
    # (R/callHSCN.R:150-167); that is only self-consistent when each chromosome's rows are contiguous
    if len(df) and df["chr"].nunique() != len(_segments(df, by_cell=False)) - 1:
        raise ValueError("rows of each chromosome must be contiguous")
    seg = _segments(df, by_cell=False)
    path = decode_arrays(df["totalcounts"].to_numpy(), df["alleleB"].to_numpy(), df["state"].to_numpy(), seg, model,
                         likelihood, S, loherror, eps, rho, Abias, selftransitionprob)
    df["state_min"] = path.astype(np.float64)
    return df


def assignHaplotypeHMM(CNBAF, minor_cn, eps=1e-12, loherror=0.02, selftransitionprob=0.999, pb=None,
                       likelihood="binomial", rho=0.0, Abias=0.0, viterbiver="cpp"):
    return _assign("hscn", CNBAF, minor_cn, eps, loherror, selftransitionprob, likelihood, rho, Abias)


def assignalleleHMM(CNBAF, minor_cn, eps=1e-12, loherror=0.02, selftransitionprob=0.999, pb=None, rho=0.0,
                    likelihood="binomial"):
    return _assign("ascn", CNBAF, minor_cn, eps, loherror, selftransitionprob, likelihood, rho, 0.0)


_PHASE = np.array(["Balanced", "A", "B"], dtype=object)
_STATE_PHASE = np.array(["Balanced", "A-Gained", "B-Gained", "A-Hom", "B-Hom"], dtype=object)


def _attach_states(df, cols):
    """materialise the string columns of add_states from the device-computed integer codes"""
    A, B, st = cols["A"], cols["B"], df["state"].to_numpy().astype(np.int64)
    df["A"] = A; df["B"] = B
    a_s, b_s = A.astype(str), B.astype(str)
    df["state_AS_phased"] = np.char.add(np.char.add(a_s, "|"), b_s)
    hi = np.maximum(st - B, B).astype(str); lo = np.minimum(st - B, B).astype(str)
    state_as = np.char.add(np.char.add(hi, "|"), lo).astype(object)
    big = st > 4
    state_as[big] = st[big].astype(str)
    df["state_AS"] = state_as
    df["state_min"] = cols["state_min"]
    df["LOH"] = np.where(cols["LOH"] == 1, "LOH", "NO")
    df["phase"] = _PHASE[cols["phase"]]
    df["state_phase"] = _STATE_PHASE[cols["state_phase"]]
    df["state_BAF"] = cols["state_BAF"]
    return df


def add_states(df):
    """R/util.R:1078-1097 for a table that already has A, B and state"""
    out = df.copy()
    path = np.ascontiguousarray(out["B"].to_numpy(), dtype=np.uint8)
    cols = postprocess_arrays(out["state"].to_numpy(), path)
    # add_states does not re-clamp: keep the caller's A / B
    cols["A"] = out["A"].to_numpy(); cols["B"] = out["B"].to_numpy()
    cols["state_min"] = np.minimum(cols["A"], cols["B"])
    return _attach_states(out, cols)


def _batch(model, CNBAF, eps, loherror, maxCN, selftransitionprob, likelihood, rho, Abias, n_devices):
    if likelihood not in ("binomial", "betabinomial"):
        raise ValueError("Likelihood model must be one of binomial or betabinomial")
    minor_cn = np.arange(0, maxCN + 1)
    S = _check_minor_cn(minor_cn)
    # "Make sure dataframe is in chromosome position order"  R/callHSCN.R:241-242 (chr is character)
    df = CNBAF.copy()
    df["chr"] = df["chr"].astype(str)
    df = df.sort_values(["cell_id", "chr", "start"], kind="stable").reset_index(drop=True)
    seg = _segments(df, by_cell=True)
    path = decode_arrays(df["totalcounts"].to_numpy(), df["alleleB"].to_numpy(), df["state"].to_numpy(), seg, model,
                         likelihood, S, loherror, eps, rho, Abias, selftransitionprob, n_devices=n_devices)
    cols = postprocess_arrays(df["state"].to_numpy(), path)
    return _attach_states(df, cols)


def callHaplotypeSpecificCN_(CNBAF, eps=1e-12, loherror=0.02, maxCN=10, selftransitionprob=0.999,
                             progressbar=True, ncores=1, likelihood="binomial", rho=0.0, Abias=0.0,
                             viterbiver="cpp", n_devices=1):
    """.callHaplotypeSpecificCN_ (R/callHSCN.R:217-295): one batched launch instead of the per-cell loop.
    `progressbar`, `ncores` and `viterbiver` are accepted for signature parity and ignored."""
    return _batch("hscn", CNBAF, eps, loherror, maxCN, selftransitionprob, likelihood, rho, Abias, n_devices)


def callAlleleSpecificCN_(CNBAF, eps=1e-12, loherror=0.02, maxCN=10, selftransitionprob=0.999, progressbar=True,
                          ncores=1, likelihood="binomial", rho=0.0, n_devices=1):
    """pass-2 loop of callAlleleSpecificCN (R/callASCN.R:382-414) plus its post block (:416-447)"""
    return _batch("ascn", CNBAF, eps, loherror, maxCN, selftransitionprob, likelihood, rho, 0.0, n_devices)
