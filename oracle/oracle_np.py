"""NumPy/SciPy twin of the reference's HMM functions -- TEST INFRASTRUCTURE ONLY.

A readable, loop-for-loop transliteration of the R sources used to cross-check
oracle/oracle.c (which is the faster C restatement).  Function names and argument
order follow the reference:

    logspace_addcpp   src/code.cpp:5-8
    viterbi           src/code.cpp:11-45
    viterbiR          R/viterbi.R:1-34
    HaplotypeHMM      R/callHSCN.R:36-108
    alleleHMM         R/callASCN.R:41-103
    dbinom_log        stats::dbinom(log=TRUE)  (third party, restated; lgamma form)
    dbetabinom_log    VGAM::dbetabinom(log=TRUE) (third party, restated)
    dbinom_log_mp / dbetabinom_log_mp : 50-digit mpmath versions for pinning accuracy
"""
import math
import numpy as np
from scipy.special import gammaln


def logspace_addcpp(logx, logy):
    with np.errstate(invalid="ignore", over="ignore"):
        return np.fmax(logx, logy) + np.log1p(np.exp(-np.abs(logx - logy)))


def dbinom_log(x, n, p):
    """elementwise over broadcastable arrays; NaN for p outside [0,1], -Inf outside support"""
    x, n, p = np.broadcast_arrays(np.asarray(x, float), np.asarray(n, float), np.asarray(p, float))
    out = np.full(x.shape, np.nan)
    okp = (p >= 0) & (p <= 1)
    with np.errstate(divide="ignore", invalid="ignore"):
        lc = gammaln(n + 1) - gammaln(x + 1) - gammaln(n - x + 1)
        body = lc + x * np.log(p) + (n - x) * np.log1p(-p)
    body = np.where(p == 0, np.where(x == 0, 0.0, -np.inf), body)
    body = np.where(p == 1, np.where(x == n, 0.0, -np.inf), body)
    body = np.where((x < 0) | (x > n), -np.inf, body)
    out[okp] = body[okp]
    return out


def dbetabinom_log(x, n, p, rho):
    if not rho > 0:
        return dbinom_log(x, n, p)
    x, n, p = np.broadcast_arrays(np.asarray(x, float), np.asarray(n, float), np.asarray(p, float))
    s = (1.0 - rho) / rho
    with np.errstate(divide="ignore", invalid="ignore"):
        a, b = p * s, (1.0 - p) * s
        lc = gammaln(n + 1) - gammaln(x + 1) - gammaln(n - x + 1)
        body = lc + (gammaln(a + x) + gammaln(b + n - x) - gammaln(s + n)) - (gammaln(a) + gammaln(b) - gammaln(a + b))
    big = math.exp(20)
    body = np.where((a > big) & (b > big), dbinom_log(x, n, p), body)
    body = np.where(p == 0, np.where(x == 0, 0.0, -np.inf), body)
    body = np.where(p == 1, np.where(x == n, 0.0, -np.inf), body)
    body = np.where((x < 0) | (x > n), -np.inf, body)
    out = np.full(x.shape, np.nan)
    okp = (p >= 0) & (p <= 1)
    out[okp] = body[okp]
    return out


def _density(x, n, p, likelihood, rho):
    return dbinom_log(x, n, p) if likelihood == "binomial" else dbetabinom_log(x, n, p, rho)


def _transition(minor_cn, selftransitionprob):
    S = len(minor_cn)
    if selftransitionprob == 0.0:
        tp = np.full((S, S), 1.0 / S)
    else:
        with np.errstate(divide="ignore", invalid="ignore"):
            tp = np.full((S, S), (1.0 - selftransitionprob) / (S - 1))
        np.fill_diagonal(tp, selftransitionprob)
    with np.errstate(divide="ignore"):
        return np.log(tp)


def viterbi(emission, transition, observations, _initial=None):
    """returns 0-based states; emission is bins x states; observations are 1-based"""
    emission = np.asarray(emission, float).T          # states x bins   (:14)
    S = transition.shape[0]
    obs = np.asarray(observations, dtype=np.int64) - 1
    T = len(obs)
    initial = -math.log(S) if _initial is None else _initial
    T1 = np.zeros((S, T)); T2 = np.zeros((S, T), dtype=np.int64)
    T1[:, 0] = initial + emission[:, obs[0]]
    for j in range(1, T):
        for i in range(S):
            probs = (T1[:, j - 1] + transition[:, i]) + emission[i, obs[j]]
            T2[i, j] = int(np.argmax(probs))           # first maximum, like which_max
            T1[i, j] = probs[T2[i, j]]
    MLP = np.zeros(T)
    MLP[T - 1] = T2[int(np.argmax(T1[:, T - 1])), T - 1]
    for i in range(T - 1, 0, -1):
        MLP[i - 1] = T2[int(np.argmax(T1[:, i])), i]
    return MLP


def viterbiR(emission, transition, observations):
    S = transition.shape[0]
    return viterbi(emission, transition, observations, _initial=math.log(1.0 / S))


def viterbi_chained(emission, transition, observations):
    """textbook Viterbi with backtrace (the opt-in decode='chained' mode)"""
    emission = np.asarray(emission, float).T
    S = transition.shape[0]
    obs = np.asarray(observations, dtype=np.int64) - 1
    T = len(obs)
    T1 = np.zeros((S, T)); T2 = np.zeros((S, T), dtype=np.int64)
    T1[:, 0] = -math.log(S) + emission[:, obs[0]]
    for j in range(1, T):
        for i in range(S):
            probs = (T1[:, j - 1] + transition[:, i]) + emission[i, obs[j]]
            T2[i, j] = int(np.argmax(probs)); T1[i, j] = probs[T2[i, j]]
    path = np.zeros(T)
    cur = int(np.argmax(T1[:, T - 1])); path[T - 1] = cur
    for i in range(T - 1, 0, -1):
        cur = T2[cur, i]; path[i - 1] = cur
    return path


def HaplotypeHMM(n, x, binstates, minor_cn, loherror=0.02, selftransitionprob=0.999, eps=1e-12,
                 likelihood="binomial", rho=0.0, Abias=0.0, viterbiver="cpp"):
    n = np.asarray(n, float); x = np.asarray(x, float)
    binstates = np.asarray(binstates, float).copy()
    minor_cn = np.asarray(minor_cn, float)
    binstates[binstates == 0] = 1
    mm = np.tile(minor_cn, (len(binstates), 1))
    tt = np.tile(binstates[:, None], (1, len(minor_cn)))
    with np.errstate(divide="ignore", invalid="ignore"):
        p = mm / tt
    p[mm < tt / 2] += loherror
    p[mm > tt / 2] -= loherror
    l = _density(x[:, None], n[:, None], p, likelihood, rho)
    if eps > 0.0:
        l = logspace_addcpp(l, math.log(eps))
    if Abias > 0.0:
        sel = mm < tt / 2
        l[sel] = logspace_addcpp(l[sel], math.log(Abias))
    l[np.isnan(l)] = -np.inf
    l[mm > tt] = -np.inf
    tr = _transition(minor_cn, selftransitionprob)
    obs = np.arange(1, len(binstates) + 1)
    res = viterbi(l, tr, obs)
    if viterbiver == "R":
        res = viterbiR(l, tr, obs)
    return {"minorcn": res, "l": l}


def alleleHMM(n, x, binstates, minor_cn, loherror=0.02, selftransitionprob=0.999, eps=1e-12, rho=0.0,
              likelihood="binomial"):
    n = np.asarray(n, float); x = np.asarray(x, float)
    binstates = np.asarray(binstates, float)
    minor_cn = np.asarray(minor_cn, float)
    mm = np.tile(minor_cn, (len(binstates), 1))
    tt = np.tile(binstates[:, None], (1, len(minor_cn)))
    with np.errstate(divide="ignore", invalid="ignore"):
        p = mm / tt
    p[mm < tt / 2] += loherror
    p[mm > tt / 2] -= loherror
    l1 = _density(x[:, None], n[:, None], p, likelihood, rho)
    l2 = _density((n - x)[:, None], n[:, None], p, likelihood, rho)
    l = logspace_addcpp(l1, l2)
    if eps > 0.0:
        l = logspace_addcpp(l, math.log(eps))
    l[np.isnan(l)] = -np.inf
    l[mm > tt / 2] = -np.inf
    tr = _transition(minor_cn, selftransitionprob)
    res = viterbi(l, tr, np.arange(1, len(binstates) + 1))
    return {"minorcn": res, "l": l}


# ---------------------------------------------------------------- 50-digit pins
def dbinom_log_mp(x, n, p):
    import mpmath as mp
    mp.mp.dps = 50
    x, n, p = int(x), int(n), mp.mpf(p)
    if p < 0 or p > 1:
        return float("nan")
    if x < 0 or x > n:
        return -math.inf
    if p == 0:
        return 0.0 if x == 0 else -math.inf
    if p == 1:
        return 0.0 if x == n else -math.inf
    return mp.log(mp.binomial(n, x)) + x * mp.log(p) + (n - x) * mp.log1p(-p)


def dbetabinom_log_mp(x, n, p, rho):
    import mpmath as mp
    mp.mp.dps = 50
    x, n = int(x), int(n)
    p, rho = mp.mpf(p), mp.mpf(rho)
    s = (1 - rho) / rho
    a, b = p * s, (1 - p) * s
    lbeta = lambda u, v: mp.loggamma(u) + mp.loggamma(v) - mp.loggamma(u + v)
    return mp.log(mp.binomial(n, x)) + lbeta(a + x, b + n - x) - lbeta(a, b)
