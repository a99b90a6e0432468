/* oracle.c -- CPU restatement of the signals HSCN/ASCN HMM hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or as
 * the timed CPU baseline.  The product (signals_b200/) never links, imports
 * or calls it and has no CPU fallback.
 *
 * PARITY: the reference ships no golden vectors / known-answer tests for
 * viterbi, logspace_addcpp, HaplotypeHMM or alleleHMM and neither
 * R, Rcpp nor VGAM exist in this image, so the reference as a whole cannot be run.
 *   - decode side (orc_viterbi, orc_logspace_add): PINNED against oracle/_ref, the
 *     reference's own src/code.cpp compiled verbatim against a mock of the few Rcpp
 *     classes it uses (oracle/rcpp_mock/Rcpp.h); tests/test_oracle.py::test_ref_*.
 *   - emission side (HaplotypeHMM / alleleHMM are R code over stats::dbinom and
 *     VGAM::dbetabinom): PARITY UNPINNED -- restated line by line below (citations on
 *     every function) and pinned against 50-digit mpmath densities instead.
 *
 * Third-party arithmetic that is NOT under /root/reference (restated from the
 * published algorithms; versions unpinned by the reference's DESCRIPTION):
 *   stats::dbinom    (base R >= 3.5)  -- log-density of Binomial(n,p) at x;
 *                    argument checking order and limits as in R's dbinom.c.
 *   VGAM::dbetabinom (Suggests: VGAM) -- lchoose(n,x) + lbeta(a+x, b+n-x) - lbeta(a,b),
 *                    a = p(1-rho)/rho, b = (1-p)(1-rho)/rho; rho = 0 and shapes
 *                    above exp(20) fall back to dbinom.
 *   Rcpp sugar max / which_max -- first strict maximum (same as R which.max,
 *                    R/viterbi.R:18,25,29).
 *
 * Two math flavours are provided behind one code path:
 *   ORC_MATH_LIBM (0): glibc exp/log/log1p/lgamma, plain a*b+c arithmetic --
 *                      the natural C restatement; also the timed CPU baseline.
 *   ORC_MATH_DET  (1): the deterministic functions of signals_b200/csrc/sgb_math.h
 *                      (header-only, shared with the device) so that emissions and
 *                      therefore state paths are bit-identical to the CUDA kernels.
 *   ORC_MATH_RMATH (2): like LIBM, but the binomial log-density is computed the way base R does
 *                      (nmath/dbinom.c dbinom_raw with bd0 and stirlerr, C. Loader 2000), restated from the
 *                      published algorithm -- used to show that the lgamma form stays within the 1e-9
 *                      tolerance of what stats::dbinom itself returns.
 * The emission logic itself (probabilities, masks, special cases, recursion,
 * decode rule) is written here independently of the CUDA code.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include "../signals_b200/csrc/sgb_math.h"

#define ORC_MATH_LIBM 0
#define ORC_MATH_DET 1
#define ORC_MATH_RMATH 2   /* binomial density by R's own algorithm (Loader's saddle-point dbinom_raw); rest as LIBM */
#define ORC_MODEL_HSCN 0
#define ORC_MODEL_ASCN 1
#define ORC_LIK_BINOMIAL 0
#define ORC_LIK_BETABINOMIAL 1
#define ORC_DECODE_REFERENCE 0
#define ORC_DECODE_CHAINED 1

typedef struct {
    int32_t model;      /* ORC_MODEL_* */
    int32_t likelihood; /* ORC_LIK_*   */
    int32_t S;          /* number of states = maxCN + 1; minor_cn = 0..S-1 (R/callHSCN.R:232) */
    int32_t decode;     /* ORC_DECODE_* */
    double loherror;
    double eps;
    double rho;
    double Abias;
    double selftransitionprob;
} orc_params;

static const double ORC_LT[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
static const double ORC_ET[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
static const double ORC_LF[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

/* ------------------------------------------------------------------ math flavours */
static double m_log(double x, int math) { return math == ORC_MATH_DET ? sgb_log(x, ORC_LT) : log(x); }
static double m_log1p(double x, int math) { return math == ORC_MATH_DET ? sgb_log1p(x, ORC_LT) : log1p(x); }
static double m_lgamma(double x, int math) { return math == ORC_MATH_DET ? sgb_lgamma(x, ORC_LT) : lgamma(x); }
static double m_logfact(int64_t k, int math)
{
    return math == ORC_MATH_DET ? sgb_logfact(k, ORC_LF, ORC_LT) : lgamma((double)k + 1.0);
}
/* lc + k*lp + (n-k)*lq ; the det flavour fixes the rounding sequence to two nested fmas */
static double m_combine(double lc, double kd, double lp, double nkd, double lq, int math)
{
    if (math == ORC_MATH_DET) return sgb_fma(nkd, lq, sgb_fma(kd, lp, lc));
    return lc + kd * lp + nkd * lq;
}

/* src/code.cpp:5-8  logspace_addcpp */
double orc_logspace_add(double logx, double logy, int math)
{
    if (math == ORC_MATH_DET) return sgb_logspace_add(logx, logy, ORC_LT, ORC_ET);
    return fmax(logx, logy) + log1p(exp(-fabs(logx - logy)));
}

/* evaluate one det/libm primitive over a vector (used by tests to pin host == device bits) */
void orc_math_eval(int fn, const double* in, const double* in2, double* out, int64_t n, int math)
{
    for (int64_t i = 0; i < n; i++) {
        double x = in[i];
        switch (fn) {
        case 0: out[i] = math == ORC_MATH_DET ? sgb_exp(x, ORC_ET) : exp(x); break;
        case 1: out[i] = m_log(x, math); break;
        case 2: out[i] = math == ORC_MATH_DET ? sgb_log1p_01(x, ORC_LT) : log1p(x); break;
        case 3: out[i] = m_log1p(x, math); break;
        case 4: out[i] = m_lgamma(x, math); break;
        case 5: out[i] = orc_logspace_add(x, in2[i], math); break;
        case 6: out[i] = m_logfact((int64_t)x, math); break;
        default: out[i] = NAN;
        }
    }
}

/* returns 0 when this translation unit was compiled without fp contraction */
int orc_selftest(void)
{
    volatile double a = 1.0 + 0x1p-30, b = 1.0 - 0x1p-30, c = -1.0;
    double unfused = a * b + c;          /* a*b rounds to 1 - 2^-60 -> 1.0 ; + c = 0 */
    double fused = __builtin_fma(a, b, c);
    return (unfused == 0.0 && fused != 0.0) ? 0 : 1;
}

/* ------------------------------------------------------------------ R's dbinom_raw (restated) */
/* stirlerr(n) = log(n!) - log( sqrt(2 pi n) (n/e)^n ); exact table for n = 0, 0.5, ..., 15 (generated with
 * 60-digit mpmath), Stirling series above (nmath/stirlerr.c). */
static const double ORC_SFERR_HALVES[31] = {
    0x0.0p+0, 0x1.3a37a020b8c22p-3, 0x1.4c071bcda0a5bp-4, 0x1.c1098b28dcf33p-5, 0x1.52a9b923ea649p-5,
    0x1.0fab9626b44ffp-5, 0x1.c579a268d80b3p-6, 0x1.850ea113caf0ep-6, 0x1.54a2662fd78a9p-6, 0x1.2eea2e990f134p-6,
    0x1.10b4e513fcbedp-6, 0x1.eff15b81c9cc5p-7, 0x1.c6b167bebdf36p-7, 0x1.a3c5f8a1e7d1dp-7, 0x1.85d4d612e4a86p-7,
    0x1.6bdfcc7fbdb0ap-7, 0x1.552805e7b3076p-7, 0x1.411b75e41049cp-7, 0x1.2f4871b12ab64p-7, 0x1.1f553026fbce1p-7,
    0x1.10f9d4c0743a7p-7, 0x1.03fc2d49c8fb9p-7, 0x1.f0593088014f8p-8, 0x1.dac773cc5b3cdp-8, 0x1.c7018733aa9c6p-8,
    0x1.b4d04a067629fp-8, 0x1.a40514700f36cp-8, 0x1.947826547fe6cp-8, 0x1.86076c002d4a7p-8, 0x1.789583ffc5e8cp-8,
    0x1.6c08f6f194a10p-8};

static double orc_stirlerr(double n)
{
    const double S0 = 1.0 / 12.0, S1 = 1.0 / 360.0, S2 = 1.0 / 1260.0, S3 = 1.0 / 1680.0, S4 = 1.0 / 1188.0;
    if (n <= 15.0) {
        double nn = n + n;
        if (nn == (double)(int)nn) return ORC_SFERR_HALVES[(int)nn];
        return lgamma(n + 1.0) - (n + 0.5) * log(n) + n - 0.918938533204672741780329736406; /* log(sqrt(2 pi)) */
    }
    double nn = n * n;
    if (n > 500) return (S0 - S1 / nn) / n;
    if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
    if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
    return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

/* bd0(x, np) = x log(x/np) + np - x, evaluated without cancellation when x ~ np (nmath/bd0.c) */
static double orc_bd0(double x, double np)
{
    if (fabs(x - np) < 0.1 * (x + np)) {
        double v = (x - np) / (x + np);
        double s = (x - np) * v;
        if (fabs(s) < 2.2250738585072014e-308) return s;
        double ej = 2 * x * v;
        v = v * v;
        for (int j = 1; j < 1000; j++) {
            ej *= v;
            double s1 = s + ej / ((j << 1) + 1);
            if (s1 == s) return s1;
            s = s1;
        }
    }
    return x * log(x / np) + np - x;
}

/* dbinom_raw(x, n, p, q, give_log = TRUE) (nmath/dbinom.c) */
static double orc_dbinom_raw_log(double x, double n, double p, double q)
{
    if (p == 0) return x == 0 ? 0.0 : -INFINITY;
    if (q == 0) return x == n ? 0.0 : -INFINITY;
    if (x == 0) {
        if (n == 0) return 0.0;
        return (p < 0.1) ? -orc_bd0(n, n * q) - n * p : n * log(q);
    }
    if (x == n) return (q < 0.1) ? -orc_bd0(n, n * p) - n * q : n * log(p);
    if (x < 0 || x > n) return -INFINITY;
    double lc = orc_stirlerr(n) - orc_stirlerr(x) - orc_stirlerr(n - x) - orc_bd0(x, n * p) - orc_bd0(n - x, n * q);
    double lf = 1.837877066409345483560659472811 /* log(2 pi) */ + log(x) + log1p(-x / n);
    return lc - 0.5 * lf;
}

/* ------------------------------------------------------------------ emission densities */
/* stats::dbinom(k, n, p, log=TRUE) restated: argument checks in R's order
 * (p outside [0,1] or NaN -> NaN; k outside 0..n -> -Inf; p==0 / p==1 limits). */
static double orc_dbinom_log(int64_t k, int64_t n, double p, int math)
{
    if (p != p || p < 0.0 || p > 1.0) return NAN;
    if (k < 0 || k > n) return -INFINITY;
    if (p == 0.0) return k == 0 ? 0.0 : -INFINITY;
    if (p == 1.0) return k == n ? 0.0 : -INFINITY;
    if (math == ORC_MATH_RMATH) return orc_dbinom_raw_log((double)k, (double)n, p, 1.0 - p);
    double lc = (m_logfact(n, math) - m_logfact(k, math)) - m_logfact(n - k, math);
    return m_combine(lc, (double)k, m_log(p, math), (double)(n - k), m_log1p(-p, math), math);
}

/* VGAM::dbetabinom(k, n, p, rho, log=TRUE) restated (see header). */
static double orc_dbetabinom_log(int64_t k, int64_t n, double p, double rho, int math)
{
    if (!(rho > 0.0)) return orc_dbinom_log(k, n, p, math);
    if (p != p || p < 0.0 || p > 1.0) return NAN;
    if (k < 0 || k > n) return -INFINITY;
    if (p == 0.0) return k == 0 ? 0.0 : -INFINITY;
    if (p == 1.0) return k == n ? 0.0 : -INFINITY;
    double s = (1.0 - rho) / rho;
    double a = p * s;
    double b = (1.0 - p) * s;
    const double big = 485165195.40979028; /* exp(20), VGAM's Inf.shape */
    if (a > big && b > big) return orc_dbinom_log(k, n, p, math);
    double lc = (m_logfact(n, math) - m_logfact(k, math)) - m_logfact(n - k, math);
    double t1 = m_lgamma(a + (double)k, math) + m_lgamma(b + (double)(n - k), math);
    t1 = t1 - m_lgamma(s + (double)n, math);
    double t0 = (m_lgamma(a, math) + m_lgamma(b, math)) - m_lgamma(a + b, math);
    return lc + (t1 - t0);
}

static double orc_density(int64_t k, int64_t n, double p, const orc_params* P, int math)
{
    if (P->likelihood == ORC_LIK_BETABINOMIAL) return orc_dbetabinom_log(k, n, p, P->rho, math);
    return orc_dbinom_log(k, n, p, math);
}

/* The bins x states log-likelihood matrix `l` of HaplotypeHMM (R/callHSCN.R:48-80)
 * or alleleHMM (R/callASCN.R:50-81).  l_out is T x S column-major like the R matrix. */
void orc_emission(const int32_t* n, const int32_t* x, const int32_t* cn, int64_t T,
                  const orc_params* P, int math, double* l_out)
{
    const int S = P->S;
    const double logeps = P->eps > 0.0 ? m_log(P->eps, math) : 0.0;
    const double logAb = P->Abias > 0.0 ? m_log(P->Abias, math) : 0.0;
    for (int64_t t = 0; t < T; t++) {
        int64_t c = cn[t];
        if (P->model == ORC_MODEL_HSCN && c == 0) c = 1;              /* R/callHSCN.R:49 */
        for (int m = 0; m < S; m++) {
            double p = (double)m / (double)c;                          /* :53-55 ; 0/0 = NaN, m/0 = Inf */
            if (2 * (int64_t)m < c) p = p + P->loherror;               /* :56-57 */
            if (2 * (int64_t)m > c) p = p - P->loherror;               /* :58-59 */
            double l;
            if (P->model == ORC_MODEL_HSCN) {
                l = orc_density(x[t], n[t], p, P, math);               /* :61-66 */
            } else {
                double l1 = orc_density(x[t], n[t], p, P, math);       /* R/callASCN.R:62,67 */
                double l2 = orc_density((int64_t)n[t] - x[t], n[t], p, P, math); /* :63,69 */
                l = orc_logspace_add(l1, l2, math);                    /* :64,71 */
            }
            if (P->eps > 0.0) l = orc_logspace_add(l, logeps, math);   /* R/callHSCN.R:67-71, callASCN.R:74-79 */
            if (P->model == ORC_MODEL_HSCN && P->Abias > 0.0 && 2 * (int64_t)m < c)
                l = orc_logspace_add(l, logAb, math);                  /* R/callHSCN.R:72-78 */
            if (l != l) l = -INFINITY;                                 /* :79  / callASCN.R:80 */
            if (P->model == ORC_MODEL_HSCN) {
                if ((int64_t)m > c) l = -INFINITY;                     /* R/callHSCN.R:80 */
            } else {
                if (2 * (int64_t)m > c) l = -INFINITY;                 /* R/callASCN.R:81 */
            }
            l_out[t + T * (int64_t)m] = l;
        }
    }
}

/* log transition matrix, S x S column-major [from,to] (R/callHSCN.R:82-97, callASCN.R:83-96) */
void orc_transition(const orc_params* P, int math, double* tr)
{
    const int S = P->S;
    for (int to = 0; to < S; to++)
        for (int from = 0; from < S; from++) {
            double pr;
            if (P->selftransitionprob == 0.0) pr = 1.0 / (double)S;
            else pr = (from == to) ? P->selftransitionprob
                                   : (1.0 - P->selftransitionprob) / (double)(S - 1);
            tr[from + S * to] = m_log(pr, math);
        }
}

/* ------------------------------------------------------------------ viterbi (src/code.cpp:11-45) */
/* first strict maximum, like Rcpp sugar which_max / R which.max */
static int first_argmax(const double* v, int n, int stride)
{
    double best = v[0];
    int idx = 0;
    for (int i = 1; i < n; i++) {
        double cur = v[(int64_t)i * stride];
        if (cur > best) { best = cur; idx = i; }
    }
    return idx;
}

/* emission: nrow_e x S column-major (rows = bins), transition S x S column-major [from,to],
 * obs: 1-based row indices (length T).  path_out: 0-based states as doubles.
 * init_mode 0: initial = -log(S) (src/code.cpp:20), 1: log(1/S) (R/viterbi.R:3).
 * NaN emissions are treated as -Inf (the reference would index with NA: undefined). */
int orc_viterbi(const double* emission, int64_t nrow_e, const double* trans, const int32_t* obs,
                int64_t T, int S, int init_mode, int decode, int math, double* path_out)
{
    if (T <= 0) return 0;
    double* T1 = (double*)malloc(sizeof(double) * S * T);
    int32_t* T2 = (int32_t*)calloc((size_t)S * T, sizeof(int32_t));
    double* probs = (double*)malloc(sizeof(double) * S);
    if (!T1 || !T2 || !probs) { free(T1); free(T2); free(probs); return -1; }
    const double initial = init_mode == 0 ? -m_log((double)S, math) : m_log(1.0 / (double)S, math);
#define EM(state, j) (emission[(int64_t)(obs[j] - 1) + nrow_e * (int64_t)(state)])
    for (int i = 0; i < S; i++) {
        double e = EM(i, 0);
        if (e != e) e = -INFINITY;
        T1[i] = initial + e;                                           /* :25 */
    }
    for (int64_t j = 1; j < T; j++) {
        for (int i = 0; i < S; i++) {
            double e = EM(i, j);
            if (e != e) e = -INFINITY;
            for (int k = 0; k < S; k++)
                probs[k] = (T1[k + S * (j - 1)] + trans[k + S * i]) + e;   /* :29 */
            int am = first_argmax(probs, S, 1);
            T1[i + S * j] = probs[am];                                 /* :30 */
            T2[i + S * j] = am;                                        /* :31 */
        }
    }
#undef EM
    if (decode == ORC_DECODE_REFERENCE) {
        /* :35-42  the reference's decode: state emitted for bin i-1 is the best
         * predecessor of the column-wise best state of bin i; the last bin repeats it. */
        path_out[T - 1] = (double)T2[first_argmax(T1 + S * (T - 1), S, 1) + S * (T - 1)];
        for (int64_t i = T - 1; i > 0; i--) {
            int zm = first_argmax(T1 + S * i, S, 1);
            path_out[i - 1] = (double)T2[zm + S * i];
        }
    } else {
        /* textbook backtrace (opt-in extra, not what the reference does) */
        int cur = first_argmax(T1 + S * (T - 1), S, 1);
        path_out[T - 1] = (double)cur;
        for (int64_t i = T - 1; i > 0; i--) {
            cur = T2[cur + S * i];
            path_out[i - 1] = (double)cur;
        }
    }
    free(T1); free(T2); free(probs);
    return 0;
}

/* HaplotypeHMM (R/callHSCN.R:36-108) / alleleHMM (R/callASCN.R:41-103) for one sequence.
 * l_out (T x S col-major) may be NULL; path_out is T doubles. */
int orc_hmm_sequence(const int32_t* n, const int32_t* x, const int32_t* cn, int64_t T,
                     const orc_params* P, int math, double* path_out, double* l_out)
{
    if (T <= 0) return 0;
    const int S = P->S;
    double* l = l_out ? l_out : (double*)malloc(sizeof(double) * S * T);
    double* tr = (double*)malloc(sizeof(double) * S * S);
    int32_t* obs = (int32_t*)malloc(sizeof(int32_t) * T);
    if (!l || !tr || !obs) return -1;
    orc_emission(n, x, cn, T, P, math, l);
    orc_transition(P, math, tr);
    for (int64_t t = 0; t < T; t++) obs[t] = (int32_t)(t + 1);         /* seq_len(length(binstates)) */
    int rc = orc_viterbi(l, T, tr, obs, T, S, 0, P->decode, math, path_out);
    if (!l_out) free(l);
    free(tr); free(obs);
    return rc;
}

/* ------------------------------------------------------------------ batch (R/callHSCN.R:217-295) */
typedef struct {
    const orc_params* P; int math;
    const int32_t *n, *x, *cn; const int64_t* seg_off; int64_t nseg;
    uint8_t* state_out; int tid, nthreads; int rc;
} orc_job;

static void* orc_worker(void* arg)
{
    orc_job* J = (orc_job*)arg;
    int64_t maxT = 0;
    for (int64_t s = 0; s < J->nseg; s++) {
        int64_t T = J->seg_off[s + 1] - J->seg_off[s];
        if (T > maxT) maxT = T;
    }
    double* path = (double*)malloc(sizeof(double) * (maxT > 0 ? maxT : 1));
    for (int64_t s = J->tid; s < J->nseg; s += J->nthreads) {
        int64_t o = J->seg_off[s], T = J->seg_off[s + 1] - o;
        if (T <= 0) continue;
        if (orc_hmm_sequence(J->n + o, J->x + o, J->cn + o, T, J->P, J->math, path, NULL)) { J->rc = -1; break; }
        for (int64_t t = 0; t < T; t++) J->state_out[o + t] = (uint8_t)path[t];
    }
    free(path);
    return NULL;
}

/* Per-cell / per-chromosome loop of assignHaplotypeHMM (R/callHSCN.R:150-167) and
 * assignalleleHMM (R/callASCN.R:143-156) over a table already sorted by (cell_id, chr, start):
 * segment s covers rows seg_off[s] .. seg_off[s+1]-1.  nthreads mirrors mclapply(mc.cores). */
int orc_decode_batch(const orc_params* P, int math, int64_t N, const int32_t* n, const int32_t* x,
                     const int32_t* cn, int64_t nseg, const int64_t* seg_off, uint8_t* state_out, int nthreads)
{
    (void)N;
    if (nthreads < 1) nthreads = 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
    orc_job* jobs = (orc_job*)malloc(sizeof(orc_job) * nthreads);
    for (int i = 0; i < nthreads; i++) {
        jobs[i] = (orc_job){P, math, n, x, cn, seg_off, nseg, state_out, i, nthreads, 0};
        if (nthreads > 1) pthread_create(&th[i], NULL, orc_worker, &jobs[i]);
    }
    int rc = 0;
    if (nthreads == 1) orc_worker(&jobs[0]);
    for (int i = 0; i < nthreads; i++) {
        if (nthreads > 1) pthread_join(th[i], NULL);
        rc |= jobs[i].rc;
    }
    free(th); free(jobs);
    return rc;
}

/* ------------------------------------------------------------------ table-level post-processing */
/* HSCN: R/callHSCN.R:282-290 then add_states (R/util.R:1078-1097).
 * ASCN: R/callASCN.R:416-447 (the swap at :419-420 is overwritten by :424-429, so the
 * arithmetic is the same as HSCN; kept as the reference has it).
 * Outputs are small integer codes; strings are materialised by the host wrappers:
 *   phase: 0 Balanced, 1 A, 2 B ; state_phase: 0 Balanced 1 A-Gained 2 B-Gained 3 A-Hom 4 B-Hom
 *   state_BAF: round((B/state)/0.1)*0.1, NaN -> 0.5 (R rounds half to even) */
void orc_postprocess(int64_t N, const int32_t* cn, const uint8_t* path, int32_t* A, int32_t* B,
                     int32_t* state_min, uint8_t* loh, uint8_t* phase, uint8_t* state_phase,
                     double* state_baf)
{
    for (int64_t i = 0; i < N; i++) {
        int32_t st = cn[i];
        int32_t sm = path[i];
        if (sm < 0) sm = 0;
        int32_t a = st - sm, b = sm;
        if (b < 0) b = 0;
        if (a < 0) a = 0;
        if (b > st) b = st;
        if (a > st) a = st;
        A[i] = a; B[i] = b;
        int32_t mn = a < b ? a : b;
        state_min[i] = mn;
        loh[i] = mn == 0;
        phase[i] = (uint8_t)((b < a) ? 1 : (b > a) ? 2 : 0);
        state_phase[i] = (uint8_t)(1 * ((b < a) && (b != 0)) + 2 * ((b > a) && (a != 0)) +
                                   3 * ((b < a) && (b == 0)) + 4 * ((b > a) && (a == 0)));
        double q = ((double)b / (double)st) / 0.1;
        double baf = nearbyint(q) * 0.1;
        if (baf != baf) baf = 0.5;
        state_baf[i] = baf;
    }
}

/* ------------------------------------------------------------------ fix_assignments (R/callHSCN.R:1072-1286), as written */
/* dbinom(k, k + nk, p) / VGAM::dbetabinom(k, k + nk, p, rho) as DENSITIES (the reference does not pass log = TRUE).
 * The det flavour exponentiates with sgb_exp, which flushes below exp(-708) to 0 (R keeps subnormals down to exp(-745)). */
static double orc_fix_density(int64_t k, int64_t nk, double p, int lik_bb, double rho, int math)
{
    orc_params P;
    memset(&P, 0, sizeof P);
    P.likelihood = lik_bb ? ORC_LIK_BETABINOMIAL : ORC_LIK_BINOMIAL;
    P.rho = rho;
    double l = orc_density(k, k + nk, p, &P, math);
    if (l != l) return l;
    return math == ORC_MATH_DET ? sgb_exp(l, ORC_ET) : exp(l);
}
static double orc_fix_p(int num, int state, double loh)
{
    double p = (double)num / (double)state;
    if (p != p) p = 0.0;                    /* ifelse(is.nan(B / state), 0.0, B / state)   :1083 */
    if (p == 0.0) p = p + loh;              /* :1087-1091 */
    if (p == 1.0) p = p - loh;              /* :1092-1096 */
    return p;
}
static void orc_fix_clamp(int state, int smin, int32_t* A, int32_t* B)
{
    int a = state - smin, b = smin;         /* :1129-1134 */
    if (b < 0) b = 0;
    if (a < 0) a = 0;
    if (b > state) b = state;
    if (a > state) a = state;
    *A = a; *B = b;
}
/* rows in table order; seg_off delimits the (cell_id, chr) groups of the rleid(..., by = .(cell_id, chr)) calls */
int orc_fix_assignments(int64_t N, const int32_t* state, const int32_t* A, const int32_t* B, const int32_t* alleleA,
                        const int32_t* alleleB, int64_t nseg, const int64_t* seg_off, double loherror, int lik_bb, double rho,
                        int math, int32_t* A_out, int32_t* B_out)
{
    if (N <= 0) return 0;
    uint8_t* head = (uint8_t*)calloc((size_t)N + 1, 1);
    int32_t* A1 = (int32_t*)malloc(sizeof(int32_t) * N);
    int32_t* B1 = (int32_t*)malloc(sizeof(int32_t) * N);
    if (!head || !A1 || !B1) { free(head); free(A1); free(B1); return -1; }
    for (int64_t s = 0; s < nseg; s++) if (seg_off[s + 1] > seg_off[s]) head[seg_off[s]] = 1;
    head[N] = 1;
    /* ---- first block (:1074-1145 binomial, :1196-1262 beta-binomial): singletons */
    for (int64_t i = 0; i < N; i++) {
        /* nbins := .N by (rlid, chr, cell_id), rlid := rleid(state_AS_phased) by (cell_id, chr) */
        int same_prev = !head[i] && A[i - 1] == A[i] && B[i - 1] == B[i];
        int same_next = !head[i + 1] && A[i + 1] == A[i] && B[i + 1] == B[i];
        int nbins1 = !same_prev && !same_next;
        int Bprev = i > 0 ? B[i - 1] : 0;                  /* shift(B, fill = 0): over the whole table */
        int Bafter = i + 1 < N ? B[i + 1] : 0;             /* shift(B, fill = 0, type = "lead") */
        double pBafter = orc_fix_p(Bafter, state[i], loherror), pBprev = orc_fix_p(Bprev, state[i], loherror);
        double LLafter = orc_fix_density(alleleA[i], alleleB[i], pBafter, lik_bb, rho, math);
        if (LLafter != LLafter) LLafter = 0.0;             /* fifelse(is.na(LLafter), 0, LLafter) */
        double LLprev = orc_fix_density(alleleA[i], alleleB[i], pBprev, lik_bb, rho, math);
        LLprev = (LLprev != LLprev) ? 0.0 : LLafter;       /* fifelse(is.na(LLprev), 0, LLafter)   [sic, :1114] */
        int smin = (nbins1 && LLafter < LLprev) ? Bprev : B[i];      /* :1115-1119 */
        smin = (nbins1 && LLafter >= LLprev) ? Bafter : B[i];        /* :1120-1124 overwrites it */
        orc_fix_clamp(state[i], smin, &A1[i], &B1[i]);
    }
    /* ---- second block (:1147-1194 binomial: rleid by (cell_id, chr); :1264-1280 beta-binomial: rleid over the table) */
    for (int64_t i = 0; i < N;) {
        int64_t j = i;
        int64_t ta = 0, tb = 0;
        while (j < N && A1[j] == A1[i] && B1[j] == B1[i] && (j == i || lik_bb || !head[j])) { ta += alleleA[j]; tb += alleleB[j]; j++; }
        for (int64_t r = i; r < j; r++) {
            double pB = orc_fix_p(B1[r], state[r], loherror), pA = orc_fix_p(A1[r], state[r], loherror);
            double LLassigned = orc_fix_density(ta, tb, pB, lik_bb, rho, math);
            double LLother = orc_fix_density(ta, tb, pA, lik_bb, rho, math);
            int smin = (LLother < LLassigned) ? A1[r] : B1[r];       /* fifelse(LLother < LLassigned, A, B) */
            orc_fix_clamp(state[r], smin, &A_out[r], &B_out[r]);
        }
        i = j;
    }
    free(head); free(A1); free(B1);
    return 0;
}
/* margin report for the second block: |LLother - LLassigned| relative to their maximum, per row (0 when both are 0) */
int orc_fix_margins(int64_t N, const int32_t* state, const int32_t* A1, const int32_t* B1, const int32_t* alleleA,
                    const int32_t* alleleB, int64_t nseg, const int64_t* seg_off, double loherror, int lik_bb, double rho,
                    int math, double* margin_out)
{
    uint8_t* head = (uint8_t*)calloc((size_t)N + 1, 1);
    if (!head) return -1;
    for (int64_t s = 0; s < nseg; s++) if (seg_off[s + 1] > seg_off[s]) head[seg_off[s]] = 1;
    for (int64_t i = 0; i < N;) {
        int64_t j = i, ta = 0, tb = 0;
        while (j < N && A1[j] == A1[i] && B1[j] == B1[i] && (j == i || lik_bb || !head[j])) { ta += alleleA[j]; tb += alleleB[j]; j++; }
        for (int64_t r = i; r < j; r++) {
            double la = orc_fix_density(ta, tb, orc_fix_p(B1[r], state[r], loherror), lik_bb, rho, math);
            double lo = orc_fix_density(ta, tb, orc_fix_p(A1[r], state[r], loherror), lik_bb, rho, math);
            double mx = fmax(la, lo);
            margin_out[r] = mx > 0.0 ? fabs(la - lo) / mx : 0.0;
        }
        i = j;
    }
    free(head);
    return 0;
}
