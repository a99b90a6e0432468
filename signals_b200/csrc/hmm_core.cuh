// hmm_core.cuh -- per-sequence arithmetic of the HSCN / ASCN HMM, written once as
// host+device inline functions.  The CUDA kernels (hmm_kernels.cu) call these per lane;
// tests/hostemu compiles the very same functions with g++ so the control flow can be
// checked against the oracle on a machine without a GPU.  There is no CPU execution
// path in the product: nothing in signals_b200/ calls these functions on the host.
//
// Reference behaviour implemented here (all citations into /root/reference):
//   emission probabilities and masks      R/callHSCN.R:48-80 (HaplotypeHMM), R/callASCN.R:50-81 (alleleHMM)
//   transition matrix                     R/callHSCN.R:82-97, R/callASCN.R:83-96
//   forward max-plus recursion + decode   src/code.cpp:20-42 (viterbi)
//   logspace_add                          src/code.cpp:5-8
// Third-party densities (stats::dbinom, VGAM::dbetabinom) are restated in lgamma form; the
// exact operation order below is the contract shared with oracle/oracle.c (det flavour).
#pragma once
#include "sgb_math.h"

#define SGB_MODEL_HSCN 0
#define SGB_MODEL_ASCN 1
#define SGB_LIK_BINOMIAL 0
#define SGB_LIK_BETABINOMIAL 1
#define SGB_DECODE_REFERENCE 0
#define SGB_DECODE_CHAINED 1

#define SGB_MAX_STATES 32      // S = maxCN+1 supported by the fused kernels
#define SGB_CN_TABLE 16        // total-CN values 0..15 have their per-state constants tabulated
#define SGB_COUNT_BITS 28      // read counts must be < 2^28 (packed record n:28 | x:28 | cn:8)
#define SGB_LGK 10             // beta-binomial: lgamma(shape + k) is tabulated per (total CN, state) for k < SGB_LGK

// kind of a (total CN, state) pair
#define SGB_KIND_NORMAL 0      // 0 < p < 1, use the density formula
#define SGB_KIND_NAN 1         // p outside [0,1] or NaN -> dbinom returns NaN -> -Inf after the NA mask
#define SGB_KIND_P0 2          // p == 0 : point mass at k == 0
#define SGB_KIND_P1 3          // p == 1 : point mass at k == n
#define SGB_KIND_BB_AS_BINOM 4 // beta-binomial with both shapes > exp(20): VGAM falls back to dbinom

struct SgbStateConst {
    double lp;    // log(p)
    double lq;    // log1p(-p)
    double a;     // beta-binomial shape1 = p (1-rho)/rho
    double b;     // beta-binomial shape2 = (1-p)(1-rho)/rho
    double l0;    // lgamma(a) + lgamma(b) - lgamma(a+b)
    int kind;
};

// scalar parameters derived once per call (device set-up kernel / oracle computes the same values)
struct SgbScalars {
    double log_eps;     // log(eps)           (used when eps > 0)
    double log_abias;   // log(Abias)         (used when Abias > 0, HSCN only)
    double tr_self;     // log transition, from == to
    double tr_off;      // log transition, from != to
    double init;        // -log(S)            (src/code.cpp:20)
    double bb_s;        // (1-rho)/rho
    int use_eps;
    int use_abias;
    int fast_ok;        // the lean emission path is valid (no Abias, eps > 0 with log(eps) finite: the default set-up)
    int diag_dominant;  // tr_self >= tr_off: the O(S) recursion needs only the column maximum
};

// total CN as the model sees it: HaplotypeHMM remaps 0 -> 1 (R/callHSCN.R:49), alleleHMM does not
SGB_HD int sgb_cprime(int model, int c) { return (model == SGB_MODEL_HSCN && c == 0) ? 1 : c; }

// number of leading states that are not masked to -Inf:
//   HSCN  l[m > c'] = -Inf      (R/callHSCN.R:80)  -> states 0..min(c',S-1)
//   ASCN  l[m > c/2] = -Inf     (R/callASCN.R:81)  -> states 0..min(floor(c/2),S-1)
SGB_HD int sgb_live_states(int model, int cprime, int S)
{
    int top = (model == SGB_MODEL_HSCN) ? cprime : (cprime >> 1);
    return (top < S - 1 ? top : S - 1) + 1;
}

// per-(total CN, state) constants; p follows R/callHSCN.R:53-59 / R/callASCN.R:53-59
SGB_HD SgbStateConst sgb_state_const(int cprime, int m, double loherror, int lik, double rho, double bb_s,
                                     const double* __restrict__ LT)
{
    SgbStateConst sc;
    double p = sgb_div((double)m, (double)cprime);          // 0/0 = NaN, m/0 = Inf (ASCN, CN 0)
    if (2 * m < cprime) p = sgb_add(p, loherror);
    if (2 * m > cprime) p = sgb_sub(p, loherror);
    sc.lp = 0.0; sc.lq = 0.0; sc.a = 0.0; sc.b = 0.0; sc.l0 = 0.0;
    if (p != p || p < 0.0 || p > 1.0) { sc.kind = SGB_KIND_NAN; return sc; }
    if (p == 0.0) { sc.kind = SGB_KIND_P0; return sc; }
    if (p == 1.0) { sc.kind = SGB_KIND_P1; return sc; }
    sc.kind = SGB_KIND_NORMAL;
    sc.lp = sgb_log(p, LT);
    sc.lq = sgb_log1p(-p, LT);
    if (lik == SGB_LIK_BETABINOMIAL && rho > 0.0) {
        sc.a = sgb_mul(p, bb_s);
        sc.b = sgb_mul(sgb_sub(1.0, p), bb_s);
        const double big = 485165195.40979028;              // exp(20), VGAM's Inf.shape
        if (sc.a > big && sc.b > big) sc.kind = SGB_KIND_BB_AS_BINOM;
        else sc.l0 = sgb_sub(sgb_add(sgb_lgamma(sc.a, LT), sgb_lgamma(sc.b, LT)), sgb_lgamma(sgb_add(sc.a, sc.b), LT));
    }
    return sc;
}

SGB_HD SgbScalars sgb_scalars(int S, double eps, double Abias, double stp, double rho, const double* __restrict__ LT)
{
    SgbScalars s;
    s.use_eps = eps > 0.0;
    s.use_abias = Abias > 0.0;
    s.log_eps = s.use_eps ? sgb_log(eps, LT) : 0.0;
    s.log_abias = s.use_abias ? sgb_log(Abias, LT) : 0.0;
    if (stp == 0.0) {
        s.tr_self = s.tr_off = sgb_log(sgb_div(1.0, (double)S), LT);
    } else {
        s.tr_self = sgb_log(stp, LT);
        s.tr_off = sgb_log(sgb_div(sgb_sub(1.0, stp), (double)(S - 1)), LT);
    }
    if (S == 1) s.tr_off = -sgb_inf();                       // a 1 x 1 transition matrix has no off-diagonal entry
    s.init = -sgb_log((double)S, LT);
    s.bb_s = (rho > 0.0) ? sgb_div(sgb_sub(1.0, rho), rho) : 0.0;
    s.fast_ok = !s.use_abias && s.use_eps && s.log_eps == s.log_eps && sgb_fabs(s.log_eps) != sgb_inf();
    s.diag_dominant = (s.tr_self >= s.tr_off) ? 1 : 0;
    return s;
}

// quantities shared by every state of one bin
struct SgbBin {
    int n, x, cprime;
    double kd, nkd;     // (double)x, (double)(n-x)
    double lc_x;        // log choose(n, x)      = (LF(n) - LF(x)) - LF(n-x)
    double lc_y;        // log choose(n, n-x)    = (LF(n) - LF(n-x)) - LF(x)   (ASCN second density)
    double lg_sn;       // lgamma(s + n)         (beta-binomial)
    int in_support;     // 0 <= x <= n
};

SGB_HD SgbBin sgb_bin_prepare(int model, int lik_bb, int n, int x, int c, double bb_s,
                              const double* __restrict__ LF, const double* __restrict__ LT)
{
    SgbBin b;
    b.n = n; b.x = x; b.cprime = sgb_cprime(model, c);
    b.in_support = (x >= 0) && (x <= n);
    b.kd = (double)x; b.nkd = (double)(n - x);
    b.lc_x = 0.0; b.lc_y = 0.0; b.lg_sn = 0.0;
    if (b.in_support) {
        double fn = sgb_logfact(n, LF, LT), fx = sgb_logfact(x, LF, LT), fy = sgb_logfact(n - x, LF, LT);
        b.lc_x = sgb_sub(sgb_sub(fn, fx), fy);
        b.lc_y = sgb_sub(sgb_sub(fn, fy), fx);
        if (lik_bb) b.lg_sn = sgb_lgamma(sgb_add(bb_s, (double)n), LT);
    }
    return b;
}

// log density of k successes out of k + nk trials for one state (dbinom / dbetabinom restated)
SGB_HD double sgb_density(const SgbStateConst& sc, int lik_bb, double kd, double nkd, double lc, double lg_sn,
                          int in_support, const double* __restrict__ LT)
{
    if (sc.kind == SGB_KIND_NAN) return sgb_nan();
    if (!in_support) return -sgb_inf();
    if (sc.kind == SGB_KIND_P0) return (kd == 0.0) ? 0.0 : -sgb_inf();
    if (sc.kind == SGB_KIND_P1) return (nkd == 0.0) ? 0.0 : -sgb_inf();
    if (!lik_bb || sc.kind == SGB_KIND_BB_AS_BINOM) return sgb_fma(nkd, sc.lq, sgb_fma(kd, sc.lp, lc));
    double t1 = sgb_add(sgb_lgamma(sgb_add(sc.a, kd), LT), sgb_lgamma(sgb_add(sc.b, nkd), LT));
    t1 = sgb_sub(t1, lg_sn);
    return sgb_add(lc, sgb_sub(t1, sc.l0));
}

// emission log-likelihood l[bin, m] of a LIVE state m (masked states are -Inf and never reach here)
SGB_HD double sgb_emission(int model, int lik_bb, const SgbBin& b, int m, const SgbStateConst& sc,
                           const SgbScalars& sp, const double* __restrict__ LT, const double* __restrict__ ET)
{
    double l = sgb_density(sc, lik_bb, b.kd, b.nkd, b.lc_x, b.lg_sn, b.in_support, LT);
    if (model == SGB_MODEL_ASCN) {
        double l2 = sgb_density(sc, lik_bb, b.nkd, b.kd, b.lc_y, b.lg_sn, b.in_support, LT);
        l = sgb_logspace_add(l, l2, LT, ET);                                  // R/callASCN.R:64,71
    }
    if (sp.use_eps) l = sgb_logspace_add(l, sp.log_eps, LT, ET);              // R/callHSCN.R:67-71
    if (model == SGB_MODEL_HSCN && sp.use_abias && 2 * m < b.cprime)
        l = sgb_logspace_add(l, sp.log_abias, LT, ET);                        // R/callHSCN.R:72-78
    if (l != l) l = -sgb_inf();                                               // l[is.na(l)] <- log(0)
    return l;
}

// lgamma(shape + k) on the lean beta-binomial path.  Counts below SGB_LGK come from the per-call table (a cache
// of sgb_lgamma_pos(shape + k), filled by the same function); larger ones have shape + k >= 10, where the
// Stirling part alone gives the same bits.  Branch-free: both are evaluated and one is selected.
SGB_HD double sgb_lg_shape(double shape, double kd, const double* __restrict__ lgk, const double* __restrict__ LT)
{
    const int small = kd < (double)SGB_LGK;
    const double zs = small ? 16.0 : sgb_add(shape, kd);         // any harmless argument for the discarded lane
    const double st = sgb_lgamma_ge10(zs, LT);
    const double tb = lgk[small ? (int)kd : 0];
    return small ? tb : st;
}

// Lean emission for the common case: tabulated total CN, 0 < p < 1 (kind NORMAL), 0 <= x <= n, no Abias,
// finite log(eps).  Same operations in the same order as sgb_density / sgb_emission, minus the branches
// that cannot be taken under these conditions -- so the bits are identical (tests/test_hostemu.py).
template <int MODEL, int LIKBB>
SGB_HD double sgb_emission_fast(double kd, double nkd, double lc_x, double lc_y, double lg_sn, double lp, double lq,
                                double a, double b, double l0, const double* __restrict__ lgk, const SgbScalars& sp,
                                const double* __restrict__ LT, const double* __restrict__ ET)
{
    double l;
    if (!LIKBB) {
        l = sgb_fma(nkd, lq, sgb_fma(kd, lp, lc_x));
    } else {
        double t1 = sgb_add(sgb_lg_shape(a, kd, lgk, LT), sgb_lg_shape(b, nkd, lgk + SGB_LGK, LT));
        t1 = sgb_sub(t1, lg_sn);
        l = sgb_add(lc_x, sgb_sub(t1, l0));
    }
    if (MODEL == SGB_MODEL_ASCN) {
        double l2;
        if (!LIKBB) {
            l2 = sgb_fma(kd, lq, sgb_fma(nkd, lp, lc_y));
        } else {
            double t1 = sgb_add(sgb_lg_shape(a, nkd, lgk, LT), sgb_lg_shape(b, kd, lgk + SGB_LGK, LT));
            t1 = sgb_sub(t1, lg_sn);
            l2 = sgb_add(lc_y, sgb_sub(t1, l0));
        }
        l = sgb_logspace_add_lean(l, l2, LT, ET);
    }
    return sgb_logspace_add_lean(l, sp.log_eps, LT, ET);        // sp.fast_ok implies sp.use_eps
}

// Two lean emissions at once: identical arithmetic per item (so identical bits), but written so that both
// dependency chains sit in the same basic blocks and the scheduler can interleave them.
template <int MODEL, int LIKBB>
SGB_HD void sgb_emission_fast2(double kd0, double nk0, double lx0, double ly0, double gs0, double lp0, double lq0,
                               double a0, double b0, double z0, double kd1, double nk1, double lx1, double ly1,
                               double gs1, double lp1, double lq1, double a1, double b1, double z1,
                               const double* __restrict__ g0, const double* __restrict__ g1, const SgbScalars& sp, const double* __restrict__ LT, const double* __restrict__ ET,
                               double& r0, double& r1)
{
    double l0, l1;
    if (!LIKBB) {
        l0 = sgb_fma(nk0, lq0, sgb_fma(kd0, lp0, lx0));
        l1 = sgb_fma(nk1, lq1, sgb_fma(kd1, lp1, lx1));
    } else {
        double t0 = sgb_add(sgb_lg_shape(a0, kd0, g0, LT), sgb_lg_shape(b0, nk0, g0 + SGB_LGK, LT));
        double t1 = sgb_add(sgb_lg_shape(a1, kd1, g1, LT), sgb_lg_shape(b1, nk1, g1 + SGB_LGK, LT));
        l0 = sgb_add(lx0, sgb_sub(sgb_sub(t0, gs0), z0));
        l1 = sgb_add(lx1, sgb_sub(sgb_sub(t1, gs1), z1));
    }
    if (MODEL == SGB_MODEL_ASCN) {
        double m0, m1;
        if (!LIKBB) {
            m0 = sgb_fma(kd0, lq0, sgb_fma(nk0, lp0, ly0));
            m1 = sgb_fma(kd1, lq1, sgb_fma(nk1, lp1, ly1));
        } else {
            double t0 = sgb_add(sgb_lg_shape(a0, nk0, g0, LT), sgb_lg_shape(b0, kd0, g0 + SGB_LGK, LT));
            double t1 = sgb_add(sgb_lg_shape(a1, nk1, g1, LT), sgb_lg_shape(b1, kd1, g1 + SGB_LGK, LT));
            m0 = sgb_add(ly0, sgb_sub(sgb_sub(t0, gs0), z0));
            m1 = sgb_add(ly1, sgb_sub(sgb_sub(t1, gs1), z1));
        }
        l0 = sgb_logspace_add_lean(l0, m0, LT, ET);
        l1 = sgb_logspace_add_lean(l1, m1, LT, ET);
    }
    const double le = sp.log_eps;                              // sp.fast_ok implies sp.use_eps
    r0 = sgb_logspace_add_lean(l0, le, LT, ET);
    r1 = sgb_logspace_add_lean(l1, le, LT, ET);
}

#if defined(__CUDA_ARCH__)
#define SGB_NOINLINE __device__ __noinline__
#else
#define SGB_NOINLINE static __attribute__((noinline))
#endif

// Tables of the per-(total CN, state) constants for total CN < SGB_CN_TABLE, row pitch S
// (shared memory in the kernel, plain arrays in the host emulation), plus what the
// untabulated path needs to recompute an entry on the fly.
struct SgbTables {
    const double *LT, *ET, *LP, *LQ, *A, *B, *L0;
    int pstride;          // element stride of LP / LQ: 1 = two plain arrays, 2 = interleaved {lp, lq} pairs (LQ = LP + 1)
    const double* LGK;    // [entry][2][SGB_LGK]: lgamma(a + k), lgamma(b + k) for k < SGB_LGK (beta-binomial; same bits as sgb_lgamma_pos)
    const int* KIND;
    int S, lik;
    double loherror, rho;
};

// rare cases: untabulated total CN (>= 16), p at or outside {0,1}, x outside 0..n, Abias, odd eps
template <int MODEL, int LIKBB>
SGB_NOINLINE double sgb_emit_item_slow(int cprime, int in_support, double kd, double nkd, double lc_x, double lc_y,
                                       double lg_sn, int m, const SgbTables& T, const SgbScalars& sp)
{
    SgbStateConst sc;
    if (cprime < SGB_CN_TABLE) {
        const int ti = cprime * T.S + m;
        sc.lp = T.LP[ti * T.pstride]; sc.lq = T.LQ[ti * T.pstride]; sc.kind = T.KIND[ti];
        sc.a = LIKBB ? T.A[ti] : 0.0; sc.b = LIKBB ? T.B[ti] : 0.0; sc.l0 = LIKBB ? T.L0[ti] : 0.0;
    } else {
        sc = sgb_state_const(cprime, m, T.loherror, T.lik, T.rho, sp.bb_s, T.LT);
    }
    SgbBin b;
    b.n = 0; b.x = 0; b.cprime = cprime; b.kd = kd; b.nkd = nkd; b.lc_x = lc_x; b.lc_y = lc_y; b.lg_sn = lg_sn;
    b.in_support = in_support;
    return sgb_emission(MODEL, LIKBB, b, m, sc, sp, T.LT, T.ET);
}

// emission of live state m of one bin: lean path when its preconditions hold, else the general one
template <int MODEL, int LIKBB>
SGB_HD double sgb_emit_item(int cprime, int in_support, double kd, double nkd, double lc_x, double lc_y,
                            double lg_sn, int m, const SgbTables& T, const SgbScalars& sp)
{
    const int ti = (cprime < SGB_CN_TABLE) ? cprime * T.S + m : 0;
    const bool fast = (cprime < SGB_CN_TABLE) && in_support && sp.fast_ok && (T.KIND[ti] == SGB_KIND_NORMAL);
    if (fast)
        return sgb_emission_fast<MODEL, LIKBB>(kd, nkd, lc_x, lc_y, lg_sn, T.LP[ti * T.pstride], T.LQ[ti * T.pstride], LIKBB ? T.A[ti] : 0.0,
                                               LIKBB ? T.B[ti] : 0.0, LIKBB ? T.L0[ti] : 0.0,
                                               LIKBB ? T.LGK + (long)ti * 2 * SGB_LGK : T.LT, sp, T.LT, T.ET);
    return sgb_emit_item_slow<MODEL, LIKBB>(cprime, in_support, kd, nkd, lc_x, lc_y, lg_sn, m, T, sp);
}

// ------------------------------------------------------------------ forward recursion
// State of one sequence between bins.  Because the reference's transition matrix has one
// value on the diagonal (tr_self) and one off it (tr_off), and rounding is monotone,
//     max_k fl(fl(T1[k] + tr[k,i]) + e_i) = max( fl(fl(M_i + tr_off) + e_i), fl(fl(T1[i] + tr_self) + e_i) )
// with M_i = max_{k != i} T1[k], which needs only the largest value m1 (first index i1) and the
// largest value m2 among the other states.  That gives every T1[i,t] bit-exactly in O(S).
// The reference's decode (src/code.cpp:35-42) only ever reads T2 at the column-wise best
// state, so the first-index argmax over predecessors is evaluated for that one state alone,
// by the reference's own expression (T1[k] + tr[k,i]) + e_i and strict '>' (first maximum).
template <int SMAX>
struct SgbSeqState {
    double T1[SMAX];
    double m1, m2;
    int i1;
};

template <int SMAX>
SGB_HD void sgb_top2(SgbSeqState<SMAX>& st, int S)
{
    double m1 = st.T1[0], m2 = -sgb_inf();
    int i1 = 0;
#pragma unroll
    for (int i = 1; i < SMAX; i++) {
        if (i < S) {
            double v = st.T1[i];
            if (v > m1) { m2 = m1; m1 = v; i1 = i; }
            else if (v > m2) m2 = v;
        }
    }
    st.m1 = m1; st.m2 = m2; st.i1 = i1;
}

// first bin of a sequence:  T1[,0] = initial + emission   (src/code.cpp:25)
template <int SMAX>
SGB_HD void sgb_seq_init(SgbSeqState<SMAX>& st, const double* e, int S, double init)
{
#pragma unroll
    for (int i = 0; i < SMAX; i++) st.T1[i] = (i < S) ? sgb_add(init, e[i]) : -sgb_inf();
    sgb_top2<SMAX>(st, S);
}

SGB_HD int sgb_ffs32(unsigned int m)
{
#if defined(__CUDA_ARCH__)
    return __ffs((int)m);
#else
    return __builtin_ffs((int)m);
#endif
}

// General variant (tr_self < tr_off, i.e. selftransitionprob below 1/S): needs the two largest values.
template <int SMAX>
SGB_HD int sgb_seq_step_general(SgbSeqState<SMAX>& st, const double* e, int S, double tr_self, double tr_off)
{
    double Tn[SMAX];
    const double u1 = sgb_add(st.m1, tr_off), u2 = sgb_add(st.m2, tr_off);
    double m1 = -sgb_inf(), m2 = -sgb_inf();
    int i1 = 0;
#pragma unroll
    for (int i = 0; i < SMAX; i++) {
        if (i < S) {
            double voff = sgb_add((i == st.i1) ? u2 : u1, e[i]);
            double vdia = sgb_add(sgb_add(st.T1[i], tr_self), e[i]);
            double v = (S > 1 && voff > vdia) ? voff : vdia;
            Tn[i] = v;
            if (i == 0) { m1 = v; }
            else if (v > m1) { m2 = m1; m1 = v; i1 = i; }
            else if (v > m2) m2 = v;
        } else {
            Tn[i] = -sgb_inf();
        }
    }
    double ej = e[0];
#pragma unroll
    for (int i = 1; i < SMAX; i++) if (i == i1) ej = e[i];
    int back = 0;
#pragma unroll
    for (int k = SMAX - 1; k >= 0; k--) {
        if (k < S) {
            double val = sgb_add(sgb_add(st.T1[k], (k == i1) ? tr_self : tr_off), ej);
            if (val == m1) back = k;
        }
    }
#pragma unroll
    for (int i = 0; i < SMAX; i++) st.T1[i] = Tn[i];
    st.m1 = m1; st.m2 = m2; st.i1 = i1;
    return back;
}

// One bin (j >= 1).  Returns the state the reference emits for bin j-1:
// T2[argmax_i T1[i,j], j]  (src/code.cpp:39-42).  e[i] must be -Inf for masked states.
//
// Diagonal-dominant transition (tr_self >= tr_off, every realistic selftransitionprob): for the state that
// holds the column maximum m1 the diagonal candidate fl(fl(m1+tr_self)+e) already dominates any
// off-diagonal one, so fl(fl(m1+tr_off)+e) can stand in for "max over k != i" for EVERY i without
// changing a bit.  Branch-free: 3 adds + 1 compare per state, one max chain, one equality mask.
// The predecessor scan runs only when the winner was not reached strictly through its own diagonal.
// States i >= S (when S < SMAX) are carried as permanently masked states: their e[i] and T1[i] are -Inf,
// they can never win a strict '>' or an equality against a finite maximum, and when everything is -Inf the
// first-index rule picks state 0 -- so the loops below run over all SMAX slots without guards.
// ecol/estride: the same emissions addressable by a run-time index (shared-memory column in the kernel).
template <int SMAX>
SGB_HD int sgb_seq_step(SgbSeqState<SMAX>& st, const double* e, const double* ecol, int estride, int S,
                        double tr_self, double tr_off, int diag_dominant)
{
    if (!diag_dominant) return sgb_seq_step_general<SMAX>(st, e, S, tr_self, tr_off);
    double Tn[SMAX];
    const double u1 = sgb_add(st.m1, tr_off);
#pragma unroll
    for (int i = 0; i < SMAX; i++) {
        const double voff = sgb_add(u1, e[i]);
        const double vdia = sgb_add(sgb_add(st.T1[i], tr_self), e[i]);
        Tn[i] = (vdia > voff) ? vdia : voff;
    }
    // first maximum by a tournament (depth log2 SMAX instead of a chain of SMAX-1 dependent compares):
    // the left operand always holds the lower indices, so '>' (strict) keeps the lowest index on ties
    double tv[SMAX];
    int ti[SMAX];
#pragma unroll
    for (int i = 0; i < SMAX; i++) { tv[i] = Tn[i]; ti[i] = i; }
#pragma unroll
    for (int stride = 1; stride < SMAX; stride *= 2) {
#pragma unroll
        for (int i = 0; i + stride < SMAX; i += 2 * stride) {
            const bool gt = tv[i + stride] > tv[i];
            tv[i] = gt ? tv[i + stride] : tv[i];
            ti[i] = gt ? ti[i + stride] : ti[i];
        }
    }
    const double m1 = tv[0];
    const int i1 = ti[0];
    int back = i1;
    const double ej = ecol[i1 * estride];
    if (m1 == sgb_add(u1, ej)) {
        // The winner's value equals its off-diagonal candidate (its diagonal one is not strictly larger), so a
        // predecessor other than i1 may hold the first maximum: find the first k with
        // (T1[k] + tr[k,i1]) + e[i1] == m1 -- the reference's own expression (src/code.cpp:29-31).
        double tj = st.T1[0];
#pragma unroll
        for (int i = 1; i < SMAX; i++) if (i == i1) tj = st.T1[i];
        unsigned int pm = 0;
#pragma unroll
        for (int k = 0; k < SMAX; k++) {
            const double val = sgb_add(sgb_add(st.T1[k], tr_off), ej);
            pm |= (val == m1) ? (1u << k) : 0u;
        }
        const double wd = sgb_add(sgb_add(tj, tr_self), ej);
        pm = (pm & ~(1u << i1)) | ((wd == m1) ? (1u << i1) : 0u);
        back = sgb_ffs32(pm) - 1;
    }
#pragma unroll
    for (int i = 0; i < SMAX; i++) st.T1[i] = Tn[i];
    st.m1 = m1; st.i1 = i1;
    return back;
}

// The same step with the T1 column kept in (shared) memory instead of registers: T1[i*ts], E[i*es].
// Frees 2*SMAX registers per thread for occupancy; arithmetic and results are identical to sgb_seq_step.
template <int SMAX>
SGB_HD void sgb_seq_init_mem(double* T1, int ts, const double* E, int es, int S, double init, double& m1)
{
    double best = -sgb_inf();
#pragma unroll
    for (int i = 0; i < SMAX; i++) {
        const double v = (i < S) ? sgb_add(init, E[i * es]) : -sgb_inf();
        T1[i * ts] = v;
        best = (v > best) ? v : best;
    }
    m1 = best;
}

// NL (compile time) bounds the states that can be alive: states >= NL are -Inf both in the previous column and
// in this bin's emissions, so they are neither loaded, computed nor stored (they stay -Inf in memory).
template <int NL>
SGB_HD int sgb_seq_step_mem_nl(double* T1, int ts, const double* E, int es, double& m1io, double tr_self, double tr_off)
{
    double Tn[NL];
    const double u1 = sgb_add(m1io, tr_off);
#pragma unroll
    for (int i = 0; i < NL; i++) {
        // max(fl(u1 + e), fl(fl(T1[i] + tr_self) + e)) == fl(max(u1, fl(T1[i] + tr_self)) + e): rounding is monotone,
        // so taking the maximum before the common addend gives the same bits with one addition less
        const double td = sgb_add(T1[i * ts], tr_self);
        Tn[i] = sgb_add((td > u1) ? td : u1, E[i * es]);
    }
    // first maximum by a tournament: the left operand always holds the lower indices, so a strict '>' keeps
    // the lowest index on ties (what which_max does)
    double tv[NL];
    int ti[NL];
#pragma unroll
    for (int i = 0; i < NL; i++) { tv[i] = Tn[i]; ti[i] = i; }
#pragma unroll
    for (int stride = 1; stride < NL; stride *= 2) {
#pragma unroll
        for (int i = 0; i + stride < NL; i += 2 * stride) {
            const bool gt = tv[i + stride] > tv[i];
            tv[i] = gt ? tv[i + stride] : tv[i];
            ti[i] = gt ? ti[i + stride] : ti[i];
        }
    }
    const double m1 = tv[0];
    const int i1 = ti[0];
    int back = i1;
    const double ej = E[i1 * es];
    if (m1 == sgb_add(u1, ej)) {
        // The winner's value equals its off-diagonal candidate (its diagonal one is not strictly larger), so a
        // predecessor other than i1 may hold the first maximum: find the first k with
        // (T1[k] + tr[k,i1]) + e[i1] == m1 -- the reference's own expression (src/code.cpp:29-31).
        // (everything -Inf: every predecessor "matches" and the first one is state 0, as in the full scan)
        unsigned int pm = 0;
#pragma unroll
        for (int k = 0; k < NL; k++) {
            const double val = sgb_add(sgb_add(T1[k * ts], tr_off), ej);
            pm |= (val == m1) ? (1u << k) : 0u;
        }
        const double wd = sgb_add(sgb_add(T1[i1 * ts], tr_self), ej);
        pm = (pm & ~(1u << i1)) | ((wd == m1) ? (1u << i1) : 0u);
        back = sgb_ffs32(pm) - 1;
    }
#pragma unroll
    for (int i = 0; i < NL; i++) T1[i * ts] = Tn[i];
    m1io = m1;
    return back;
}

// run-time nlive (uniform across a warp) -> the smallest compiled bound that covers it
template <int SMAX>
SGB_HD int sgb_seq_step_mem(double* T1, int ts, const double* E, int es, double& m1io, double tr_self, double tr_off,
                            int nlive)
{
    if (SMAX <= 12) {
        if (nlive <= 2) return sgb_seq_step_mem_nl<(SMAX < 2 ? SMAX : 2)>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 4) return sgb_seq_step_mem_nl<(SMAX < 4 ? SMAX : 4)>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 6) return sgb_seq_step_mem_nl<(SMAX < 6 ? SMAX : 6)>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 8) return sgb_seq_step_mem_nl<(SMAX < 8 ? SMAX : 8)>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 10) return sgb_seq_step_mem_nl<(SMAX < 10 ? SMAX : 10)>(T1, ts, E, es, m1io, tr_self, tr_off);
        return sgb_seq_step_mem_nl<SMAX>(T1, ts, E, es, m1io, tr_self, tr_off);
    } else {
        if (nlive <= 8) return sgb_seq_step_mem_nl<8>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 16) return sgb_seq_step_mem_nl<(SMAX < 16 ? SMAX : 16)>(T1, ts, E, es, m1io, tr_self, tr_off);
        if (nlive <= 24) return sgb_seq_step_mem_nl<(SMAX < 24 ? SMAX : 24)>(T1, ts, E, es, m1io, tr_self, tr_off);
        return sgb_seq_step_mem_nl<SMAX>(T1, ts, E, es, m1io, tr_self, tr_off);
    }
}

// Chained (textbook) mode: backpointers of ALL states for bin j, bp[i] = first argmax_k.
// nl: states >= nl are dead in this bin AND the previous one (e[i] = T1[i] = -Inf there), so their column entry is -Inf with
// backpointer 0 (every candidate is -Inf: the first one wins) and they never beat a candidate of a live state -- the S x S
// candidate loop shrinks to nl x nl (nl = warp-wide maximum of the live-state counts of the two bins, 5-6 of 12 on DLP data).
template <int SMAX>
SGB_HD void sgb_seq_step_chained(SgbSeqState<SMAX>& st, const double* e, int S, double tr_self, double tr_off,
                                 unsigned char* bp, int nl = SMAX)
{
    double Tn[SMAX];
    const int Sl = S < nl ? S : nl;
#pragma unroll
    for (int i = 0; i < SMAX; i++) {
        Tn[i] = -sgb_inf();
        if (i < S) bp[i] = 0;
        if (i < Sl) {
            double best = sgb_add(sgb_add(st.T1[0], (0 == i) ? tr_self : tr_off), e[i]);
            int arg = 0;
#pragma unroll
            for (int k = 1; k < SMAX; k++) {
                if (k < Sl) {
                    double val = sgb_add(sgb_add(st.T1[k], (k == i) ? tr_self : tr_off), e[i]);
                    if (val > best) { best = val; arg = k; }
                }
            }
            Tn[i] = best;
            bp[i] = (unsigned char)arg;
        }
    }
#pragma unroll
    for (int i = 0; i < SMAX; i++) st.T1[i] = Tn[i];
    sgb_top2<SMAX>(st, S);
}

// ------------------------------------------------------------------ register-resident recursion (split schedule)
// The split schedule (k_emit + k_recur) hands every lane the emissions of its sequence's bin as NL doubles, NL a
// multiple of 4 that bounds the live states of the WHOLE sequence (states >= NL are -Inf in every bin and therefore
// in every forward column).  Same arithmetic, in the same order, as sgb_seq_init_mem / sgb_seq_step_mem_nl above, with
// the forward column in registers.
template <int NL>
SGB_HD void sgb_seq_init_r(double (&T1)[NL], const double (&e)[NL], double init, double& m1)
{
    double best = -sgb_inf();
#pragma unroll
    for (int i = 0; i < NL; i++) {
        const double v = sgb_add(init, e[i]);        // e[i] = -Inf for masked states and for states >= S
        T1[i] = v;
        best = (v > best) ? v : best;
    }
    m1 = best;
}

template <int NL>
SGB_HD int sgb_seq_step_r(double (&T1)[NL], const double (&e)[NL], double& m1io, double tr_self, double tr_off)
{
    double Tn[NL];
    const double u1 = sgb_add(m1io, tr_off);
#pragma unroll
    for (int i = 0; i < NL; i++) {
        const double td = sgb_add(T1[i], tr_self);
        Tn[i] = sgb_add((td > u1) ? td : u1, e[i]);
    }
    double tv[NL];
    int ti[NL];
#pragma unroll
    for (int i = 0; i < NL; i++) { tv[i] = Tn[i]; ti[i] = i; }
#pragma unroll
    for (int stride = 1; stride < NL; stride *= 2) {
#pragma unroll
        for (int i = 0; i + stride < NL; i += 2 * stride) {
            const bool gt = tv[i + stride] > tv[i];
            tv[i] = gt ? tv[i + stride] : tv[i];
            ti[i] = gt ? ti[i + stride] : ti[i];
        }
    }
    const double m1 = tv[0];
    const int i1 = ti[0];
    int back = i1;
    double ej = e[0];
#pragma unroll
    for (int i = 1; i < NL; i++) ej = (i == i1) ? e[i] : ej;
    if (m1 == sgb_add(u1, ej)) {
        // see sgb_seq_step_mem_nl: first k with (T1[k] + tr[k,i1]) + e[i1] == m1 (src/code.cpp:29-31)
        double tj = T1[0];
#pragma unroll
        for (int i = 1; i < NL; i++) tj = (i == i1) ? T1[i] : tj;
        unsigned int pm = 0;
#pragma unroll
        for (int k = 0; k < NL; k++) {
            const double val = sgb_add(sgb_add(T1[k], tr_off), ej);
            pm |= (val == m1) ? (1u << k) : 0u;
        }
        const double wd = sgb_add(sgb_add(tj, tr_self), ej);
        pm = (pm & ~(1u << i1)) | ((wd == m1) ? (1u << i1) : 0u);
        back = sgb_ffs32(pm) - 1;
    }
#pragma unroll
    for (int i = 0; i < NL; i++) T1[i] = Tn[i];
    m1io = m1;
    return back;
}

// The same step without the tie analysis: the forward column advances exactly as in sgb_seq_step_r and the column-best state
// i1 is returned; `tie` tells whether the backpointer of i1 may differ from i1 (the best off-diagonal candidate reaches
// the maximum), in which case the caller has to redo the step with sgb_seq_step_r from the column it started with.
// Branch-free, so that consecutive steps of an unrolled loop form one basic block and overlap.
template <int NL>
SGB_HD int sgb_seq_step_fast(double (&T1)[NL], const double (&e)[NL], double& m1io, double tr_self, double tr_off, bool& tie)
{
    const double u1 = sgb_add(m1io, tr_off);
    double tv[NL];
    int ti[NL];
#pragma unroll
    for (int i = 0; i < NL; i++) {
        const double td = sgb_add(T1[i], tr_self);
        T1[i] = sgb_add((td > u1) ? td : u1, e[i]);
        tv[i] = T1[i]; ti[i] = i;
    }
#pragma unroll
    for (int stride = 1; stride < NL; stride *= 2) {
#pragma unroll
        for (int i = 0; i + stride < NL; i += 2 * stride) {
            const bool gt = tv[i + stride] > tv[i];
            tv[i] = gt ? tv[i + stride] : tv[i];
            ti[i] = gt ? ti[i + stride] : ti[i];
        }
    }
    const double m1 = tv[0];
    const int i1 = ti[0];
    double ej = e[0];
#pragma unroll
    for (int i = 1; i < NL; i++) ej = (i == i1) ? e[i] : ej;
    tie = (m1 == sgb_add(u1, ej));
    m1io = m1;
    return i1;
}

// number of 4-state planes a sequence whose largest live-state count is L needs; the recursion kernels are compiled
// for 1, 2, 3, 4, 6 and 8 planes
SGB_HD int sgb_planes_for(int L)
{
    const int p = (L + 3) >> 2;
    return p <= 4 ? p : (p <= 6 ? 6 : 8);
}
