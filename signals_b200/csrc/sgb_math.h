// sgb_math.h -- deterministic fp64 transcendental functions shared by the CUDA
// kernels (device) and by the CPU oracle's "det" flavour (host).
//
// Why this exists: the reference computes its emission log-likelihoods with
// libm / Rmath (exp, log1p in src/code.cpp:5-8; dbinom / lgamma-class code
// behind R/callHSCN.R:61-66 and R/callASCN.R:61-73).  CUDA's libdevice and
// glibc differ in the last ulp, which is enough to flip a first-index argmax
// tie in the Viterbi recursion (src/code.cpp:27-33).  Every function below is
// built ONLY from IEEE-754 correctly rounded primitives (add, sub, mul, fma,
// div), integer bit manipulation and table lookups of constants emitted as hex
// floats, so the host and the device produce bit-identical results and the
// GPU state paths can be compared bit-for-bit with the oracle.
//
// Compile rules:  device code uses the __d*_rn intrinsics, which nvcc never
// contracts; host code MUST be built with -ffp-contract=off (the oracle's
// Makefile does this and orc_selftest() verifies it at run time).
//
// Accuracy (measured against 50-digit mpmath in tests/test_detmath.py):
//   sgb_exp       <= 1 ulp            on [-708, 709]  (0 below, +Inf above)
//   sgb_log       <= 1 ulp            all positive normals (near-1 branch is fdlibm-style)
//   sgb_log_tab   <= 2e-16 absolute near 1, <= 1.5 ulp elsewhere
//   sgb_log1p_01  <= 1.5 ulp          on [0, 1]
//   sgb_lgamma    <= 6e-16*|lgamma| + 2e-14 absolute,  z > 0
#pragma once
#include <stdint.h>
#include <string.h>
#include "sgb_tables.h"

#if defined(__CUDACC__)
#define SGB_HD static __host__ __device__ __forceinline__
#else
#define SGB_HD static inline __attribute__((always_inline))
#endif

#define SGB_SHIFT (0x1.8p52)

// Where entry j of the exp table / pair i of the log table sits relative to the pointer a function is given.  The
// default is the plain packed layout.  A translation unit may define both before including this header to read
// lane-private, bank-conflict-free copies of the tables in shared memory (split_kernels.cu): the VALUES read are the
// same, so the results are the same bits.
#ifndef SGB_ETI
#define SGB_ETI(j) (j)
#endif
#ifndef SGB_LTI
#define SGB_LTI(i) (2 * (i))
#endif

// Scalar constants: host code uses the hex-float literals; device code reads the same values from
// the constant bank (operand-folded into DFMA/DADD instead of two 32-bit immediate moves per use).
#if defined(__CUDACC__)
static __constant__ double sgb_kdev[SGB_NK] = SGB_KTAB_INIT;
#endif
#if defined(__CUDA_ARCH__)
#define SGB_C(name) (sgb_kdev[SGB_KI_##name])
#else
#define SGB_C(name) (SGB_##name)
#endif

// ---------------------------------------------------------------- primitives
SGB_HD uint64_t sgb_d2u(double x)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u; memcpy(&u, &x, 8); return u;
#endif
}
SGB_HD double sgb_u2d(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double x; memcpy(&x, &u, 8); return x;
#endif
}
SGB_HD double sgb_add(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
SGB_HD double sgb_sub(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, -b);
#else
    return a - b;
#endif
}
SGB_HD double sgb_mul(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
}
SGB_HD double sgb_fma(double a, double b, double c)
{
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
SGB_HD double sgb_div(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    return a / b;
#endif
}
SGB_HD double sgb_inf(void) { return sgb_u2d(0x7ff0000000000000ULL); }
SGB_HD double sgb_nan(void) { return sgb_u2d(0x7ff8000000000000ULL); }
SGB_HD double sgb_fabs(double x) { return sgb_u2d(sgb_d2u(x) & 0x7fffffffffffffffULL); }
// C fmax semantics (src/code.cpp:7 uses fmax): a NaN operand yields the other one.
SGB_HD double sgb_fmax(double a, double b)
{
    if (a != a) return b;
    if (b != b) return a;
    return (a > b) ? a : b;
}

// ---------------------------------------------------------------- exp
// exp(x) = 2^q * T[j] * (1 + p(r)),  x = (64 q + j) ln2/64 + r,  |r| <= ln2/128.
// Results below 2^-1022 are flushed to +0 (x <= -708): the only caller that can
// get there is logspace_add, where exp(d) is added to 1 -- irrelevant there.
// core: valid for x in [-708, 709], no special-value handling
SGB_HD double sgb_exp_core(double x, const double* __restrict__ ET)
{
    double t  = sgb_fma(x, SGB_C(INVLN2_N), SGB_SHIFT);
    double kd = sgb_sub(t, SGB_SHIFT);
    int32_t k = (int32_t)(uint32_t)sgb_d2u(t);
    double r  = sgb_fma(kd, -SGB_C(LN2N_HI), x);
    r = sgb_fma(kd, -SGB_C(LN2N_LO), r);
    int32_t j = k & (SGB_NEXP - 1);
    int64_t q = (int64_t)(k >> 6);
    double s = sgb_u2d(sgb_d2u(ET[SGB_ETI(j)]) + ((uint64_t)q << 52));
    // Estrin evaluation (3 dependent levels instead of 6): 1 + C2 r + ... + C6 r^5
    double r2 = sgb_mul(r, r);
    double p01 = sgb_fma(r, SGB_C(EXPC2), 1.0);
    double p23 = sgb_fma(r, SGB_C(EXPC4), SGB_C(EXPC3));
    double p45 = sgb_fma(r, SGB_C(EXPC6), SGB_C(EXPC5));
    double p = sgb_fma(r2, p45, p23);
    p = sgb_fma(r2, p, p01);
    p = sgb_mul(r, p);
    return sgb_fma(s, p, s);
}

SGB_HD double sgb_exp(double x, const double* __restrict__ ET)
{
    if (!(x > -708.0)) return (x != x) ? x : 0.0;
    if (x > 709.0) return sgb_inf();      // (overflow threshold rounded down from 709.78; never reached: callers pass x <= 0)
    return sgb_exp_core(x, ET);
}

// ---------------------------------------------------------------- log
// Table kernel: x = 2^k * w, w in [1,2); interval i = top 7 mantissa bits,
// r = w*invc_i - 1 (one fma, |r| <= 2^-7), log x = k ln2 + logc_i + log1p(r).
// *rcp_seed (optional) receives invc_i * 2^-k ~ 1/x to 8 bits for sgb_rcp().
SGB_HD double sgb_log_core(uint64_t ix, int32_t kadj, const double* __restrict__ LT, double* rcp_seed)
{
    int32_t k = (int32_t)(ix >> 52) - 1023;
    uint64_t m = ix & 0x000fffffffffffffULL;
    int32_t i = (int32_t)(m >> 45);
    double w = sgb_u2d(m | 0x3ff0000000000000ULL);
    double invc = LT[SGB_LTI(i)], logc = LT[SGB_LTI(i) + 1];
    if (rcp_seed) *rcp_seed = sgb_u2d(sgb_d2u(invc) - ((uint64_t)(int64_t)k << 52));
    double r = sgb_fma(w, invc, -1.0);
    double kd = (double)(k + kadj);
    // Estrin evaluation of C2 + C3 r + ... + C9 r^7 (3 dependent levels instead of 7)
    double r2 = sgb_mul(r, r);
    double r4 = sgb_mul(r2, r2);
    double q0 = sgb_fma(r, SGB_C(LOGC3), SGB_C(LOGC2));
    double q1 = sgb_fma(r, SGB_C(LOGC5), SGB_C(LOGC4));
    double q2 = sgb_fma(r, SGB_C(LOGC7), SGB_C(LOGC6));
    double q3 = sgb_fma(r, SGB_C(LOGC9), SGB_C(LOGC8));
    double p = sgb_fma(r4, sgb_fma(r2, q3, q2), sgb_fma(r2, q1, q0));
    double lo = sgb_fma(r2, p, sgb_mul(kd, SGB_C(LN2_LO)));
    double hi = sgb_fma(kd, SGB_C(LN2_HI), logc);
    return sgb_add(hi, sgb_add(r, lo));
}

// log of a positive finite x (normal or subnormal); no special-value handling.
SGB_HD double sgb_log_tab(double x, const double* __restrict__ LT)
{
    uint64_t ix = sgb_d2u(x);
    int32_t kadj = 0;
    if (ix < 0x0010000000000000ULL) { ix = sgb_d2u(sgb_mul(x, 0x1p54)); kadj = -54; }
    return sgb_log_core(ix, kadj, LT, 0);
}

// log(1+f) for |f| <= 0.5 by the fdlibm identity log(1+f) = f - (f^2/2 - s*(f^2/2 + R)),
// s = f/(2+f), R = sum_k 2/(2k+1) s^(2k).  Division + 14 terms: cold path only.
SGB_HD double sgb_log1pf_series(double f)
{
    double s = sgb_div(f, sgb_add(2.0, f));
    double z = sgb_mul(s, s);
    double R = sgb_fma(z, SGB_C(ATHC14), SGB_C(ATHC13));
    R = sgb_fma(z, R, SGB_C(ATHC12));
    R = sgb_fma(z, R, SGB_C(ATHC11));
    R = sgb_fma(z, R, SGB_C(ATHC10));
    R = sgb_fma(z, R, SGB_C(ATHC9));
    R = sgb_fma(z, R, SGB_C(ATHC8));
    R = sgb_fma(z, R, SGB_C(ATHC7));
    R = sgb_fma(z, R, SGB_C(ATHC6));
    R = sgb_fma(z, R, SGB_C(ATHC5));
    R = sgb_fma(z, R, SGB_C(ATHC4));
    R = sgb_fma(z, R, SGB_C(ATHC3));
    R = sgb_fma(z, R, SGB_C(ATHC2));
    R = sgb_fma(z, R, SGB_C(ATHC1));
    R = sgb_mul(z, R);
    double hfsq = sgb_mul(0.5, sgb_mul(f, f));
    return sgb_sub(f, sgb_sub(hfsq, sgb_mul(s, sgb_add(hfsq, R))));
}

// Full log with IEEE special cases and an fdlibm-style branch near 1 (used for
// table set-up: log(p), log(eps), log(transition) -- cold path).
SGB_HD double sgb_log(double x, const double* __restrict__ LT)
{
    if (x != x) return x;
    if (x < 0.0) return sgb_nan();
    if (x == 0.0) return -sgb_inf();
    if (x == sgb_inf()) return x;
    if (x >= 0.75 && x < 1.5) return sgb_log1pf_series(sgb_sub(x, 1.0));   // x-1 exact (Sterbenz)
    return sgb_log_tab(x, LT);
}

// core of log1p(u) for u in [0,1] given w = fl(1+u) in [1,2] (no checks).  w == 2.0 lands on table
// entry 128 (invc = 1/2, logc = ln 2), so the result there is exactly SGB_LN2.
SGB_HD double sgb_log1p01_core(double u, double w, const double* __restrict__ LT)
{
    double c = sgb_sub(u, sgb_sub(w, 1.0));          // exact rounding residual of 1+u
    int32_t i = (int32_t)(((uint32_t)(sgb_d2u(w) >> 32) - 0x3ff00000u) >> 13);     // top 7 mantissa bits (w == 2 -> 128)
    double invc = LT[SGB_LTI(i)], logc = LT[SGB_LTI(i) + 1];
    double r = sgb_fma(w, invc, -1.0);
    // Estrin evaluation of C2 + C3 r + ... + C9 r^7 (3 dependent levels instead of 7)
    double r2 = sgb_mul(r, r);
    double r4 = sgb_mul(r2, r2);
    double q0 = sgb_fma(r, SGB_C(LOGC3), SGB_C(LOGC2));
    double q1 = sgb_fma(r, SGB_C(LOGC5), SGB_C(LOGC4));
    double q2 = sgb_fma(r, SGB_C(LOGC7), SGB_C(LOGC6));
    double q3 = sgb_fma(r, SGB_C(LOGC9), SGB_C(LOGC8));
    double p = sgb_fma(r4, sgb_fma(r2, q3, q2), sgb_fma(r2, q1, q0));
    double lo = sgb_fma(r2, p, sgb_mul(c, invc));
    return sgb_add(logc, sgb_add(r, lo));
}

// log1p(u) for u in [0,1] (hot path: u = exp(-|a-b|) inside logspace_add).
SGB_HD double sgb_log1p_01(double u, const double* __restrict__ LT)
{
    if (!(u <= 1.0)) return (u != u) ? u : sgb_nan();
    return sgb_log1p01_core(u, sgb_add(1.0, u), LT);
}

// General log1p for x > -1 (cold path: log1p(-p) for the per-(CN,state) tables).
SGB_HD double sgb_log1p(double x, const double* __restrict__ LT)
{
    if (x != x) return x;
    if (x < -1.0) return sgb_nan();
    if (x == -1.0) return -sgb_inf();
    if (x == sgb_inf()) return x;
    if (x >= -0.25 && x < 0.5) return sgb_log1pf_series(x);
    double w = sgb_add(1.0, x);
    double c = (x < 1.0) ? sgb_sub(x, sgb_sub(w, 1.0)) : sgb_add(sgb_sub(x, w), 1.0);
    return sgb_add(sgb_log(w, LT), sgb_div(c, w));
}

// ---------------------------------------------------------------- logspace_add
// src/code.cpp:5-8  fmax(a,b) + log1p(exp(-fabs(a-b))), same NaN/Inf behaviour.
SGB_HD double sgb_max(double a, double b)            // operands are never NaN here (fmax would add NaN fix-ups)
{
    return (a > b) ? a : b;
}

// Branch-free body, valid when neither operand is NaN and they are not the same infinity.
// The exponent is clamped at -708 (exp(-708) ~ 3e-308 instead of an underflowing value: it vanishes
// against any max(a,b) that is not itself below 1e-290 in magnitude).
SGB_HD double sgb_logspace_add_lean(double a, double b, const double* __restrict__ LT, const double* __restrict__ ET)
{
    double mx = sgb_max(a, b);
    // -fabs(a-b) with its magnitude clamped near 708, both on the high word only (sign bit set, unsigned minimum
    // with the high word of -708.0): one integer min instead of an FP64 compare-and-select.  A clamped value
    // lands in (-708.0005, -708], where exp is ~3e-308 and vanishes against max(a,b) just as exp(-708) does.
    uint64_t db = sgb_d2u(sgb_sub(a, b));
    uint32_t hi = (uint32_t)(db >> 32) | 0x80000000u;
    hi = hi < 0xC0862000u ? hi : 0xC0862000u;
    double d = sgb_u2d(((uint64_t)hi << 32) | (uint64_t)(uint32_t)db);
    double u = sgb_exp_core(d, ET);
    return sgb_add(mx, sgb_log1p01_core(u, sgb_add(1.0, u), LT));
}

SGB_HD double sgb_logspace_add(double a, double b, const double* __restrict__ LT, const double* __restrict__ ET)
{
    if (a != a || b != b) return sgb_nan();
    if (a == b && sgb_fabs(a) == sgb_inf()) return sgb_nan();               // Inf - Inf
    return sgb_logspace_add_lean(a, b, LT, ET);
}

// ---------------------------------------------------------------- lgamma (z > 0)
// z < 10 is shifted up with the recurrence (P = z(z+1)...(z+k-1)), then
// Stirling with 7 correction terms; 1/y by two cubic Newton steps from the
// table seed (no division instruction on the hot path).
// core: finite z > 0, no special-value handling.  Branch-free on purpose (several calls sit in one basic
// block of the emission kernel and are interleaved by the scheduler): arguments below 10 are shifted up by
// exactly 10 with the recurrence lgamma(z) = lgamma(z+10) - log(z (z+1) ... (z+9)); larger ones use P = 1,
// whose logarithm the table kernel returns as an exact 0.
SGB_HD double sgb_lgamma_pos(double z, const double* __restrict__ LT)
{
    const int sm = z < 10.0;
    double P = z;
    P = sgb_mul(P, sgb_add(z, 1.0));
    P = sgb_mul(P, sgb_add(z, 2.0));
    P = sgb_mul(P, sgb_add(z, 3.0));
    P = sgb_mul(P, sgb_add(z, 4.0));
    P = sgb_mul(P, sgb_add(z, 5.0));
    P = sgb_mul(P, sgb_add(z, 6.0));
    P = sgb_mul(P, sgb_add(z, 7.0));
    P = sgb_mul(P, sgb_add(z, 8.0));
    P = sgb_mul(P, sgb_add(z, 9.0));
    P = sm ? P : 1.0;
    const double sh = sm ? 10.0 : 0.0;
    const double y = sgb_add(z, sh);
    double seed;
    const double ly = sgb_log_core(sgb_d2u(y), 0, LT, &seed);
    const double lP = sgb_log_core(sgb_d2u(P), 0, LT, 0);
    // 1/y by two cubic Newton steps from the 8-bit table seed (for y beyond 2^500 the series vanishes anyway:
    // the seed's exponent arithmetic is then meaningless but every term below is multiplied by rr ~ 0 or NaN-free)
    double e = sgb_fma(-y, seed, 1.0);
    double t = sgb_fma(e, e, e);
    double rr = sgb_fma(seed, t, seed);
    e = sgb_fma(-y, rr, 1.0);
    t = sgb_fma(e, e, e);
    rr = sgb_fma(rr, t, rr);
    rr = (y < 0x1p500) ? rr : 0.0;
    const double r2 = sgb_mul(rr, rr);
    const double r4 = sgb_mul(r2, r2);
    const double c01 = sgb_fma(r2, SGB_C(STIR2), SGB_C(STIR1));
    const double c23 = sgb_fma(r2, SGB_C(STIR4), SGB_C(STIR3));
    const double c45 = sgb_fma(r2, SGB_C(STIR6), SGB_C(STIR5));
    const double c = sgb_fma(r4, sgb_fma(r4, sgb_fma(r4, SGB_C(STIR7), c45), c23), c01);
    double corr = sgb_mul(rr, c);
    // y = fl(z + shift): fold the rounding residual back in with lgamma'(y) ~ log y - 1/(2y)
    const double dy = sgb_sub(z, sgb_sub(y, sh));
    corr = sgb_fma(dy, sgb_fma(-0.5, rr, ly), corr);
    const double res = sgb_add(sgb_fma(sgb_sub(y, 0.5), ly, -y), sgb_add(SGB_C(HALF_LN_2PI), corr));
    return sgb_sub(res, lP);
}

// Stirling part alone: for z >= 10 this performs exactly the operations sgb_lgamma_pos performs there
// (P = 1 -> log P = +0, shift 0 -> residual 0), hence the same bits, without the product and the second log.
SGB_HD double sgb_lgamma_ge10(double z, const double* __restrict__ LT)
{
    double seed;
    const double ly = sgb_log_core(sgb_d2u(z), 0, LT, &seed);
    double e = sgb_fma(-z, seed, 1.0);
    double t = sgb_fma(e, e, e);
    double rr = sgb_fma(seed, t, seed);
    e = sgb_fma(-z, rr, 1.0);
    t = sgb_fma(e, e, e);
    rr = sgb_fma(rr, t, rr);
    rr = (z < 0x1p500) ? rr : 0.0;
    const double r2 = sgb_mul(rr, rr);
    const double r4 = sgb_mul(r2, r2);
    const double c01 = sgb_fma(r2, SGB_C(STIR2), SGB_C(STIR1));
    const double c23 = sgb_fma(r2, SGB_C(STIR4), SGB_C(STIR3));
    const double c45 = sgb_fma(r2, SGB_C(STIR6), SGB_C(STIR5));
    const double c = sgb_fma(r4, sgb_fma(r4, sgb_fma(r4, SGB_C(STIR7), c45), c23), c01);
    const double corr = sgb_mul(rr, c);      // (the residual term is fma(0, finite, corr) = corr there, and log 1 = +0)
    return sgb_add(sgb_fma(sgb_sub(z, 0.5), ly, -z), sgb_add(SGB_C(HALF_LN_2PI), corr));
}

SGB_HD double sgb_lgamma(double z, const double* __restrict__ LT)
{
    if (z != z) return z;
    if (!(z > 0.0)) return (z == 0.0) ? sgb_inf() : sgb_nan();
    if (z == sgb_inf()) return z;
    return sgb_lgamma_pos(z, LT);
}

// ---------------------------------------------------------------- log-factorial
// LF(k) = log(k!) : correctly rounded table for k < SGB_NLF, Stirling above.
SGB_HD double sgb_logfact(int64_t k, const double* __restrict__ LF, const double* __restrict__ LT)
{
    if (k < 0) return sgb_nan();
    if (k < SGB_NLF) return LF[k];
    return sgb_lgamma(sgb_add((double)k, 1.0), LT);
}
