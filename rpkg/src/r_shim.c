/* r_shim.c -- .Call glue between the signals R package and libsignals_b200.so.
 *
 * Neither R nor its headers exist where this repository is built and tested, so this file has never been
 * built inside R.  It is written against R's C API only (no Rcpp); tests/test_r_shim_mock.py compiles it
 * unmodified against a mock of that API (tests/r_mock/) and drives every entry point -- registration,
 * argument handling, error unwinding, and on a GPU box the results -- so what remains unverified is R itself.
 *
 * It keeps the two routines the reference registers (src/RcppExports.cpp:39-48) under their original
 * names so R/RcppExports.R works unchanged, and adds the batched entry points the modified R functions in
 * rpkg/R/hmm_gpu.R call.  Build: rpkg/src/Makevars (in place of src/code.cpp + src/RcppExports.cpp).
 *
 * Data columns are handed to the library AS THEY ARE: an INTSXP column as int32, a REALSXP column as double
 * (data.table columns of counts are often numeric) -- no coercion, no copy and no element-wise check on the R thread;
 * the library stages pageable R memory through page-locked buffers with worker threads, validates on the way
 * (NA / NaN / fractions / negative counts are an error: the reference would yield NaN or -Inf rows) and writes the
 * decoded states straight into a numeric vector.
 */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <stdint.h>
#include <string.h>
#include "signals_b200.h"

static void check_rc(int rc)
{
    if (rc != 0) Rf_error("signals_b200: %s", sgb_last_error());     /* BEGIN_RCPP/END_RCPP equivalent */
}

/* element type code of a data column (no conversion is made) */
static int col_type(SEXP x, const char* what)
{
    if (TYPEOF(x) == INTSXP) return SGB_T_I32;
    if (TYPEOF(x) == REALSXP) return SGB_T_F64;
    Rf_error("%s must be an integer or numeric vector", what);
    return 0;
}
static const void* col_ptr(SEXP x) { return TYPEOF(x) == INTSXP ? (const void*)INTEGER(x) : (const void*)REAL(x); }

/* seg_off (0-based row offsets, length nseg + 1, integer or numeric) as int64 */
static int64_t* seg_offsets(SEXP seg_off, R_xlen_t* nseg)
{
    SEXP so = PROTECT(Rf_coerceVector(seg_off, REALSXP));
    *nseg = XLENGTH(so) - 1;
    if (*nseg < 0) Rf_error("seg_off must have at least one element");
    int64_t* off = (int64_t*)R_alloc((size_t)*nseg + 1, sizeof(int64_t));
    for (R_xlen_t i = 0; i <= *nseg; i++) {
        const double v = REAL(so)[i];
        if (ISNAN(v) || v < 0 || v > 9007199254740992.0) Rf_error("seg_off holds NA or a value out of range");
        off[i] = (int64_t)v;
    }
    UNPROTECT(1);
    return off;
}

/* logspace_addcpp(logx, logy)  -- src/code.cpp:5-8.  NOTE: one scalar call is a GPU launch plus two copies; R code
 * that loops over elements should call logspace_add (vectorised, _signals_logspace_add_vec) instead. */
SEXP _signals_logspace_addcpp(SEXP logxSEXP, SEXP logySEXP)
{
    double out = NA_REAL;
    check_rc(sgb_logspace_add(Rf_asReal(logxSEXP), Rf_asReal(logySEXP), &out));
    return Rf_ScalarReal(out);
}

/* logspace_add(logx, logy), vectorised (R/callASCN.R:9-11); both arguments numeric of the same length */
SEXP _signals_logspace_add_vec(SEXP logxSEXP, SEXP logySEXP)
{
    SEXP a = PROTECT(Rf_coerceVector(logxSEXP, REALSXP));
    SEXP b = PROTECT(Rf_coerceVector(logySEXP, REALSXP));
    R_xlen_t n = XLENGTH(a);
    if (XLENGTH(b) != n) Rf_error("logx and logy must have the same length");
    SEXP out = PROTECT(Rf_allocVector(REALSXP, n));
    check_rc(sgb_logspace_add_vec(REAL(a), REAL(b), (int64_t)n, REAL(out)));
    UNPROTECT(3);
    return out;
}

/* viterbi(emission, transition, observations)  -- src/code.cpp:11-45 ; returns 0-based states as doubles */
SEXP _signals_viterbi(SEXP emissionSEXP, SEXP transitionSEXP, SEXP observationsSEXP)
{
    SEXP em = PROTECT(Rf_coerceVector(emissionSEXP, REALSXP));
    SEXP tr = PROTECT(Rf_coerceVector(transitionSEXP, REALSXP));
    SEXP ob = PROTECT(Rf_coerceVector(observationsSEXP, INTSXP));
    SEXP edim = Rf_getAttrib(emissionSEXP, R_DimSymbol), tdim = Rf_getAttrib(transitionSEXP, R_DimSymbol);
    if (Rf_length(edim) != 2 || Rf_length(tdim) != 2) Rf_error("emission and transition must be matrices");
    int64_t nrow_e = INTEGER(edim)[0];
    int S = INTEGER(tdim)[0];
    if (INTEGER(edim)[1] != S || INTEGER(tdim)[1] != S) Rf_error("emission / transition dimensions disagree");
    R_xlen_t T = XLENGTH(ob);
    SEXP pathI = PROTECT(Rf_allocVector(INTSXP, T));
    check_rc(sgb_viterbi(REAL(em), nrow_e, REAL(tr), INTEGER(ob), (int64_t)T, S, 0, SGB_DECODE_REFERENCE, INTEGER(pathI)));
    SEXP path = PROTECT(Rf_allocVector(REALSXP, T));
    for (R_xlen_t i = 0; i < T; i++) REAL(path)[i] = (double)INTEGER(pathI)[i];
    UNPROTECT(5);
    return path;
}

static void fill_params(sgb_params* P, SEXP model, SEXP likelihood, SEXP maxCN, SEXP loherror, SEXP eps, SEXP rho,
                        SEXP Abias, SEXP stp)
{
    const char* lik = CHAR(STRING_ELT(likelihood, 0));
    if (strcmp(lik, "binomial") != 0 && strcmp(lik, "betabinomial") != 0)
        Rf_error("Likelihood model must be one of binomial or betabinomial");        /* R/callHSCN.R:778-780 */
    P->model = Rf_asInteger(model);
    P->likelihood = strcmp(lik, "betabinomial") == 0 ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL;
    P->n_states = Rf_asInteger(maxCN) + 1;                                            /* minor_cn = 0:maxCN */
    P->decode = SGB_DECODE_REFERENCE;
    P->loherror = Rf_asReal(loherror);
    P->eps = Rf_asReal(eps);
    P->rho = Rf_asReal(rho);
    P->Abias = Rf_asReal(Abias);
    P->selftransitionprob = Rf_asReal(stp);
}

/* Batched decode of a table sorted by (cell_id, chr, start): replaces the per-cell loops of
 * .callHaplotypeSpecificCN_ (R/callHSCN.R:244-280) and callAlleleSpecificCN (R/callASCN.R:382-414, :556-588).
 * seg_off: 0-based row offsets, length nseg+1.  ngpus: CUDA devices to shard the cells over (the role of ncores /
 * mclapply).  Returns list(state_min (numeric), l (numeric or NULL)). */
SEXP _signals_hmm_decode_batch(SEXP model, SEXP totalcounts, SEXP alleleB, SEXP state, SEXP seg_off,
                               SEXP likelihood, SEXP maxCN, SEXP loherror, SEXP eps, SEXP rho, SEXP Abias,
                               SEXP stp, SEXP want_l, SEXP ngpus)
{
    sgb_params P;
    fill_params(&P, model, likelihood, maxCN, loherror, eps, rho, Abias, stp);
    const int tn = col_type(totalcounts, "totalcounts"), tx = col_type(alleleB, "alleleB"), tc = col_type(state, "state");
    R_xlen_t N = XLENGTH(totalcounts);
    if (XLENGTH(alleleB) != N || XLENGTH(state) != N) Rf_error("totalcounts, alleleB and state must have the same length");
    R_xlen_t nseg = 0;
    int64_t* off = seg_offsets(seg_off, &nseg);
    int nd = Rf_asInteger(ngpus);
    if (nd == NA_INTEGER || nd < 1) nd = 1;
    const int avail = sgb_device_count();
    if (avail < 1) Rf_error("signals_b200: %s", sgb_last_error());
    if (nd > avail) nd = avail;
    SEXP res = PROTECT(Rf_allocVector(REALSXP, N));
    SEXP l = R_NilValue;
    int np = 1;
    if (Rf_asLogical(want_l) == TRUE) {
        /* the `l` matrix of a single HaplotypeHMM / alleleHMM call: small, so the int32 / uint8 entry point (with copies) */
        SEXP n = PROTECT(Rf_coerceVector(totalcounts, INTSXP));
        SEXP x = PROTECT(Rf_coerceVector(alleleB, INTSXP));
        SEXP c = PROTECT(Rf_coerceVector(state, INTSXP));
        np += 3;
        for (R_xlen_t i = 0; i < N; i++)
            if (INTEGER(n)[i] == NA_INTEGER || INTEGER(x)[i] == NA_INTEGER || INTEGER(c)[i] == NA_INTEGER)
                Rf_error("signals_b200: NA in totalcounts / alleleB / state");
        uint8_t* path = (uint8_t*)R_alloc((size_t)(N > 0 ? N : 1), 1);
        l = PROTECT(Rf_allocVector(REALSXP, N * P.n_states)); np++;
        check_rc(sgb_hmm_decode(&P, (int64_t)N, INTEGER(n), INTEGER(x), INTEGER(c), 4, (int64_t)nseg, off, path, REAL(l)));
        for (R_xlen_t i = 0; i < N; i++) REAL(res)[i] = (double)path[i];
    } else {
        check_rc(sgb_hmm_decode_typed(&P, (int64_t)N, col_ptr(totalcounts), tn, col_ptr(alleleB), tx, col_ptr(state), tc,
                                      (int64_t)nseg, off, REAL(res), SGB_T_F64, nd));
    }
    SEXP out = PROTECT(Rf_allocVector(VECSXP, 2)); np++;
    SET_VECTOR_ELT(out, 0, res);
    SET_VECTOR_ELT(out, 1, l);
    UNPROTECT(np);
    return out;
}

/* The table kept on the device between passes (sgb_cols_*): callHaplotypeSpecificCN runs .callHaplotypeSpecificCN_ twice on
 * the same CNBAF (R/callHSCN.R:883-935), callAlleleSpecificCN once more (R/callASCN.R:372-414); with a handle the second and
 * third pass cost the kernels and the copy of the states, not another conversion + upload of the three columns.
 * _signals_cols_upload returns an external pointer (finalizer: sgb_cols_free); _signals_cols_decode runs one pass and
 * returns state_min as numeric; _signals_cols_free releases the device memory at once. */
static void cols_finalizer(SEXP h)
{
    sgb_cols* c = (sgb_cols*)R_ExternalPtrAddr(h);
    if (c) {
        sgb_cols_free(c);
        R_ClearExternalPtr(h);
    }
}

SEXP _signals_cols_upload(SEXP totalcounts, SEXP alleleB, SEXP state, SEXP seg_off)
{
    const int tn = col_type(totalcounts, "totalcounts"), tx = col_type(alleleB, "alleleB"), tc = col_type(state, "state");
    R_xlen_t N = XLENGTH(totalcounts);
    if (XLENGTH(alleleB) != N || XLENGTH(state) != N) Rf_error("totalcounts, alleleB and state must have the same length");
    R_xlen_t nseg = 0;
    int64_t* off = seg_offsets(seg_off, &nseg);
    sgb_cols* c = NULL;
    check_rc(sgb_cols_upload((int64_t)N, col_ptr(totalcounts), tn, col_ptr(alleleB), tx, col_ptr(state), tc, (int64_t)nseg, off, &c));
    SEXP h = PROTECT(R_MakeExternalPtr(c, R_NilValue, R_NilValue));
    R_RegisterCFinalizerEx(h, cols_finalizer, TRUE);
    SEXP len = PROTECT(Rf_allocVector(REALSXP, 1));
    REAL(len)[0] = (double)N;
    SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));       /* list(handle, number of rows) */
    SET_VECTOR_ELT(out, 0, h);
    SET_VECTOR_ELT(out, 1, len);
    UNPROTECT(3);
    return out;
}

SEXP _signals_cols_decode(SEXP handle, SEXP nrows, SEXP model, SEXP likelihood, SEXP maxCN, SEXP loherror, SEXP eps, SEXP rho,
                          SEXP Abias, SEXP stp)
{
    if (TYPEOF(handle) != EXTPTRSXP) Rf_error("handle must come from the table upload");
    sgb_cols* c = (sgb_cols*)R_ExternalPtrAddr(handle);
    if (!c) Rf_error("signals_b200: the table has been released");
    sgb_params P;
    fill_params(&P, model, likelihood, maxCN, loherror, eps, rho, Abias, stp);
    const double nr = Rf_asReal(nrows);
    if (ISNAN(nr) || nr < 0) Rf_error("bad row count");
    SEXP res = PROTECT(Rf_allocVector(REALSXP, (R_xlen_t)nr));
    check_rc(sgb_cols_decode(c, &P, REAL(res), SGB_T_F64));
    UNPROTECT(1);
    return res;
}

SEXP _signals_cols_free(SEXP handle)
{
    if (TYPEOF(handle) != EXTPTRSXP) Rf_error("handle must come from the table upload");
    cols_finalizer(handle);
    return R_NilValue;
}

/* Post-decode columns on the device: the clamp block R/callHSCN.R:282-290 and the numeric part of add_states
 * (R/util.R:1078-1097).  state: integer vector; state_min: numeric / integer vector of decoded states (0..255).
 * Returns list(A, B, state_min, LOH (1/0), phase (0/1/2), state_phase (0..4), state_BAF). */
SEXP _signals_postprocess(SEXP state, SEXP state_min)
{
    SEXP c = PROTECT(Rf_coerceVector(state, INTSXP));
    SEXP pm = PROTECT(Rf_coerceVector(state_min, INTSXP));
    R_xlen_t N = XLENGTH(c);
    if (XLENGTH(pm) != N) Rf_error("state and state_min must have the same length");
    uint8_t* path = (uint8_t*)R_alloc((size_t)(N > 0 ? N : 1), 1);
    for (R_xlen_t i = 0; i < N; i++) {
        const int v = INTEGER(pm)[i];
        if (v == NA_INTEGER || v < 0 || v > 255) Rf_error("state_min must be in 0..255");
        path[i] = (uint8_t)v;
    }
    uint8_t* b3 = (uint8_t*)R_alloc((size_t)(N > 0 ? N : 1), 3);
    SEXP A = PROTECT(Rf_allocVector(INTSXP, N)), B = PROTECT(Rf_allocVector(INTSXP, N)), SM = PROTECT(Rf_allocVector(INTSXP, N));
    SEXP BAF = PROTECT(Rf_allocVector(REALSXP, N));
    check_rc(sgb_postprocess((int64_t)N, INTEGER(c), 4, path, INTEGER(A), INTEGER(B), INTEGER(SM), b3, b3 + N, b3 + 2 * N, REAL(BAF)));
    SEXP LOH = PROTECT(Rf_allocVector(INTSXP, N)), PH = PROTECT(Rf_allocVector(INTSXP, N)), SP = PROTECT(Rf_allocVector(INTSXP, N));
    for (R_xlen_t i = 0; i < N; i++) { INTEGER(LOH)[i] = b3[i]; INTEGER(PH)[i] = b3[N + i]; INTEGER(SP)[i] = b3[2 * N + i]; }
    SEXP out = PROTECT(Rf_allocVector(VECSXP, 7));
    SET_VECTOR_ELT(out, 0, A); SET_VECTOR_ELT(out, 1, B); SET_VECTOR_ELT(out, 2, SM); SET_VECTOR_ELT(out, 3, LOH);
    SET_VECTOR_ELT(out, 4, PH); SET_VECTOR_ELT(out, 5, SP); SET_VECTOR_ELT(out, 6, BAF);
    UNPROTECT(10);
    return out;
}

/* Between-pass sums on the device (R/callHSCN.R:898-902, :620-621): list(c(sum alleleB, sum totalcounts, rows) over the
 * A-Hom rows, A x B histogram (maxCN+1 square, column-major [A, B] as an R matrix indexed [A+1, B+1])) */
SEXP _signals_pass_stats(SEXP totalcounts, SEXP alleleB, SEXP state, SEXP state_min, SEXP maxCN)
{
    SEXP n = PROTECT(Rf_coerceVector(totalcounts, INTSXP));
    SEXP x = PROTECT(Rf_coerceVector(alleleB, INTSXP));
    SEXP c = PROTECT(Rf_coerceVector(state, INTSXP));
    SEXP pm = PROTECT(Rf_coerceVector(state_min, INTSXP));
    R_xlen_t N = XLENGTH(n);
    if (XLENGTH(x) != N || XLENGTH(c) != N || XLENGTH(pm) != N) Rf_error("columns must have the same length");
    const int S = Rf_asInteger(maxCN) + 1;
    if (S < 1 || S > 32) Rf_error("maxCN must be in 0..31");
    uint8_t* path = (uint8_t*)R_alloc((size_t)(N > 0 ? N : 1), 1);
    for (R_xlen_t i = 0; i < N; i++) {
        const int v = INTEGER(pm)[i];
        if (v == NA_INTEGER || v < 0 || v > 255) Rf_error("state_min must be in 0..255");
        path[i] = (uint8_t)v;
    }
    int64_t sums[3];
    int64_t* hist = (int64_t*)R_alloc((size_t)S * S, sizeof(int64_t));
    check_rc(sgb_pass_stats((int64_t)N, INTEGER(n), INTEGER(x), INTEGER(c), 4, path, S, sums, hist));
    SEXP s = PROTECT(Rf_allocVector(REALSXP, 3));
    for (int i = 0; i < 3; i++) REAL(s)[i] = (double)sums[i];
    SEXP h = PROTECT(Rf_allocVector(REALSXP, (R_xlen_t)S * S));
    for (int a = 0; a < S; a++)
        for (int b = 0; b < S; b++) REAL(h)[a + (R_xlen_t)S * b] = (double)hist[a * S + b];
    SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));
    SET_VECTOR_ELT(out, 0, s); SET_VECTOR_ELT(out, 1, h);
    UNPROTECT(7);
    return out;
}

/* number of CUDA devices visible (for mapping ncores / options(signals.gpus)) */
SEXP _signals_device_count(void)
{
    SEXP r = PROTECT(Rf_allocVector(INTSXP, 1));
    INTEGER(r)[0] = sgb_device_count();
    UNPROTECT(1);
    return r;
}

static const R_CallMethodDef CallEntries[] = {
    {"_signals_logspace_addcpp", (DL_FUNC)&_signals_logspace_addcpp, 2},
    {"_signals_viterbi", (DL_FUNC)&_signals_viterbi, 3},
    {"_signals_hmm_decode_batch", (DL_FUNC)&_signals_hmm_decode_batch, 14},
    {"_signals_logspace_add_vec", (DL_FUNC)&_signals_logspace_add_vec, 2},
    {"_signals_postprocess", (DL_FUNC)&_signals_postprocess, 2},
    {"_signals_pass_stats", (DL_FUNC)&_signals_pass_stats, 5},
    {"_signals_device_count", (DL_FUNC)&_signals_device_count, 0},
    {"_signals_cols_upload", (DL_FUNC)&_signals_cols_upload, 4},
    {"_signals_cols_decode", (DL_FUNC)&_signals_cols_decode, 10},
    {"_signals_cols_free", (DL_FUNC)&_signals_cols_free, 1},
    {NULL, NULL, 0}};

void R_init_signals(DllInfo* dll)
{
    R_registerRoutines(dll, NULL, CallEntries, NULL, NULL);
    R_useDynamicSymbols(dll, FALSE);
}
