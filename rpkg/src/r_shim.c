/* r_shim.c -- .Call glue between the signals R package and libsignals_b200.so.
 *
 * Neither R nor its headers exist where this repository is built and tested, so this file has never been
 * built inside R.  It is written against R's C API only (no Rcpp); tests/test_r_shim_mock.py compiles it
 * unmodified against a mock of that API (tests/r_mock/) and drives the three entry points -- registration,
 * coercion, error unwinding, and on a GPU box the results -- so what remains unverified is R itself.
 *
 * It keeps the two routines the reference registers (src/RcppExports.cpp:39-48) under their original
 * names so R/RcppExports.R works unchanged, and adds the batched entry points the modified R functions in
 * rpkg/R/hmm_gpu.R call.
 *
 * Build (inside the signals package, in place of src/code.cpp + src/RcppExports.cpp):
 *     PKG_CPPFLAGS = -I<repo>/include
 *     PKG_LIBS     = -L<repo>/signals_b200 -lsignals_b200 -Wl,-rpath,<repo>/signals_b200
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

/* counts arrive as integer or (integer-valued) double vectors */
static SEXP as_int_vec(SEXP x, const char* what)
{
    if (TYPEOF(x) == INTSXP) return x;
    if (TYPEOF(x) != REALSXP) Rf_error("%s must be numeric", what);
    R_xlen_t n = XLENGTH(x);
    const double* p = REAL(x);
    for (R_xlen_t i = 0; i < n; i++)
        if (ISNAN(p[i]) || p[i] != (double)(int)p[i]) Rf_error("%s must be integer valued and not NA", what);
    return Rf_coerceVector(x, INTSXP);
}

/* logspace_addcpp(logx, logy)  -- src/code.cpp:5-8 */
SEXP _signals_logspace_addcpp(SEXP logxSEXP, SEXP logySEXP)
{
    double out = NA_REAL;
    check_rc(sgb_logspace_add(Rf_asReal(logxSEXP), Rf_asReal(logySEXP), &out));
    return Rf_ScalarReal(out);
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
 * .callHaplotypeSpecificCN_ (R/callHSCN.R:244-280) and callAlleleSpecificCN (R/callASCN.R:382-414).
 * seg_off: double or integer vector of 0-based row offsets, length nseg+1.  Returns state_min (numeric). */
SEXP _signals_hmm_decode_batch(SEXP model, SEXP totalcounts, SEXP alleleB, SEXP state, SEXP seg_off,
                               SEXP likelihood, SEXP maxCN, SEXP loherror, SEXP eps, SEXP rho, SEXP Abias,
                               SEXP stp, SEXP want_l)
{
    sgb_params P;
    fill_params(&P, model, likelihood, maxCN, loherror, eps, rho, Abias, stp);
    SEXP n = PROTECT(as_int_vec(totalcounts, "totalcounts"));
    SEXP x = PROTECT(as_int_vec(alleleB, "alleleB"));
    SEXP c = PROTECT(as_int_vec(state, "state"));
    R_xlen_t N = XLENGTH(n);
    if (XLENGTH(x) != N || XLENGTH(c) != N) Rf_error("totalcounts, alleleB and state must have the same length");
    SEXP so = PROTECT(Rf_coerceVector(seg_off, REALSXP));
    R_xlen_t nseg = XLENGTH(so) - 1;
    int64_t* off = (int64_t*)R_alloc((size_t)nseg + 1, sizeof(int64_t));
    for (R_xlen_t i = 0; i <= nseg; i++) off[i] = (int64_t)REAL(so)[i];
    uint8_t* path = (uint8_t*)R_alloc((size_t)(N > 0 ? N : 1), 1);
    SEXP l = R_NilValue;
    int np = 4;
    if (Rf_asLogical(want_l)) { l = PROTECT(Rf_allocVector(REALSXP, N * P.n_states)); np++; }
    check_rc(sgb_hmm_decode(&P, (int64_t)N, INTEGER(n), INTEGER(x), INTEGER(c), 4, (int64_t)nseg, off, path,
                            l == R_NilValue ? NULL : REAL(l)));
    SEXP res = PROTECT(Rf_allocVector(REALSXP, N)); np++;
    for (R_xlen_t i = 0; i < N; i++) REAL(res)[i] = (double)path[i];
    SEXP out = PROTECT(Rf_allocVector(VECSXP, 2)); np++;
    SET_VECTOR_ELT(out, 0, res);
    SET_VECTOR_ELT(out, 1, l);
    UNPROTECT(np);
    return out;
}

static const R_CallMethodDef CallEntries[] = {
    {"_signals_logspace_addcpp", (DL_FUNC)&_signals_logspace_addcpp, 2},
    {"_signals_viterbi", (DL_FUNC)&_signals_viterbi, 3},
    {"_signals_hmm_decode_batch", (DL_FUNC)&_signals_hmm_decode_batch, 13},
    {NULL, NULL, 0}};

void R_init_signals(DllInfo* dll)
{
    R_registerRoutines(dll, NULL, CallEntries, NULL, NULL);
    R_useDynamicSymbols(dll, FALSE);
}
