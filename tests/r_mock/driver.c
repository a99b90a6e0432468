/* driver.c -- calls the .Call entry points of rpkg/src/r_shim.c the way R would, on mock SEXPs, and hands the
 * results back through plain C arrays (loaded with ctypes by tests/test_r_shim_mock.py).  Test infrastructure. */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <setjmp.h>
#include <string.h>
#include <time.h>

extern jmp_buf mock_error_jmp;
extern int mock_error_armed;
extern char mock_error_msg[1024];
extern const R_CallMethodDef* mock_registered_calls;
extern int mock_dynamic_symbols;

void R_init_signals(DllInfo* dll);
SEXP _signals_logspace_addcpp(SEXP, SEXP);
SEXP _signals_viterbi(SEXP, SEXP, SEXP);
SEXP _signals_hmm_decode_batch(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
SEXP _signals_logspace_add_vec(SEXP, SEXP);
SEXP _signals_postprocess(SEXP, SEXP);
SEXP _signals_pass_stats(SEXP, SEXP, SEXP, SEXP, SEXP);
SEXP _signals_device_count(void);
SEXP _signals_cols_upload(SEXP, SEXP, SEXP, SEXP);
SEXP _signals_cols_decode(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
SEXP _signals_cols_free(SEXP);

static int fail(char* err)
{
    if (err) { strncpy(err, mock_error_msg, 1023); err[1023] = 0; }
    mock_error_armed = 0;
    mock_reset();
    return 1;
}
#define GUARDED(err) mock_error_armed = 1; if (setjmp(mock_error_jmp)) return fail(err)
#define DONE() mock_error_armed = 0; mock_reset(); return 0

/* names (';'-separated) and arities of what R_init_signals registers; returns the number of routines */
int drv_registered(char* names, int names_cap, int* nargs, int nargs_cap, int* dynamic)
{
    DllInfo info = {NULL, -1};
    R_init_signals(&info);
    int k = 0;
    names[0] = 0;
    for (const R_CallMethodDef* d = mock_registered_calls; d && d->name; d++, k++) {
        if ((int)(strlen(names) + strlen(d->name) + 2) < names_cap) { strcat(names, d->name); strcat(names, ";"); }
        if (k < nargs_cap) nargs[k] = d->numArgs;
    }
    *dynamic = mock_dynamic_symbols;
    return k;
}

int drv_logspace_add(double a, double b, double* out, char* err)
{
    GUARDED(err);
    SEXP r = _signals_logspace_addcpp(Rf_ScalarReal(a), Rf_ScalarReal(b));
    *out = REAL(r)[0];
    DONE();
}

int drv_viterbi(const double* emission, int nrow, int S, const double* transition, const double* observations, int T,
                double* path_out, char* err)
{
    GUARDED(err);
    /* R passes `observations` as a numeric (double) vector: 1:nrow */
    SEXP r = _signals_viterbi(mock_real_matrix(emission, nrow, S), mock_real_matrix(transition, S, S),
                              mock_real_vector(observations, T));
    if (XLENGTH(r) != T) Rf_error("driver: viterbi returned %ld values for T = %d", (long)XLENGTH(r), T);
    memcpy(path_out, REAL(r), (size_t)T * sizeof(double));
    DONE();
}

/* counts_as_double: hand totalcounts / alleleB / state over as double vectors (as data.table columns often are) */
int drv_decode_batch(int model, const int* n, const int* x, const int* state, long N, const double* seg_off, long nseg,
                     const char* likelihood, int maxCN, double loherror, double eps, double rho, double Abias, double stp,
                     int want_l, int counts_as_double, int ngpus, double* path_out, double* l_out, char* err)
{
    GUARDED(err);
    SEXP sn, sx, sc;
    if (counts_as_double) {
        sn = Rf_coerceVector(mock_int_vector(n, N), REALSXP);
        sx = Rf_coerceVector(mock_int_vector(x, N), REALSXP);
        sc = Rf_coerceVector(mock_int_vector(state, N), REALSXP);
    } else {
        sn = mock_int_vector(n, N); sx = mock_int_vector(x, N); sc = mock_int_vector(state, N);
    }
    int mdl = model, mcn = maxCN;
    SEXP r = _signals_hmm_decode_batch(mock_int_vector(&mdl, 1), sn, sx, sc, mock_real_vector(seg_off, nseg + 1),
                                       mock_string(likelihood), mock_int_vector(&mcn, 1), Rf_ScalarReal(loherror),
                                       Rf_ScalarReal(eps), Rf_ScalarReal(rho), Rf_ScalarReal(Abias), Rf_ScalarReal(stp),
                                       mock_logical(want_l), mock_int_vector(&ngpus, 1));
    if (TYPEOF(r) != VECSXP || XLENGTH(r) != 2) Rf_error("driver: decode_batch must return a list of 2");
    SEXP p = VECTOR_ELT(r, 0), l = VECTOR_ELT(r, 1);
    if (XLENGTH(p) != N) Rf_error("driver: path length %ld != N", (long)XLENGTH(p));
    memcpy(path_out, REAL(p), (size_t)N * sizeof(double));
    if (want_l) {
        if (l == R_NilValue || XLENGTH(l) != N * (maxCN + 1)) Rf_error("driver: l has the wrong length");
        memcpy(l_out, REAL(l), (size_t)XLENGTH(l) * sizeof(double));
    } else if (l != R_NilValue) {
        Rf_error("driver: l should be NULL when not requested");
    }
    DONE();
}

int drv_logspace_add_vec(const double* a, const double* b, long n, double* out, char* err)
{
    GUARDED(err);
    SEXP r = _signals_logspace_add_vec(mock_real_vector(a, n), mock_real_vector(b, n));
    if (XLENGTH(r) != n) Rf_error("driver: logspace_add_vec returned the wrong length");
    memcpy(out, REAL(r), (size_t)n * sizeof(double));
    DONE();
}

/* out columns: A, B, state_min, LOH, phase, state_phase as ints (6 x N), state_BAF as doubles */
int drv_postprocess(const int* state, const double* state_min, long N, int* ints_out, double* baf_out, char* err)
{
    GUARDED(err);
    SEXP r = _signals_postprocess(mock_int_vector(state, N), mock_real_vector(state_min, N));
    if (TYPEOF(r) != VECSXP || XLENGTH(r) != 7) Rf_error("driver: postprocess must return a list of 7");
    for (int k = 0; k < 6; k++) memcpy(ints_out + (size_t)k * N, INTEGER(VECTOR_ELT(r, k)), (size_t)N * sizeof(int));
    memcpy(baf_out, REAL(VECTOR_ELT(r, 6)), (size_t)N * sizeof(double));
    DONE();
}

int drv_pass_stats(const int* n, const int* x, const int* state, const double* state_min, long N, int maxCN,
                   double* sums3, double* hist, char* err)
{
    GUARDED(err);
    int mcn = maxCN;
    SEXP r = _signals_pass_stats(mock_int_vector(n, N), mock_int_vector(x, N), mock_int_vector(state, N),
                                 mock_real_vector(state_min, N), mock_int_vector(&mcn, 1));
    memcpy(sums3, REAL(VECTOR_ELT(r, 0)), 3 * sizeof(double));
    memcpy(hist, REAL(VECTOR_ELT(r, 1)), (size_t)(maxCN + 1) * (maxCN + 1) * sizeof(double));
    DONE();
}

int drv_device_count(int* out, char* err)
{
    GUARDED(err);
    *out = INTEGER(_signals_device_count())[0];
    DONE();
}

/* Times ONLY the .Call body (the vectors are built before the clock starts, as they already exist in an R session):
 * `reps` calls of _signals_hmm_decode_batch on pageable INTSXP (as_double = 0) or REALSXP (1) columns, numeric states
 * back.  ms_out = mean milliseconds per call; checksum_out = sum of the decoded states of the last call. */
int drv_decode_batch_timed(int model, const int* n, const int* x, const int* state, long N, const double* seg_off, long nseg,
                           const char* likelihood, int maxCN, double loherror, double eps, double rho, double Abias, double stp,
                           int as_double, int ngpus, int reps, double* ms_out, double* checksum_out, char* err)
{
    GUARDED(err);
    SEXP sn = mock_int_vector(n, N), sx = mock_int_vector(x, N), sc = mock_int_vector(state, N);
    if (as_double) { sn = Rf_coerceVector(sn, REALSXP); sx = Rf_coerceVector(sx, REALSXP); sc = Rf_coerceVector(sc, REALSXP); }
    int mdl = model, mcn = maxCN;
    SEXP a0 = mock_int_vector(&mdl, 1), so = mock_real_vector(seg_off, nseg + 1), lk = mock_string(likelihood);
    SEXP a6 = mock_int_vector(&mcn, 1), wl = mock_logical(0), ng = mock_int_vector(&ngpus, 1);
    struct timespec t0, t1;
    SEXP r = R_NilValue;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int k = 0; k < reps; k++)
        r = _signals_hmm_decode_batch(a0, sn, sx, sc, so, lk, a6, Rf_ScalarReal(loherror), Rf_ScalarReal(eps), Rf_ScalarReal(rho),
                                      Rf_ScalarReal(Abias), Rf_ScalarReal(stp), wl, ng);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    *ms_out = ((t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) * 1e-6) / (reps > 0 ? reps : 1);
    double cs = 0.0;
    const double* p = REAL(VECTOR_ELT(r, 0));
    for (long i = 0; i < N; i++) cs += p[i];
    *checksum_out = cs;
    DONE();
}

/* ---- the table handle: upload once (handle returned to the caller as an opaque pointer), decode passes, free / gc */
int drv_cols_upload(const int* n, const int* x, const int* state, long N, const double* seg_off, long nseg, int as_double,
                    void** handle_out, double* nrows_out, char* err)
{
    GUARDED(err);
    SEXP sn = mock_int_vector(n, N), sx = mock_int_vector(x, N), sc = mock_int_vector(state, N);
    if (as_double) { sn = Rf_coerceVector(sn, REALSXP); sx = Rf_coerceVector(sx, REALSXP); sc = Rf_coerceVector(sc, REALSXP); }
    SEXP r = _signals_cols_upload(sn, sx, sc, mock_real_vector(seg_off, nseg + 1));
    if (TYPEOF(r) != VECSXP || XLENGTH(r) != 2 || TYPEOF(VECTOR_ELT(r, 0)) != EXTPTRSXP) Rf_error("driver: upload must return list(handle, nrows)");
    *handle_out = (void*)VECTOR_ELT(r, 0);
    *nrows_out = REAL(VECTOR_ELT(r, 1))[0];
    DONE();
}

int drv_cols_decode(void* handle, double nrows, int model, const char* likelihood, int maxCN, double loherror, double eps, double rho,
                    double Abias, double stp, double* path_out, char* err)
{
    GUARDED(err);
    int mdl = model, mcn = maxCN;
    SEXP r = _signals_cols_decode((SEXP)handle, Rf_ScalarReal(nrows), mock_int_vector(&mdl, 1), mock_string(likelihood),
                                  mock_int_vector(&mcn, 1), Rf_ScalarReal(loherror), Rf_ScalarReal(eps), Rf_ScalarReal(rho),
                                  Rf_ScalarReal(Abias), Rf_ScalarReal(stp));
    if (TYPEOF(r) != REALSXP || (double)XLENGTH(r) != nrows) Rf_error("driver: decode must return a numeric vector of nrows");
    memcpy(path_out, REAL(r), (size_t)XLENGTH(r) * sizeof(double));
    DONE();
}

int drv_cols_free(void* handle, char* err)
{
    GUARDED(err);
    _signals_cols_free((SEXP)handle);
    DONE();
}

/* a scalar where a handle is expected (argument checking) */
int drv_cols_decode_bad_handle(char* err)
{
    GUARDED(err);
    int one = 1;
    _signals_cols_decode(Rf_ScalarReal(1.0), Rf_ScalarReal(1.0), mock_int_vector(&one, 1), mock_string("binomial"), mock_int_vector(&one, 1),
                         Rf_ScalarReal(0.02), Rf_ScalarReal(1e-12), Rf_ScalarReal(0.0), Rf_ScalarReal(0.0), Rf_ScalarReal(0.95));
    DONE();
}

int drv_gc(void) { return mock_gc(); }
