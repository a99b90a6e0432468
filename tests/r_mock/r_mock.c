/* r_mock.c -- minimal implementation of the mock R API declared in Rinternals.h (test infrastructure only) */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <setjmp.h>
#include <stdarg.h>
#include <string.h>
#include <limits.h>

static struct mock_sexp nil_obj = {NILSXP, 0, NULL, NULL};
static struct mock_sexp dim_obj = {CHARSXP, 3, (void*)"dim", NULL};
SEXP R_NilValue = &nil_obj;
SEXP R_DimSymbol = &dim_obj;
double R_NaReal = NAN;
int R_NaInt = INT_MIN;

/* allocation list, released by mock_reset() */
static void** g_blocks = NULL;
static size_t g_nblocks = 0, g_cap = 0;
static void* track(void* p)
{
    if (g_nblocks == g_cap) {
        g_cap = g_cap ? 2 * g_cap : 64;
        g_blocks = (void**)realloc(g_blocks, g_cap * sizeof(void*));
    }
    g_blocks[g_nblocks++] = p;
    return p;
}
void mock_reset(void)
{
    for (size_t i = 0; i < g_nblocks; i++) free(g_blocks[i]);
    g_nblocks = 0;
}

/* Rf_error unwinds to the driver with the formatted message */
jmp_buf mock_error_jmp;
int mock_error_armed = 0;
char mock_error_msg[1024];
void Rf_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(mock_error_msg, sizeof mock_error_msg, fmt, ap);
    va_end(ap);
    if (mock_error_armed) longjmp(mock_error_jmp, 1);
    fprintf(stderr, "mock Rf_error outside a driver call: %s\n", mock_error_msg);
    abort();
}

static size_t elt_size(int type)
{
    switch (type) {
    case INTSXP: case LGLSXP: return sizeof(int);
    case REALSXP: return sizeof(double);
    case STRSXP: case VECSXP: return sizeof(SEXP);
    case CHARSXP: return 1;
    default: return 0;
    }
}
SEXP Rf_allocVector(int type, R_xlen_t n)
{
    size_t es = elt_size(type);
    if (!es || n < 0) Rf_error("mock: cannot allocate a vector of type %d", type);
    SEXP s = (SEXP)track(calloc(1, sizeof(struct mock_sexp)));
    s->type = type; s->len = n; s->dim = R_NilValue;
    /* like R, numeric / integer vectors come back uninitialised (zeroing 456 MB on one thread would be the mock's cost, not R's) */
    s->data = track((type == REALSXP || type == INTSXP) ? malloc(((size_t)n + 1) * es) : calloc((size_t)n + 1, es));
    if (type == STRSXP || type == VECSXP)
        for (R_xlen_t i = 0; i < n; i++) ((SEXP*)s->data)[i] = R_NilValue;
    return s;
}
int TYPEOF(SEXP x) { return x->type; }
R_xlen_t XLENGTH(SEXP x) { return x->len; }
int Rf_length(SEXP x) { return (int)x->len; }
int* INTEGER(SEXP x)
{
    if (x->type != INTSXP && x->type != LGLSXP) Rf_error("mock: INTEGER() on a non-integer vector");
    return (int*)x->data;
}
double* REAL(SEXP x)
{
    if (x->type != REALSXP) Rf_error("mock: REAL() on a non-double vector");
    return (double*)x->data;
}
SEXP Rf_coerceVector(SEXP x, int type)
{
    if (x->type == type) return x;
    SEXP r = Rf_allocVector(type, x->len);
    r->dim = x->dim;
    if (x->type == REALSXP && type == INTSXP) {
        for (R_xlen_t i = 0; i < x->len; i++) {
            double v = ((double*)x->data)[i];
            ((int*)r->data)[i] = isnan(v) ? R_NaInt : (int)v;
        }
    } else if ((x->type == INTSXP || x->type == LGLSXP) && type == REALSXP) {
        for (R_xlen_t i = 0; i < x->len; i++) {
            int v = ((int*)x->data)[i];
            ((double*)r->data)[i] = v == R_NaInt ? R_NaReal : (double)v;
        }
    } else {
        Rf_error("mock: unsupported coercion %d -> %d", x->type, type);
    }
    return r;
}
SEXP Rf_getAttrib(SEXP x, SEXP name) { return name == R_DimSymbol ? x->dim : R_NilValue; }
double Rf_asReal(SEXP x)
{
    if (x->len < 1) return R_NaReal;
    if (x->type == REALSXP) return ((double*)x->data)[0];
    if (x->type == INTSXP || x->type == LGLSXP) return ((int*)x->data)[0] == R_NaInt ? R_NaReal : (double)((int*)x->data)[0];
    return R_NaReal;
}
int Rf_asInteger(SEXP x)
{
    if (x->len < 1) return R_NaInt;
    if (x->type == INTSXP || x->type == LGLSXP) return ((int*)x->data)[0];
    if (x->type == REALSXP) return isnan(((double*)x->data)[0]) ? R_NaInt : (int)((double*)x->data)[0];
    return R_NaInt;
}
int Rf_asLogical(SEXP x) { int v = Rf_asInteger(x); return v == R_NaInt ? R_NaInt : v != 0; }
SEXP Rf_ScalarReal(double v) { SEXP s = Rf_allocVector(REALSXP, 1); ((double*)s->data)[0] = v; return s; }
SEXP STRING_ELT(SEXP x, R_xlen_t i)
{
    if (x->type != STRSXP || i < 0 || i >= x->len) Rf_error("mock: STRING_ELT out of range");
    return ((SEXP*)x->data)[i];
}
const char* CHAR(SEXP x) { return (const char*)x->data; }
SEXP VECTOR_ELT(SEXP x, R_xlen_t i)
{
    if (x->type != VECSXP || i < 0 || i >= x->len) Rf_error("mock: VECTOR_ELT out of range");
    return ((SEXP*)x->data)[i];
}
SEXP SET_VECTOR_ELT(SEXP x, R_xlen_t i, SEXP v)
{
    if (x->type != VECSXP || i < 0 || i >= x->len) Rf_error("mock: SET_VECTOR_ELT out of range");
    ((SEXP*)x->data)[i] = v;
    return v;
}
char* R_alloc(size_t n, int size) { return (char*)track(calloc(n ? n : 1, (size_t)size)); }

SEXP mock_int_vector(const int* v, R_xlen_t n)
{
    SEXP s = Rf_allocVector(INTSXP, n);
    if (n) memcpy(s->data, v, (size_t)n * sizeof(int));
    return s;
}
SEXP mock_real_vector(const double* v, R_xlen_t n)
{
    SEXP s = Rf_allocVector(REALSXP, n);
    if (n) memcpy(s->data, v, (size_t)n * sizeof(double));
    return s;
}
SEXP mock_real_matrix(const double* v, int nrow, int ncol)
{
    SEXP s = mock_real_vector(v, (R_xlen_t)nrow * ncol);
    int d[2] = {nrow, ncol};
    s->dim = mock_int_vector(d, 2);
    return s;
}
SEXP mock_string(const char* str)
{
    SEXP ch = Rf_allocVector(CHARSXP, (R_xlen_t)strlen(str));
    memcpy(ch->data, str, strlen(str) + 1);
    SEXP s = Rf_allocVector(STRSXP, 1);
    ((SEXP*)s->data)[0] = ch;
    return s;
}
SEXP mock_logical(int v)
{
    SEXP s = Rf_allocVector(LGLSXP, 1);
    ((int*)s->data)[0] = v;
    return s;
}

/* registration: remember what the package's R_init_<pkg> handed over */
const R_CallMethodDef* mock_registered_calls = NULL;
int mock_dynamic_symbols = -1;
int R_registerRoutines(DllInfo* info, const void* c, const R_CallMethodDef* call, const void* f, const void* e)
{
    (void)c; (void)f; (void)e;
    if (info) info->call = call;
    mock_registered_calls = call;
    return 1;
}
int R_useDynamicSymbols(DllInfo* info, int value)
{
    if (info) info->dynamic = value;
    mock_dynamic_symbols = value;
    return 1;
}

/* ---- external pointers: persistent (a handle lives across .Call invocations), finalized by mock_gc() */
#define MOCK_MAX_EXTPTR 64
static SEXP g_ext[MOCK_MAX_EXTPTR];
static R_CFinalizer_t g_fin[MOCK_MAX_EXTPTR];
SEXP R_MakeExternalPtr(void* p, SEXP tag, SEXP prot)
{
    (void)tag; (void)prot;
    for (int i = 0; i < MOCK_MAX_EXTPTR; i++)
        if (!g_ext[i]) {
            SEXP s = (SEXP)calloc(1, sizeof(struct mock_sexp));
            s->type = EXTPTRSXP; s->len = 1; s->data = p; s->dim = R_NilValue;
            g_ext[i] = s; g_fin[i] = NULL;
            return s;
        }
    Rf_error("mock: too many external pointers");
    return R_NilValue;
}
void* R_ExternalPtrAddr(SEXP s) { return s->type == EXTPTRSXP ? s->data : NULL; }
void R_ClearExternalPtr(SEXP s) { if (s->type == EXTPTRSXP) s->data = NULL; }
void R_RegisterCFinalizerEx(SEXP s, R_CFinalizer_t fun, int onexit)
{
    (void)onexit;
    for (int i = 0; i < MOCK_MAX_EXTPTR; i++)
        if (g_ext[i] == s) g_fin[i] = fun;
}
int mock_gc(void)
{
    int ran = 0;
    for (int i = 0; i < MOCK_MAX_EXTPTR; i++)
        if (g_ext[i]) {
            if (g_fin[i]) { g_fin[i](g_ext[i]); ran++; }
            free(g_ext[i]);
            g_ext[i] = NULL; g_fin[i] = NULL;
        }
    return ran;
}
