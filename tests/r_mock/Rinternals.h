/* Rinternals.h -- MOCK of the slice of R's C API that rpkg/src/r_shim.c uses.
 *
 * Test infrastructure only (R is not installed where this repository is built): it lets the test-suite compile
 * the .Call shim unmodified and drive its entry points from C, so the shim is at least executed against the
 * real libsignals_b200.so.  Semantics follow "Writing R Extensions" for the functions below; nothing here is
 * taken from R's sources, and it is no substitute for a build inside R. */
#ifndef MOCK_RINTERNALS_H
#define MOCK_RINTERNALS_H
#include <stddef.h>
#include <math.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef ptrdiff_t R_xlen_t;
typedef struct mock_sexp* SEXP;
struct mock_sexp {
    int type;
    R_xlen_t len;
    void* data;        /* int[] / double[] / SEXP[] / char[] */
    SEXP dim;          /* the "dim" attribute or R_NilValue */
};

enum { NILSXP = 0, CHARSXP = 9, LGLSXP = 10, INTSXP = 13, REALSXP = 14, STRSXP = 16, VECSXP = 19, EXTPTRSXP = 22 };

extern SEXP R_NilValue;
extern SEXP R_DimSymbol;
extern double R_NaReal;
extern int R_NaInt;
#define NA_REAL R_NaReal
#define NA_INTEGER R_NaInt
#define ISNAN(x) (isnan(x) != 0)
#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif

int TYPEOF(SEXP x);
R_xlen_t XLENGTH(SEXP x);
int Rf_length(SEXP x);
int* INTEGER(SEXP x);
double* REAL(SEXP x);
SEXP Rf_allocVector(int type, R_xlen_t n);
SEXP Rf_coerceVector(SEXP x, int type);
SEXP Rf_getAttrib(SEXP x, SEXP name);
double Rf_asReal(SEXP x);
int Rf_asInteger(SEXP x);
int Rf_asLogical(SEXP x);
SEXP Rf_ScalarReal(double v);
SEXP STRING_ELT(SEXP x, R_xlen_t i);
const char* CHAR(SEXP x);
SEXP VECTOR_ELT(SEXP x, R_xlen_t i);
SEXP SET_VECTOR_ELT(SEXP x, R_xlen_t i, SEXP v);
void Rf_error(const char* fmt, ...) __attribute__((noreturn, format(printf, 1, 2)));
char* R_alloc(size_t n, int size);
/* external pointers: objects that outlive one .Call (not released by mock_reset); finalizers run in mock_gc() */
typedef void (*R_CFinalizer_t)(SEXP);
SEXP R_MakeExternalPtr(void* p, SEXP tag, SEXP prot);
void* R_ExternalPtrAddr(SEXP s);
void R_ClearExternalPtr(SEXP s);
void R_RegisterCFinalizerEx(SEXP s, R_CFinalizer_t fun, int onexit);
#define PROTECT(x) (x)
#define UNPROTECT(n) ((void)(n))

/* mock-only helpers (used by the test driver, not by the shim) */
SEXP mock_int_vector(const int* v, R_xlen_t n);
SEXP mock_real_vector(const double* v, R_xlen_t n);
SEXP mock_real_matrix(const double* v, int nrow, int ncol);
SEXP mock_string(const char* s);
SEXP mock_logical(int v);
void mock_reset(void);                 /* frees everything allocated since the last reset */
int mock_gc(void);                     /* runs the finalizers of all external pointers and frees them; returns how many ran */

#ifdef __cplusplus
}
#endif
#endif
