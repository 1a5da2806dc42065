/* minir.h -- a tiny stand-in for the parts of R's C API that the reference's
 * hot-path files touch. TEST INFRASTRUCTURE ONLY: it exists so that the
 * reference's own C files under /root/reference/src can be compiled, unmodified
 * and in place, into oracle/_ref/libbnref.so (see oracle/Makefile). Nothing in
 * the product (bnlearn_b200/, include/) includes or links this.
 *
 * It is NOT R: vectors are malloc'ed from an arena that is released by
 * minir_arena_reset(), PROTECT/UNPROTECT are no-ops, error() longjmps to the
 * handler armed by minir_try() (or aborts), and only the vector types the
 * reference uses on this path exist. */
#ifndef MINIR_H
#define MINIR_H

#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <float.h>
#include <limits.h>
#include <stdarg.h>
#include <stdbool.h>
#include <setjmp.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  NILSXP = 0, SYMSXP = 1, LISTSXP = 2, CLOSXP = 3, ENVSXP = 4, CHARSXP = 9,
  LGLSXP = 10, INTSXP = 13, REALSXP = 14, STRSXP = 16, VECSXP = 19
} SEXPTYPE;

typedef struct minir_sexp *SEXP;
typedef struct minir_attr { SEXP tag; SEXP value; struct minir_attr *next; } minir_attr;
struct minir_sexp {
  int type;
  int len;
  int named;
  void *data;          /* int*, double*, SEXP* or char* */
  minir_attr *attr;
};

typedef int Rboolean;
#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif
typedef ptrdiff_t R_xlen_t;
typedef int R_len_t;

/* special values. */
#define NA_INTEGER INT_MIN
#define NA_LOGICAL INT_MIN
extern double R_NaReal, R_NaN, R_PosInf, R_NegInf;
#define NA_REAL R_NaReal
#define ISNAN(x) (isnan(x) != 0)
#define ISNA(x) R_IsNA(x)
#define R_FINITE(x) isfinite(x)
int R_IsNA(double x);

extern SEXP R_NilValue, R_NamesSymbol, R_LevelsSymbol, R_ClassSymbol,
  R_DimSymbol, R_DimNamesSymbol, R_RowNamesSymbol, R_NaString, R_BlankString;

/* allocation and access. */
SEXP Rf_allocVector(SEXPTYPE type, R_xlen_t n);
SEXP Rf_allocMatrix(SEXPTYPE type, int nr, int nc);
int Rf_length(SEXP x);
#define LENGTH(x) Rf_length(x)
#define XLENGTH(x) ((R_xlen_t)Rf_length(x))
#define xlength(x) ((R_xlen_t)Rf_length(x))
int TYPEOF(SEXP x);
int *INTEGER(SEXP x);
int *LOGICAL(SEXP x);
double *REAL(SEXP x);
SEXP STRING_ELT(SEXP x, R_xlen_t i);
void SET_STRING_ELT(SEXP x, R_xlen_t i, SEXP v);
SEXP VECTOR_ELT(SEXP x, R_xlen_t i);
SEXP SET_VECTOR_ELT(SEXP x, R_xlen_t i, SEXP v);
const char *CHAR(SEXP x);
SEXP Rf_mkChar(const char *s);
SEXP Rf_mkString(const char *s);
SEXP Rf_install(const char *s);
SEXP Rf_ScalarReal(double x);
SEXP Rf_ScalarInteger(int x);
SEXP Rf_ScalarLogical(int x);
SEXP Rf_getAttrib(SEXP x, SEXP name);
SEXP Rf_setAttrib(SEXP x, SEXP name, SEXP val);
SEXP Rf_duplicate(SEXP x);
SEXP Rf_coerceVector(SEXP x, SEXPTYPE type);
SEXP Rf_match(SEXP table, SEXP x, int nomatch);
SEXP Rf_duplicated(SEXP x, Rboolean from_last);
int Rf_asInteger(SEXP x);
double Rf_asReal(SEXP x);
int Rf_isNull(SEXP x);
int Rf_isInteger(SEXP x);
int Rf_isReal(SEXP x);
int Rf_isString(SEXP x);
int Rf_isFactor(SEXP x);
int Rf_isLogical(SEXP x);
int Rf_inherits(SEXP x, const char *what);
int NAMED(SEXP x);
#define MAYBE_REFERENCED(x) (NAMED(x) > 0)

#define PROTECT(x) (x)
#define UNPROTECT(n) ((void)(n))

/* un-prefixed names, as R's Rinternals.h provides without R_NO_REMAP. */
#define allocVector Rf_allocVector
#define allocMatrix Rf_allocMatrix
#define length Rf_length
#define mkChar Rf_mkChar
#define mkString Rf_mkString
#define install Rf_install
#define ScalarReal Rf_ScalarReal
#define ScalarInteger Rf_ScalarInteger
#define ScalarLogical Rf_ScalarLogical
#define getAttrib Rf_getAttrib
#define setAttrib Rf_setAttrib
#define duplicate Rf_duplicate
#define coerceVector Rf_coerceVector
#define match Rf_match
#define duplicated Rf_duplicated
#define asInteger Rf_asInteger
#define asReal Rf_asReal
#define isNull Rf_isNull
#define isInteger Rf_isInteger
#define isReal Rf_isReal
#define isString Rf_isString
#define isFactor Rf_isFactor
#define isLogical Rf_isLogical
#define inherits Rf_inherits

/* errors, warnings, printing. */
void Rf_error(const char *fmt, ...) __attribute__((noreturn));
void Rf_warning(const char *fmt, ...);
void Rprintf(const char *fmt, ...);
void REprintf(const char *fmt, ...);
#define error Rf_error
#define warning Rf_warning
void R_CheckUserInterrupt(void);

/* arena + error trapping, driven by the test harness. */
void minir_arena_reset(void);
int minir_warning_count(void);
void minir_warning_reset(void);
const char *minir_last_error(void);
extern jmp_buf *minir_error_jmp;
void minir_init(void);

/* Rmath subset: restated in oracle/rmath_port.c (R's nmath is not in the
 * reference tree). */
double Rf_lgammafn(double x);
double Rf_gammafn(double x);
double Rf_pchisq(double x, double df, int lower_tail, int log_p);
double Rf_pgamma(double x, double alph, double scale, int lower_tail, int log_p);
double Rf_pt(double x, double n, int lower_tail, int log_p);
double Rf_pnorm5(double x, double mu, double sigma, int lower_tail, int log_p);
double Rf_dnorm4(double x, double mu, double sigma, int give_log);
double Rf_pbeta(double x, double a, double b, int lower_tail, int log_p);
double Rf_choose(double n, double k);
double Rf_lchoose(double n, double k);
int Rf_imax2(int x, int y);
int Rf_imin2(int x, int y);
double Rf_fmax2(double x, double y);
double Rf_fmin2(double x, double y);
double Rf_fround(double x, double digits);
double unif_rand(void);
void minir_rng_seed(unsigned long long seed);      /* harness only */
unsigned long long minir_rng_peek(void);
void GetRNGstate(void);
void PutRNGstate(void);
#define lgammafn Rf_lgammafn
#define gammafn Rf_gammafn
#define pchisq Rf_pchisq
#define pgamma Rf_pgamma
#define pt Rf_pt
#define pnorm Rf_pnorm5
#define dnorm Rf_dnorm4
#define pbeta Rf_pbeta
#define choose Rf_choose
#define lchoose Rf_lchoose
#define imax2 Rf_imax2
#define imin2 Rf_imin2
#define fmax2 Rf_fmax2
#define fmin2 Rf_fmin2
#define fround Rf_fround
#ifndef M_LN_SQRT_PI
#define M_LN_SQRT_PI 0.572364942924700087071713675677   /* log(sqrt(pi)) */
#endif
#ifndef M_LN_SQRT_2PI
#define M_LN_SQRT_2PI 0.918938533204672741780329736406
#endif
#ifndef M_LN2
#define M_LN2 0.693147180559945309417232121458
#endif

/* Fortran linkage: LAPACK comes from scipy's bundled OpenBLAS (symbols are
 * prefixed scipy_), LINPACK dqrdc2/dqrsl from oracle/linpack_port.c. */
#define F77_NAME(x) minir_f77_##x
#define F77_CALL(x) minir_f77_##x
#define F77_SUB(x) minir_f77_##x
#define FCONE
#define FCLEN
#define La_INT int
void minir_f77_dgesdd(const char *jobz, const int *m, const int *n, double *a,
    const int *lda, double *s, double *u, const int *ldu, double *vt,
    const int *ldvt, double *work, const int *lwork, int *iwork, int *info);
void minir_f77_dgetrf(const int *m, const int *n, double *a, const int *lda,
    int *ipiv, int *info);
void minir_f77_dgemm(const char *transa, const char *transb, const int *m,
    const int *n, const int *k, const double *alpha, const double *a,
    const int *lda, const double *b, const int *ldb, const double *beta,
    double *c, const int *ldc);
void minir_f77_dqrdc2(double *x, int *ldx, int *n, int *p, double *tol,
    int *rank, double *qraux, int *pivot, double *work);
void minir_f77_dqrsl(double *x, int *ldx, int *n, int *k, double *qraux,
    double *y, double *qy, double *qty, double *b, double *rsd, double *xb,
    int *job, int *info);

#ifdef __cplusplus
}
#endif

#endif
