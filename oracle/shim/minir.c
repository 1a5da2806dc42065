/* minir.c -- implementation of the R-API stand-in declared in minir.h.
 * TEST INFRASTRUCTURE ONLY (see the header). Written from R's documented C API
 * semantics ("Writing R Extensions"); no code from R or from the reference. */
#include "minir.h"
#include <dlfcn.h>

double R_NaReal, R_NaN, R_PosInf, R_NegInf;
SEXP R_NilValue, R_NamesSymbol, R_LevelsSymbol, R_ClassSymbol, R_DimSymbol,
  R_DimNamesSymbol, R_RowNamesSymbol, R_NaString, R_BlankString;
jmp_buf *minir_error_jmp = NULL;

/* symbols the reference defines in src/globals.c (not compiled: it registers
 * .Call routines with a real R). */
SEXP BN_ModelstringSymbol, BN_NodesSymbol, BN_ProbSymbol, BN_MethodSymbol,
  BN_WeightsSymbol, BN_DsepsetSymbol, BN_MetaDataSymbol, BN_NobsSymbol,
  BN_DfSymbol, BN_CutpointsSymbol, TRUESEXP, FALSESEXP;

static char last_error[1024];
static int n_warnings = 0;
static int initialised = 0;

/* ------------------------------ arena -------------------------------- */

typedef struct chunk { struct chunk *next; } chunk;
static chunk *arena = NULL;

static void *arena_alloc(size_t bytes) {
  chunk *c = (chunk *)calloc(1, sizeof(chunk) + bytes + 16);
  if (!c) { fprintf(stderr, "minir: out of memory\n"); abort(); }
  c->next = arena;
  arena = c;
  return (void *)(c + 1);
}

void minir_arena_reset(void) {
  while (arena) { chunk *n = arena->next; free(arena); arena = n; }
}

/* permanent objects (symbols, nil) are never released. */
static SEXP permanent(int type, int len, size_t bytes) {
  SEXP s = (SEXP)calloc(1, sizeof(struct minir_sexp));
  s->type = type; s->len = len;
  s->data = bytes ? calloc(1, bytes) : NULL;
  return s;
}

static size_t elsize(int type) {
  switch (type) {
    case LGLSXP: case INTSXP: return sizeof(int);
    case REALSXP: return sizeof(double);
    case STRSXP: case VECSXP: return sizeof(SEXP);
    default: return 0;
  }
}

SEXP Rf_install(const char *s) {
  /* tiny symbol table: linear search is fine for a dozen symbols. */
  static SEXP table[64]; static int ntable = 0;
  for (int i = 0; i < ntable; i++)
    if (!strcmp((char *)table[i]->data, s)) return table[i];
  SEXP sym = permanent(SYMSXP, 1, strlen(s) + 1);
  strcpy((char *)sym->data, s);
  if (ntable < 64) table[ntable++] = sym;
  return sym;
}

void minir_init(void) {
  if (initialised) return;
  initialised = 1;
  union { double d; unsigned long long u; } na;
  na.u = 0x7FF00000000007A2ULL;     /* R's NA_real_: a NaN with payload 1954. */
  R_NaReal = na.d;
  R_NaN = NAN; R_PosInf = INFINITY; R_NegInf = -INFINITY;
  R_NilValue = permanent(NILSXP, 0, 0);
  R_NamesSymbol = Rf_install("names");
  R_LevelsSymbol = Rf_install("levels");
  R_ClassSymbol = Rf_install("class");
  R_DimSymbol = Rf_install("dim");
  R_DimNamesSymbol = Rf_install("dimnames");
  R_RowNamesSymbol = Rf_install("row.names");
  R_NaString = permanent(CHARSXP, 2, 3); strcpy((char *)R_NaString->data, "NA");
  R_BlankString = permanent(CHARSXP, 0, 1);
  BN_ModelstringSymbol = Rf_install("modelstring");
  BN_NodesSymbol = Rf_install("nodes");
  BN_ProbSymbol = Rf_install("prob");
  BN_MethodSymbol = Rf_install("method");
  BN_WeightsSymbol = Rf_install("weights");
  BN_DsepsetSymbol = Rf_install("dsep.set");
  BN_MetaDataSymbol = Rf_install("metadata");
  BN_NobsSymbol = Rf_install("nobs");
  BN_DfSymbol = Rf_install("df");
  BN_CutpointsSymbol = Rf_install("cutpoints");
  TRUESEXP = permanent(LGLSXP, 1, sizeof(int)); ((int *)TRUESEXP->data)[0] = 1;
  FALSESEXP = permanent(LGLSXP, 1, sizeof(int)); ((int *)FALSESEXP->data)[0] = 0;
}

__attribute__((constructor)) static void minir_ctor(void) { minir_init(); }

int R_IsNA(double x) {
  union { double d; unsigned long long u; } v; v.d = x;
  return isnan(x) && ((v.u & 0xFFFFFFFFULL) == 1954);
}

/* ------------------------- vectors and accessors --------------------- */

SEXP Rf_allocVector(SEXPTYPE type, R_xlen_t n) {
  SEXP s = (SEXP)arena_alloc(sizeof(struct minir_sexp));
  s->type = type; s->len = (int)n; s->named = 0; s->attr = NULL;
  size_t es = elsize(type);
  s->data = (n > 0 && es) ? arena_alloc(es * (size_t)n) : NULL;
  if (type == STRSXP)
    for (int i = 0; i < n; i++) ((SEXP *)s->data)[i] = R_BlankString;
  if (type == VECSXP)
    for (int i = 0; i < n; i++) ((SEXP *)s->data)[i] = R_NilValue;
  return s;
}

/* a vector whose payload lives in caller-owned memory (no copy): used by the
 * Python harness to hand numpy columns to the reference. */
SEXP minir_vector_view(int type, void *data, int n) {
  SEXP s = (SEXP)arena_alloc(sizeof(struct minir_sexp));
  s->type = type; s->len = n; s->named = 2; s->attr = NULL; s->data = data;
  return s;
}

SEXP Rf_allocMatrix(SEXPTYPE type, int nr, int nc) {
  SEXP s = Rf_allocVector(type, (R_xlen_t)nr * nc);
  SEXP dim = Rf_allocVector(INTSXP, 2);
  INTEGER(dim)[0] = nr; INTEGER(dim)[1] = nc;
  Rf_setAttrib(s, R_DimSymbol, dim);
  return s;
}

int Rf_length(SEXP x) { return (x == NULL || x->type == NILSXP) ? 0 : x->len; }
int TYPEOF(SEXP x) { return x->type; }
int NAMED(SEXP x) { return x->named; }
int *INTEGER(SEXP x) { return (int *)x->data; }
int *LOGICAL(SEXP x) { return (int *)x->data; }
double *REAL(SEXP x) { return (double *)x->data; }
SEXP STRING_ELT(SEXP x, R_xlen_t i) { return ((SEXP *)x->data)[i]; }
void SET_STRING_ELT(SEXP x, R_xlen_t i, SEXP v) { ((SEXP *)x->data)[i] = v; }
SEXP VECTOR_ELT(SEXP x, R_xlen_t i) { return ((SEXP *)x->data)[i]; }
SEXP SET_VECTOR_ELT(SEXP x, R_xlen_t i, SEXP v) { ((SEXP *)x->data)[i] = v; return v; }
const char *CHAR(SEXP x) { return (const char *)x->data; }

SEXP Rf_mkChar(const char *str) {
  size_t n = strlen(str);
  SEXP s = (SEXP)arena_alloc(sizeof(struct minir_sexp));
  s->type = CHARSXP; s->len = (int)n;
  s->data = arena_alloc(n + 1);
  memcpy(s->data, str, n + 1);
  return s;
}

SEXP Rf_mkString(const char *str) {
  SEXP s = Rf_allocVector(STRSXP, 1);
  SET_STRING_ELT(s, 0, Rf_mkChar(str));
  return s;
}

SEXP Rf_ScalarReal(double x) { SEXP s = Rf_allocVector(REALSXP, 1); REAL(s)[0] = x; return s; }
SEXP Rf_ScalarInteger(int x) { SEXP s = Rf_allocVector(INTSXP, 1); INTEGER(s)[0] = x; return s; }
SEXP Rf_ScalarLogical(int x) { SEXP s = Rf_allocVector(LGLSXP, 1); LOGICAL(s)[0] = x; return s; }

SEXP Rf_getAttrib(SEXP x, SEXP name) {
  if (!x) return R_NilValue;
  for (minir_attr *a = x->attr; a; a = a->next)
    if (a->tag == name) return a->value;
  return R_NilValue;
}

SEXP Rf_setAttrib(SEXP x, SEXP name, SEXP val) {
  for (minir_attr *a = x->attr; a; a = a->next)
    if (a->tag == name) { a->value = val; return val; }
  minir_attr *a = (minir_attr *)arena_alloc(sizeof(minir_attr));
  a->tag = name; a->value = val; a->next = x->attr; x->attr = a;
  return val;
}

SEXP Rf_duplicate(SEXP x) {
  if (x == R_NilValue || x->type == SYMSXP || x->type == CHARSXP) return x;
  SEXP s = Rf_allocVector((SEXPTYPE)x->type, x->len);
  if (x->type == VECSXP) {
    for (int i = 0; i < x->len; i++)
      SET_VECTOR_ELT(s, i, Rf_duplicate(VECTOR_ELT(x, i)));
  } else if (x->len > 0) {
    memcpy(s->data, x->data, elsize(x->type) * (size_t)x->len);
  }
  for (minir_attr *a = x->attr; a; a = a->next)
    Rf_setAttrib(s, a->tag, Rf_duplicate(a->value));
  return s;
}

SEXP Rf_coerceVector(SEXP x, SEXPTYPE type) {
  if ((int)type == x->type) return x;
  SEXP s = Rf_allocVector(type, x->len);
  char buf[64];
  for (int i = 0; i < x->len; i++) {
    if (x->type == INTSXP && type == STRSXP) {
      if (INTEGER(x)[i] == NA_INTEGER) SET_STRING_ELT(s, i, R_NaString);
      else { snprintf(buf, sizeof buf, "%d", INTEGER(x)[i]); SET_STRING_ELT(s, i, Rf_mkChar(buf)); }
    } else if (x->type == INTSXP && type == REALSXP) {
      REAL(s)[i] = INTEGER(x)[i] == NA_INTEGER ? NA_REAL : (double)INTEGER(x)[i];
    } else if (x->type == REALSXP && type == INTSXP) {
      INTEGER(s)[i] = isnan(REAL(x)[i]) ? NA_INTEGER : (int)REAL(x)[i];
    } else if (x->type == LGLSXP && type == INTSXP) {
      INTEGER(s)[i] = LOGICAL(x)[i];
    } else {
      Rf_error("minir: unsupported coercion %d -> %d", x->type, (int)type);
    }
  }
  return s;
}

int Rf_asInteger(SEXP x) {
  if (x->len < 1) return NA_INTEGER;
  if (x->type == INTSXP || x->type == LGLSXP) return INTEGER(x)[0];
  if (x->type == REALSXP) return isnan(REAL(x)[0]) ? NA_INTEGER : (int)REAL(x)[0];
  return NA_INTEGER;
}

double Rf_asReal(SEXP x) {
  if (x->len < 1) return NA_REAL;
  if (x->type == REALSXP) return REAL(x)[0];
  if (x->type == INTSXP || x->type == LGLSXP)
    return INTEGER(x)[0] == NA_INTEGER ? NA_REAL : (double)INTEGER(x)[0];
  return NA_REAL;
}

int Rf_isNull(SEXP x) { return x == NULL || x->type == NILSXP; }
int Rf_isInteger(SEXP x) { return x->type == INTSXP && !Rf_inherits(x, "factor"); }
int Rf_isReal(SEXP x) { return x->type == REALSXP; }
int Rf_isString(SEXP x) { return x->type == STRSXP; }
int Rf_isLogical(SEXP x) { return x->type == LGLSXP; }
int Rf_isFactor(SEXP x) { return x->type == INTSXP && Rf_inherits(x, "factor"); }

int Rf_inherits(SEXP x, const char *what) {
  SEXP k = Rf_getAttrib(x, R_ClassSymbol);
  for (int i = 0; i < Rf_length(k); i++)
    if (!strcmp(CHAR(STRING_ELT(k, i)), what)) return 1;
  return 0;
}

/* -------------------- match() and duplicated() ----------------------- */

/* open-addressing hash on int keys -> first position (1-based). */
typedef struct { int *key; int *pos; unsigned cap; } ihash;

static unsigned ihash_slot(unsigned cap, int k) {
  unsigned h = (unsigned)k * 2654435761u;
  return h & (cap - 1);
}

static ihash ihash_new(int n) {
  ihash h; h.cap = 16;
  while (h.cap < (unsigned)n * 2u + 1u) h.cap <<= 1;
  h.key = (int *)malloc(h.cap * sizeof(int));
  h.pos = (int *)calloc(h.cap, sizeof(int));
  return h;
}

/* returns the stored position for k, inserting `pos` if absent. */
static int ihash_get_or_put(ihash *h, int k, int pos, int insert) {
  unsigned s = ihash_slot(h->cap, k);
  while (h->pos[s] != 0) {
    if (h->key[s] == k) return h->pos[s];
    s = (s + 1) & (h->cap - 1);
  }
  if (insert) { h->key[s] = k; h->pos[s] = pos; }
  return 0;
}

static void ihash_free(ihash *h) { free(h->key); free(h->pos); }

/* match(x, table): for each element of x the 1-based index of its first
 * occurrence in table, or nomatch. NA matches NA, as in R. */
SEXP Rf_match(SEXP table, SEXP x, int nomatch) {
  int nx = Rf_length(x), nt = Rf_length(table);
  SEXP res = Rf_allocVector(INTSXP, nx);
  int *r = INTEGER(res);
  if ((x->type == INTSXP || x->type == LGLSXP) &&
      (table->type == INTSXP || table->type == LGLSXP)) {
    ihash h = ihash_new(nt);
    int *t = INTEGER(table), *v = INTEGER(x);
    for (int i = 0; i < nt; i++) ihash_get_or_put(&h, t[i], i + 1, 1);
    for (int i = 0; i < nx; i++) {
      int p = ihash_get_or_put(&h, v[i], 0, 0);
      r[i] = p ? p : nomatch;
    }
    ihash_free(&h);
  } else if (x->type == STRSXP && table->type == STRSXP) {
    for (int i = 0; i < nx; i++) {
      const char *s = CHAR(STRING_ELT(x, i));
      r[i] = nomatch;
      for (int j = 0; j < nt; j++)
        if (!strcmp(s, CHAR(STRING_ELT(table, j)))) { r[i] = j + 1; break; }
    }
  } else if (nx == 0) {
    /* nothing to match. */
  } else {
    Rf_error("minir: match() on unsupported types %d, %d", x->type, table->type);
  }
  return res;
}

SEXP Rf_duplicated(SEXP x, Rboolean from_last) {
  int n = Rf_length(x);
  SEXP res = Rf_allocVector(LGLSXP, n);
  int *r = LOGICAL(res);
  if (x->type == INTSXP || x->type == LGLSXP) {
    ihash h = ihash_new(n);
    int *v = INTEGER(x);
    if (!from_last)
      for (int i = 0; i < n; i++) r[i] = ihash_get_or_put(&h, v[i], i + 1, 1) != 0;
    else
      for (int i = n - 1; i >= 0; i--) r[i] = ihash_get_or_put(&h, v[i], i + 1, 1) != 0;
    ihash_free(&h);
  } else if (x->type == STRSXP) {
    for (int i = 0; i < n; i++) {
      r[i] = 0;
      if (!from_last) {
        for (int j = 0; j < i; j++)
          if (!strcmp(CHAR(STRING_ELT(x, i)), CHAR(STRING_ELT(x, j)))) { r[i] = 1; break; }
      } else {
        for (int j = n - 1; j > i; j--)
          if (!strcmp(CHAR(STRING_ELT(x, i)), CHAR(STRING_ELT(x, j)))) { r[i] = 1; break; }
      }
    }
  } else if (n > 0) {
    Rf_error("minir: duplicated() on unsupported type %d", x->type);
  }
  return res;
}

/* ---------------------- errors, warnings, printing ------------------- */

void Rf_error(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  vsnprintf(last_error, sizeof last_error, fmt, ap);
  va_end(ap);
  if (minir_error_jmp) longjmp(*minir_error_jmp, 1);
  fprintf(stderr, "minir: uncaught error(): %s\n", last_error);
  abort();
}

void Rf_warning(const char *fmt, ...) { (void)fmt; n_warnings++; }
int minir_warning_count(void) { return n_warnings; }
void minir_warning_reset(void) { n_warnings = 0; }
const char *minir_last_error(void) { return last_error; }

static int print_enabled = 0;
void minir_set_print(int on) { print_enabled = on; }
void Rprintf(const char *fmt, ...) {
  if (!print_enabled) return;
  va_list ap; va_start(ap, fmt); vprintf(fmt, ap); va_end(ap);
}
void REprintf(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap);
}
void R_CheckUserInterrupt(void) {}

/* run fn(arg) with error() trapped; returns 0 on success, 1 if error() fired. */
int minir_try(void (*fn)(void *), void *arg) {
  jmp_buf env; jmp_buf *saved = minir_error_jmp;
  minir_error_jmp = &env;
  int failed = setjmp(env);
  if (!failed) fn(arg);
  minir_error_jmp = saved;
  return failed;
}

/* --------------------------- small Rmath bits ------------------------ */

int Rf_imax2(int x, int y) { return x > y ? x : y; }
int Rf_imin2(int x, int y) { return x < y ? x : y; }
double Rf_fmax2(double x, double y) { return (isnan(x) || isnan(y)) ? x + y : (x > y ? x : y); }
double Rf_fmin2(double x, double y) { return (isnan(x) || isnan(y)) ? x + y : (x < y ? x : y); }
void GetRNGstate(void) {}
void PutRNGstate(void) {}
/* R's Mersenne-Twister stream is not available here. The permutation tests only
 * need SOME uniform stream that the engine under test can be handed as well:
 * splitmix64, top 53 bits, zero rejected (unif_rand() never returns 0 or 1).
 * Unseeded use stays an error. */
static unsigned long long minir_rng_state = 0;
static int minir_rng_seeded = 0;
void minir_rng_seed(unsigned long long seed) { minir_rng_state = seed; minir_rng_seeded = 1; }
unsigned long long minir_rng_peek(void) { return minir_rng_state; }
double unif_rand(void) {
  if (!minir_rng_seeded) Rf_error("minir: unif_rand() before minir_rng_seed()");
  for (;;) {
    unsigned long long z = (minir_rng_state += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    const double u = (double)(z >> 11) * 0x1.0p-53;
    if (u > 0.0) return u;
  }
}

/* ------------------------------ LAPACK -------------------------------- */
/* resolved lazily from scipy's bundled OpenBLAS, which the Python harness
 * loads with RTLD_GLOBAL before this library (symbols carry a scipy_ prefix). */

static void *lapack_sym(const char *name) {
  void *p = dlsym(RTLD_DEFAULT, name);
  if (!p) Rf_error("minir: LAPACK symbol %s not loaded (load scipy's OpenBLAS with RTLD_GLOBAL first)", name);
  return p;
}

void minir_f77_dgesdd(const char *jobz, const int *m, const int *n, double *a,
    const int *lda, double *s, double *u, const int *ldu, double *vt,
    const int *ldvt, double *work, const int *lwork, int *iwork, int *info) {
  typedef void (*fn_t)(const char *, const int *, const int *, double *,
      const int *, double *, double *, const int *, double *, const int *,
      double *, const int *, int *, int *, size_t);
  static fn_t fn = NULL;
  if (!fn) fn = (fn_t)lapack_sym("scipy_dgesdd_");
  fn(jobz, m, n, a, lda, s, u, ldu, vt, ldvt, work, lwork, iwork, info, 1);
}

void minir_f77_dgetrf(const int *m, const int *n, double *a, const int *lda,
    int *ipiv, int *info) {
  typedef void (*fn_t)(const int *, const int *, double *, const int *, int *, int *);
  static fn_t fn = NULL;
  if (!fn) fn = (fn_t)lapack_sym("scipy_dgetrf_");
  fn(m, n, a, lda, ipiv, info);
}

void minir_f77_dgemm(const char *transa, const char *transb, const int *m,
    const int *n, const int *k, const double *alpha, const double *a,
    const int *lda, const double *b, const int *ldb, const double *beta,
    double *c, const int *ldc) {
  typedef void (*fn_t)(const char *, const char *, const int *, const int *,
      const int *, const double *, const double *, const int *, const double *,
      const int *, const double *, double *, const int *, size_t, size_t);
  static fn_t fn = NULL;
  if (!fn) fn = (fn_t)lapack_sym("scipy_dgemm_");
  fn(transa, transb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, 1, 1);
}

/* R_qsort_I(v, I, i, j): sort v[i-1 .. j-1] increasingly and permute I alongside
 * (R's sort.c / qsort-body.c, Singleton's CACM 347 variant; not in the tree).
 * This stand-in is a bottom-up merge sort, so the order of EQUAL keys is by
 * original position and may differ from R's unstable quicksort: the only caller
 * on the compiled path is chow_liu()'s d_sort over the mutual-information
 * coefficients (src/learning/local/mi.matrix.c:499), and the fixtures use data
 * without tied coefficients. */
void R_qsort_I(double *v, int *I, int i, int j) {
  int n = j - i + 1;
  if (n < 2) return;
  double *a = v + (i - 1), *tv = (double *)malloc(sizeof(double) * (size_t)n);
  int *b = I + (i - 1), *ti = (int *)malloc(sizeof(int) * (size_t)n);
  for (int w = 1; w < n; w *= 2) {
    for (int lo = 0; lo < n; lo += 2 * w) {
      int mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int p = lo, q = mid, o = lo;
      while (p < mid && q < hi) {
        if (a[q] < a[p]) { tv[o] = a[q]; ti[o++] = b[q++]; }
        else { tv[o] = a[p]; ti[o++] = b[p++]; }
      }
      while (p < mid) { tv[o] = a[p]; ti[o++] = b[p++]; }
      while (q < hi) { tv[o] = a[q]; ti[o++] = b[q++]; }
    }
    memcpy(a, tv, sizeof(double) * (size_t)n);
    memcpy(b, ti, sizeof(int) * (size_t)n);
  }
  free(tv); free(ti);
}
