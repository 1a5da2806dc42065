/* rmath_port.c -- CPU restatement of the R nmath entry points the reference's
 * hot path calls. TEST INFRASTRUCTURE ONLY (oracle side; never linked into the
 * product library).
 *
 * R's nmath is a dependency of the reference that is NOT in /root/reference
 * (DESCRIPTION:7 "Depends: R (>= 4.4.0)"; pulled in by src/include/rcore.h:4).
 * Call sites on the path:
 *   lgammafn : src/scores/dirichlet.posterior.c:64-65, 140-142
 *   pchisq   : src/tests/patterns/ctest.c:31, utest.c:39, allsubs.test.c:88,224
 *   pt       : src/tests/patterns/ctest.c:107, utest.c:101, allsubs.test.c:216
 *   pnorm    : src/tests/patterns/ctest.c:131, utest.c:121, allsubs.test.c:243
 *   dnorm    : src/scores/gaussian.loglikelihood.c:37, 69
 *
 * PARITY UNPINNED: the reference ships no golden vectors for these functions.
 * They are restated here from their mathematical definitions (regularised
 * incomplete gamma / beta functions by series and continued fractions, in x87
 * extended precision so that they are more accurate than, and independent of,
 * the fp64 device versions in bnlearn_b200/csrc/bnl_special.cuh) and pinned
 * against mpmath-generated fixtures in tests/golden/special_functions.json. */
#include "shim/minir.h"

typedef long double ld;

/* log|Gamma(x)|: platform libm (glibc's lgamma is accurate to ~1ulp for x>0). */
double Rf_lgammafn(double x) {
  int sign = 0;
  if (isnan(x)) return x;
  return lgamma_r(x, &sign);
}

double Rf_gammafn(double x) { return tgamma(x); }

/* ---- regularised incomplete gamma: P(a,x) lower, Q(a,x) upper ---- */

/* log of x^a e^-x / Gamma(a+1). */
static ld log_gamma_prefactor(ld a, ld x) {
  return a * logl(x) - x - lgammal(a + 1.0L);
}

/* lower series: P(a,x) = pref * sum_{n>=0} x^n / ((a+1)...(a+n)). */
static ld gamma_series_p(ld a, ld x) {
  ld term = 1.0L, sum = 1.0L, ap = a;
  for (int n = 1; n < 200000; n++) {
    ap += 1.0L;
    term *= x / ap;
    sum += term;
    if (fabsl(term) < fabsl(sum) * 1e-21L) break;
  }
  return expl(log_gamma_prefactor(a, x)) * sum;
}

/* upper continued fraction (modified Lentz): Q(a,x). */
static ld gamma_cf_q(ld a, ld x) {
  const ld tiny = 1e-4000L;
  ld b = x + 1.0L - a, c = 1.0L / tiny, d = 1.0L / b, h = d;
  for (int i = 1; i < 200000; i++) {
    ld an = -(ld)i * ((ld)i - a);
    b += 2.0L;
    d = an * d + b; if (fabsl(d) < tiny) d = tiny;
    c = b + an / c; if (fabsl(c) < tiny) c = tiny;
    d = 1.0L / d;
    ld del = d * c;
    h *= del;
    if (fabsl(del - 1.0L) < 1e-20L) break;
  }
  /* x^a e^-x / Gamma(a) = pref * a. */
  return expl(log_gamma_prefactor(a, x)) * a * h;
}

double Rf_pgamma(double x, double alph, double scale, int lower_tail, int log_p) {
  if (isnan(x) || isnan(alph) || isnan(scale)) return x + alph + scale;
  if (alph < 0 || scale <= 0) return R_NaN;
  ld p, q;
  x /= scale;
  if (isnan(x)) return x;
  if (alph == 0.0) { p = (x <= 0) ? 0.0L : 1.0L; q = 1.0L - p; }
  else if (x <= 0) { p = 0.0L; q = 1.0L; }
  else if (isinf(x)) { p = 1.0L; q = 0.0L; }
  else if ((ld)x < (ld)alph + 1.0L) { p = gamma_series_p(alph, x); q = 1.0L - p; }
  else { q = gamma_cf_q(alph, x); p = 1.0L - q; }
  ld r = lower_tail ? p : q;
  return log_p ? (double)logl(r) : (double)r;
}

double Rf_pchisq(double x, double df, int lower_tail, int log_p) {
  return Rf_pgamma(x, df / 2.0, 2.0, lower_tail, log_p);
}

/* ---- regularised incomplete beta I_x(a,b), given x and y = 1 - x ---- */

static ld beta_cf(ld a, ld b, ld x) {
  const ld tiny = 1e-4000L;
  ld qab = a + b, qap = a + 1.0L, qam = a - 1.0L;
  ld c = 1.0L, d = 1.0L - qab * x / qap;
  if (fabsl(d) < tiny) d = tiny;
  d = 1.0L / d;
  ld h = d;
  for (int m = 1; m < 2000000; m++) {
    ld m2 = 2.0L * m;
    ld aa = m * (b - m) * x / ((qam + m2) * (a + m2));
    d = 1.0L + aa * d; if (fabsl(d) < tiny) d = tiny;
    c = 1.0L + aa / c; if (fabsl(c) < tiny) c = tiny;
    d = 1.0L / d;
    h *= d * c;
    aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
    d = 1.0L + aa * d; if (fabsl(d) < tiny) d = tiny;
    c = 1.0L + aa / c; if (fabsl(c) < tiny) c = tiny;
    d = 1.0L / d;
    ld del = d * c;
    h *= del;
    if (fabsl(del - 1.0L) < 1e-20L) break;
  }
  return h;
}

/* Stirling remainder S(z) = lgamma(z) - [(z - 1/2) log z - z + log(2 pi)/2]. */
static ld stirling_rem(ld z) {
  ld r = 1.0L / (z * z);
  return (1.0L / 12 - r * (1.0L / 360 - r * (1.0L / 1260 - r * (1.0L / 1680 - r / 1188)))) / z;
}

/* lgamma(a + b) - lgamma(a) without cancellation when a is large. */
static ld lgamma_diff(ld a, ld b) {
  if (a < 1000.0L) return lgammal(a + b) - lgammal(a);
  return (a - 0.5L) * log1pl(b / a) + b * logl(a + b) - b
    + stirling_rem(a + b) - stirling_rem(a);
}

static ld ibeta(ld x, ld y, ld a, ld b) {
  if (x <= 0.0L) return 0.0L;
  if (y <= 0.0L) return 1.0L;
  ld lbt = (a >= b ? lgamma_diff(a, b) - lgammal(b) : lgamma_diff(b, a) - lgammal(a))
    + a * logl(x) + b * logl(y);
  if (x < (a + 1.0L) / (a + b + 2.0L))
    return expl(lbt) * beta_cf(a, b, x) / a;
  return 1.0L - expl(lbt) * beta_cf(b, a, y) / b;
}

double Rf_pbeta(double x, double a, double b, int lower_tail, int log_p) {
  if (isnan(x) || isnan(a) || isnan(b)) return x + a + b;
  if (a < 0 || b < 0) return R_NaN;
  ld p = (x <= 0) ? 0.0L : (x >= 1 ? 1.0L : ibeta(x, 1.0L - (ld)x, a, b));
  ld r = lower_tail ? p : 1.0L - p;
  return log_p ? (double)logl(r) : (double)r;
}

/* Student t: P(|T| > |x|) = I_{n/(n+x^2)}(n/2, 1/2). */
double Rf_pt(double x, double n, int lower_tail, int log_p) {
  if (isnan(x) || isnan(n)) return x + n;
  if (n <= 0.0) return R_NaN;
  ld two_sided;
  if (isinf(x)) two_sided = 0.0L;
  else if (isinf(n)) return Rf_pnorm5(x, 0.0, 1.0, lower_tail, log_p);
  else {
    ld xx = (ld)x * (ld)x, den = (ld)n + xx;
    two_sided = ibeta((ld)n / den, xx / den, (ld)n / 2.0L, 0.5L);
  }
  ld tail = two_sided / 2.0L;                 /* P(T > |x|) */
  int want_upper_of_abs = (x > 0) ? !lower_tail : lower_tail;
  ld r = want_upper_of_abs ? tail : 1.0L - tail;
  return log_p ? (double)logl(r) : (double)r;
}

double Rf_pnorm5(double x, double mu, double sigma, int lower_tail, int log_p) {
  if (isnan(x) || isnan(mu) || isnan(sigma)) return x + mu + sigma;
  if (sigma < 0) return R_NaN;
  ld z = ((ld)x - (ld)mu) / (ld)sigma;
  if (sigma == 0) z = (x < mu) ? -INFINITY : INFINITY;
  ld r = lower_tail ? 0.5L * erfcl(-z / sqrtl(2.0L)) : 0.5L * erfcl(z / sqrtl(2.0L));
  return log_p ? (double)logl(r) : (double)r;
}

double Rf_dnorm4(double x, double mu, double sigma, int give_log) {
  if (isnan(x) || isnan(mu) || isnan(sigma)) return x + mu + sigma;
  if (sigma < 0) return R_NaN;
  if (!isfinite(sigma)) return give_log ? -INFINITY : 0.0;
  if (!isfinite(x) && mu == x) return R_NaN;
  if (sigma == 0) return (x == mu) ? INFINITY : (give_log ? -INFINITY : 0.0);
  double z = (x - mu) / sigma;
  if (!isfinite(z)) return give_log ? -INFINITY : 0.0;
  z = fabs(z);
  if (give_log) return -(M_LN_SQRT_2PI + 0.5 * z * z + log(sigma));
  return exp(-0.5 * z * z) / (sqrt(2 * M_PI) * sigma);
}

double Rf_lchoose(double n, double k) {
  return Rf_lgammafn(n + 1) - Rf_lgammafn(k + 1) - Rf_lgammafn(n - k + 1);
}

double Rf_choose(double n, double k) {
  k = nearbyint(k);
  if (k < 0) return 0;
  if (k == 0) return 1;
  if (k < 30) {
    double r = n;
    for (int j = 2; j <= (int)k; j++) r *= (n - j + 1) / j;
    return nearbyint(r);
  }
  return nearbyint(exp(Rf_lchoose(n, k)));
}

double Rf_fround(double x, double digits) {
  double p = pow(10.0, digits);
  return nearbyint(x * p) / p;
}
