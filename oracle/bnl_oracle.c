/* bnl_oracle.c -- CPU restatement of the reference's statistics hot path
 * (SURVEY.md section 8a), on plain arrays, no R objects.
 *
 * TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this; the product
 * library (bnlearn_b200/csrc) never links or calls it.
 *
 * Pinning: the reference ships no golden vectors for this path (SURVEY.md
 * section 4), so this restatement is pinned against the reference's own C
 * compiled here (oracle/_ref/libbnref.so; tests/test_oracle_vs_ref.py) and
 * against fixtures generated from it (tests/golden/*.json, made by
 * tests/golden/make_fixtures.py). The R nmath / LAPACK / LINPACK pieces under
 * it are "parity unpinned" (see rmath_port.c, linpack_port.c).
 *
 * Citations are relative to /root/reference/. Data conventions follow the
 * reference: factor codes are int, 1-based, NA = INT_MIN; numeric columns are
 * double; everything column-major. */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <limits.h>

#define ORC_NA INT_MIN
#define ORC_TOL 1.4901161193847656e-08   /* sqrt(DBL_EPSILON): src/include/globals.h:21 */

double Rf_lgammafn(double x);
double Rf_pchisq(double x, double df, int lower_tail, int log_p);
double Rf_pt(double x, double n, int lower_tail, int log_p);
double Rf_pnorm5(double x, double mu, double sigma, int lower_tail, int log_p);
double Rf_dnorm4(double x, double mu, double sigma, int give_log);
void minir_f77_dqrdc2(double *x, int *ldx, int *n, int *p, double *tol,
    int *rank, double *qraux, int *pivot, double *work);
void minir_f77_dqrsl(double *x, int *ldx, int *n, int *k, double *qraux,
    double *y, double *qy, double *qty, double *b, double *rsd, double *xb,
    int *job, int *info);

/* test_e (src/tests/tests.h:8-36) and score_e (src/scores/scores.h:9-60). */
enum { T_MI = 1, T_MI_ADF = 2, T_X2 = 3, T_X2_ADF = 4, T_JT = 10, T_MI_SH = 40, T_COR = 20, T_ZF = 21, T_MI_G = 22, T_MI_G_SH = 41 };
enum { S_LOGLIK = 1, S_AIC = 3, S_BIC = 4, S_BDE = 6, S_BDS = 7, S_BDJ = 8, S_K2 = 9,
       S_FNML = 11, S_QNML = 12, S_NAL = 13, S_PNAL = 14, S_NAL_G = 106, S_PNAL_G = 107,
       S_LOGLIK_G = 100, S_AIC_G = 102, S_BIC_G = 103 };

/* ------------------------------------------------------------------------
 * a2: configuration index (src/core/sets.c:162-218). First column varies
 * fastest; a row is NA if any member is NA; `offset` is added to non-NA rows.
 * Returns the number of configurations, or -1 if it does not fit an int. */
long long orc_config(const int *const *cols, int nrow, int ncol, const int *levels,
    int *cfg, int offset) {
  long long nl = 1;
  for (int j = 0; j < ncol; j++) nl *= levels[j];
  if (nl >= INT_MAX) return -1;
  for (int i = 0; i < nrow; i++) {
    long long c = 0, cum = 1;
    int na = 0;
    for (int j = 0; j < ncol; j++) {
      int v = cols[j][i];
      if (v == ORC_NA) { na = 1; break; }
      c += (long long)(v - 1) * cum;
      cum *= levels[j];
    }
    cfg[i] = na ? ORC_NA : (int)c + offset;
  }
  return nl;
}

/* a4: joint counts n[z][x][y], flattened as ((z * llx) + x) * lly + y, skipping
 * rows with any NA (src/core/contingency.tables.c:20-34, 93-130, 198-223).
 * y == NULL gives a (z, x) table, z == NULL a single stratum. */
void orc_counts(const int *x, int llx, const int *y, int lly, const int *z, int llz,
    int num, int *table) {
  size_t cells = (size_t)llx * (y ? lly : 1) * (z ? llz : 1);
  memset(table, 0, cells * sizeof(int));
  int ly = y ? lly : 1;
  for (int k = 0; k < num; k++) {
    if (x[k] == ORC_NA) continue;
    if (y && y[k] == ORC_NA) continue;
    if (z && z[k] == ORC_NA) continue;
    size_t zi = z ? (size_t)(z[k] - 1) : 0;
    size_t yi = y ? (size_t)(y[k] - 1) : 0;
    table[(zi * llx + (size_t)(x[k] - 1)) * ly + yi]++;
  }
}

/* a10-a12: statistic and degrees of freedom from an n[z][x][y] table
 * (src/tests/discrete/discrete.tests.c:5-8, 11-94, 97-192;
 *  src/tests/discrete/df.adjust.c:4-90). */
double orc_discrete_stat(const int *n, int llx, int lly, int llz, int test, double *df) {
  int *ni = calloc((size_t)llz * llx, sizeof(int));
  int *nj = calloc((size_t)llz * lly, sizeof(int));
  int *nk = calloc((size_t)llz, sizeof(int));
  long long nobs = 0;
  for (int k = 0; k < llz; k++)
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) {
        int c = n[((size_t)k * llx + i) * lly + j];
        ni[k * llx + i] += c; nj[k * lly + j] += c; nk[k] += c;
      }
  for (int k = 0; k < llz; k++) nobs += nk[k];

  /* degrees of freedom. */
  double d = 0;
  if (test == T_MI || test == T_X2) {
    d = (double)(llx - 1) * (double)(lly - 1) * (double)llz;
  } else {
    for (int k = 0; k < llz; k++) {
      int alx = 0, aly = 0;
      for (int i = 0; i < llx; i++) alx += ni[k * llx + i] > 0;
      for (int j = 0; j < lly; j++) aly += nj[k * lly + j] > 0;
      if (alx < 1) alx = 1;
      if (aly < 1) aly = 1;
      d += (double)(alx - 1) * (double)(aly - 1);
    }
  }
  if (df) *df = d;

  double res = 0;
  int adf = (test == T_MI_ADF || test == T_X2_ADF);
  if (nobs == 0) goto done;
  if (adf && (double)nobs < 5.0 * llx * lly * llz) goto done;

  long double acc = 0;
  if (test == T_MI || test == T_MI_ADF) {
    for (int k = 0; k < llz; k++)
      for (int i = 0; i < llx; i++)
        for (int j = 0; j < lly; j++) {
          int c = n[((size_t)k * llx + i) * lly + j];
          if (c != 0)
            acc += (double)c * log((double)c * (double)nk[k] /
                     ((double)ni[k * llx + i] * (double)nj[k * lly + j]));
        }
    res = (double)acc / (double)nobs;
    res *= 2.0 * (double)nobs;
  } else {
    for (int k = 0; k < llz; k++) {
      if (nk[k] == 0) continue;
      for (int i = 0; i < llx; i++)
        for (int j = 0; j < lly; j++) {
          long double e = ni[k * llx + i] * (double)nj[k * lly + j] / nk[k];
          int c = n[((size_t)k * llx + i) * lly + j];
          if (e != 0) acc += (c - e) * (c - e) / e;
        }
    }
    res = (double)acc;
  }
done:
  free(ni); free(nj); free(nk);
  return res;
}

/* a11 + a13: one discrete CI test of x vs y given the columns z[0..nz)
 * (src/tests/patterns/ctest.c:11-55, utest.c:23-62): configuration of Z,
 * table, statistic, df, p = pchisq(stat, df, upper). Returns 0, or -1 when the
 * configuration count overflows (the reference raises an R error there,
 * src/core/sets.c:182-183). */
/* Jonckheere-Terpstra, standardised (src/tests/discrete/jonckheere.c:12-96 for
 * the two drivers, :98-172 for mean, variance and the statistic). tab is
 * n[z][x][y]; `conditional` selects c_cjt's handling of unusable strata. The
 * reference forms num * num in int (:101), which overflows past 46340 complete
 * rows; this restatement (like the engine) stays in double. */
static double orc_jt(const int *tab, int llx, int lly, int llz, int conditional) {
  double centred = 0, tvar = 0;
  long long nobs = 0;
  for (int k = 0; k < llz; k++) {
    const int *n = tab + (size_t)k * llx * lly;
    double *ni = calloc((size_t)llx, sizeof(double)), *nj = calloc((size_t)lly, sizeof(double));
    double nk = 0;
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) { ni[i] += n[i * lly + j]; nj[j] += n[i * lly + j]; nk += n[i * lly + j]; }
    nobs += (long long)nk;
    if (nk > 0) {
      double stat = 0;
      for (int i = 1; i < llx; i++)
        for (int j = 0; j < i; j++) {
          double w = 0;
          for (int s = 0; s < lly; s++) {
            double below = 0;
            for (int t = 0; t < s; t++) below += n[i * lly + t] + n[j * lly + t];
            w += (below + ((double)n[i * lly + s] + (double)n[j * lly + s] + 1) / 2) * (double)n[i * lly + s];
          }
          stat += w - ni[i] * (ni[i] + 1) / 2;
        }
      double mean = nk * nk;
      for (int i = 0; i < llx; i++) mean -= ni[i] * ni[i];
      mean /= 4;
      double U1 = nk * (nk - 1) * (2 * nk + 5), U2a = 0, U2b = 0, U3a = 0, U3b = 0;
      for (int i = 0; i < llx; i++) U1 -= ni[i] * (ni[i] - 1) * (2 * ni[i] + 5);
      for (int j = 0; j < lly; j++) U1 -= nj[j] * (nj[j] - 1) * (2 * nj[j] + 5);
      for (int i = 0; i < llx; i++) { U2a += ni[i] * (ni[i] - 1) * (ni[i] - 2); U3a += ni[i] * (ni[i] - 1); }
      for (int j = 0; j < lly; j++) { U2b += nj[j] * (nj[j] - 1) * (nj[j] - 2); U3b += nj[j] * (nj[j] - 1); }
      double var = U1 / 72 + U2a * U2b / (36 * nk * (nk - 1) * (nk - 2)) + U3a * U3b / (8 * nk * (nk - 1));
      centred += stat - mean;
      if (!conditional || (!isnan(var) && var > ORC_TOL)) tvar += var;
    }
    free(ni); free(nj);
  }
  if (!conditional && nobs <= 1) return 0;
  return tvar < ORC_TOL ? 0 : centred / sqrt(tvar);
}

/* shrinkage mutual information, scaled to G^2 (src/tests/discrete/shrinkage.c:
 * 15-76 unconditional, :79-161 conditional, :164-207 for lambda). */
static double orc_shmi(const int *tab, int llx, int lly, int llz, int conditional) {
  size_t cells = (size_t)llx * lly * llz;
  double num = 0, target = 1.0 / ((double)llx * lly * llz);
  for (size_t c = 0; c < cells; c++) num += tab[c];
  if (num <= 1) return 0;
  double lnum = 0, lden = 0;
  for (size_t c = 0; c < cells; c++) {
    double f = tab[c] / num;
    lnum += f * f;
    lden += (target - f) * (target - f);
  }
  double lambda = lden == 0 ? 1 : (1 - lnum) / ((num - 1) * lden);
  if (lambda > 1) lambda = 1;
  if (lambda < 0) lambda = 0;
  double res = 0;
  double *p = malloc(sizeof(double) * (size_t)llx * lly);
  double *pi = malloc(sizeof(double) * (size_t)llx), *pj = malloc(sizeof(double) * (size_t)lly);
  for (int k = 0; k < llz; k++) {
    const int *n = tab + (size_t)k * llx * lly;
    double pk = 0;
    memset(pi, 0, sizeof(double) * (size_t)llx);
    memset(pj, 0, sizeof(double) * (size_t)lly);
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) {
        double v = lambda * target + (1 - lambda) * n[i * lly + j] / num;
        p[i * lly + j] = v; pi[i] += v; pj[j] += v; pk += v;
      }
    if (conditional && pk == 0) continue;
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) {
        double v = p[i * lly + j];
        if (conditional) { if (v > 0) res += v * log(v * pk / (pi[i] * pj[j])); }
        else if (v != 0) res += v * log(v / (pi[i] * pj[j]));
      }
  }
  free(p); free(pi); free(pj);
  return res * 2 * num;
}

int orc_ci_discrete(int test, const int *x, int llx, const int *y, int lly,
    const int *const *z, const int *zlev, int nz, int num,
    double *stat, double *df, double *pvalue) {
  int llz = 1, *cfg = NULL;
  if (nz > 0) {
    cfg = malloc((size_t)(num > 0 ? num : 1) * sizeof(int));
    long long nl = orc_config(z, num, nz, zlev, cfg, 1);
    if (nl < 0) { free(cfg); return -1; }
    llz = (int)nl;
  }
  int *tab = malloc((size_t)llx * lly * llz * sizeof(int));
  orc_counts(x, llx, y, lly, cfg, llz, num, tab);
  if (test == T_JT) {
    *stat = orc_jt(tab, llx, lly, llz, nz > 0);
    *df = 0;
    *pvalue = 2 * Rf_pnorm5(fabs(*stat), 0, 1, 0, 0);
  } else if (test == T_MI_SH) {
    *stat = orc_shmi(tab, llx, lly, llz, nz > 0);
    *df = (double)(llx - 1) * (double)(lly - 1) * (double)llz;
    *pvalue = Rf_pchisq(*stat, *df, 0, 0);
  } else {
    *stat = orc_discrete_stat(tab, llx, lly, llz, test, df);
    *pvalue = Rf_pchisq(*stat, *df, 0, 0);
  }
  free(tab); free(cfg);
  return 0;
}

/* ------------------------------------------------------------------------
 * a5-a7: discrete node scores.
 *   loglik / aic / bic : src/scores/discrete.loglikelihood.c:7-74 and the
 *                        penalty at src/scores/per.node.score.c:98-114
 *   bde / bds          : src/scores/dirichlet.posterior.c:9-71, 74-149, with the
 *                        parent configurations as a factor over ALL Pi(levels)
 *                        codes (bde) or over the OBSERVED codes only (bds;
 *                        :180, src/minimal/common.c:43-83).
 * `arg` is k for aic/bic and iss for bde/bds. The uniform graph prior adds 0
 * (src/scores/graph.priors.c:35-37). Returns NAN if the configuration count
 * overflows an int. */
/* multinomial NML regret (src/scores/nml.regret.c:36-62 over the table of
 * src/scores/nml.regret.table.c:6-69 and the two large-n / large-k
 * approximations of nml.regret.c:6-33). */
static double *orc_regret_tab = NULL;
static double orc_regret(double n, double k) {
  enum { N = 1000, K = 100 };
  if (n == 1) return log(k);
  if (k == 1 || n == 0) return 0;
  if (n <= N && k <= K) {
    if (!orc_regret_tab) {
      double *t = calloc((size_t)(N + 1) * (K + 1), sizeof(double));
      for (int m = 1; m <= N; m++) {
        double sum = 0;
        for (int h = 0; h <= m; h++) {
          double term = Rf_lgammafn(m + 1.0) - Rf_lgammafn(h + 1.0) - Rf_lgammafn(m - h + 1.0);
          if (h) term += h * (log((double)h) - log((double)m));
          if (m - h) term += (m - h) * (log((double)(m - h)) - log((double)m));
          sum += exp(term);
        }
        t[2 + (size_t)m * (K + 1)] = log(sum);
        double c1 = 1, c2 = exp(t[2 + (size_t)m * (K + 1)]);
        for (int kk = 3; kk <= K; kk++) {
          double c3 = c2 + c1 / (kk - 2) * m;
          t[kk + (size_t)m * (K + 1)] = log(c3);
          c1 = c2; c2 = c3;
        }
      }
      orc_regret_tab = t;
    }
    return orc_regret_tab[(int)k + (size_t)(int)n * (K + 1)];
  }
  double hk = 0.5 * k, hkh = hk - 0.5;
  double ratio = exp(lgamma(hk) - lgamma(hkh));
  double g = M_SQRT1_2 - ratio * k / 3 / sqrt(n);
  double a1 = hkh * (log(n) - M_LN2) + 0.5 * log(M_PI) - lgamma(hk) + 0.5 - g * g +
              (3 + k * (k - 2) * (2 * k + 1)) / (36.0 * n);
  double a = k / n, ca = 0.5 * (1 + sqrt(1 + 4 / a));
  double a2 = n * (log(a) + (a + 2) * log(ca) - 1 / ca) - 0.5 * log(ca + 2 / a);
  return a1 < a2 ? a1 : a2;
}

double orc_score_discrete(int score, const int *x, int llx, const int *const *pa,
    const int *plev, int np, int num, double arg, double *nparams) {
  double res = 0, npar = 0;
  if (np == 0) {
    int *n = calloc((size_t)llx, sizeof(int));
    if (score == S_BDE || score == S_BDS || score == S_BDJ || score == S_K2) {
      /* dpost() does not skip NA: the front end refuses incomplete data. */
      for (int i = 0; i < num; i++) n[x[i] - 1]++;
      /* per-cell scores (K2: 1, BDJ: 1/2; per.node.score.c:233-246): alpha is
       * the constant and the "imaginary sample size" is llx (:19-25). */
      int per_cell = (score == S_BDJ || score == S_K2);
      double imaginary = per_cell ? llx : arg;
      double alpha = per_cell ? (score == S_K2 ? 1.0 : 0.5) : arg / llx;
      for (int i = 0; i < llx; i++)
        res += Rf_lgammafn(n[i] + alpha) - Rf_lgammafn(alpha);
      res += Rf_lgammafn(imaginary) - Rf_lgammafn(imaginary + num);
    } else {
      int nobs = 0;
      for (int i = 0; i < num; i++) if (x[i] != ORC_NA) n[x[i] - 1]++;
      for (int i = 0; i < llx; i++) nobs += n[i];
      for (int i = 0; i < llx; i++)
        if (n[i] != 0) res += (double)n[i] * log((double)n[i] / nobs);
      npar = llx - 1;
      /* node-average likelihood over the locally complete rows, penalty scaled by
       * the full sample size (src/scores/discrete.nal.c:8-37). */
      if (score == S_NAL || score == S_PNAL)
        res = nobs == 0 ? -INFINITY : res / nobs - (score == S_PNAL ? arg : 0.0) / num * npar;
      /* fnml() / qnml() of a root node (normalized.maximum.likelihood.c:26-41, 107-111). */
      if (score == S_FNML || score == S_QNML) res -= orc_regret(nobs, llx);
    }
    free(n);
  } else {
    int *cfg = malloc((size_t)(num > 0 ? num : 1) * sizeof(int));
    long long nl = orc_config(pa, num, np, plev, cfg, 1);
    if (nl < 0) { free(cfg); return NAN; }
    int lly = (int)nl;
    int *n = calloc((size_t)llx * lly, sizeof(int));      /* n[x][cfg] */
    int *nj = calloc((size_t)lly, sizeof(int));
    if (score == S_BDE || score == S_BDS || score == S_BDJ || score == S_K2) {
      for (int i = 0; i < num; i++) {
        n[(size_t)(x[i] - 1) * lly + (cfg[i] - 1)]++;
        nj[cfg[i] - 1]++;
      }
      int q = lly;
      if (score == S_BDS) { q = 0; for (int j = 0; j < lly; j++) q += nj[j] > 0; }
      double alpha = arg / ((double)llx * q);
      if (score == S_BDJ || score == S_K2) {
        /* cdpost() with per_cell: alpha fixed, imaginary = alpha * llx * lly (:81-88). */
        alpha = (score == S_K2) ? 1.0 : 0.5;
        arg = alpha * (double)llx * q;
      }
      for (int i = 0; i < llx; i++)
        for (int j = 0; j < lly; j++) {
          if (score == S_BDS && nj[j] == 0) continue;
          res += Rf_lgammafn(n[(size_t)i * lly + j] + alpha) - Rf_lgammafn(alpha);
        }
      for (int j = 0; j < lly; j++) {
        if (score == S_BDS && nj[j] == 0) continue;
        res += Rf_lgammafn(arg / q) - Rf_lgammafn(nj[j] + arg / q);
      }
    } else {
      for (int i = 0; i < num; i++)
        if (x[i] != ORC_NA && cfg[i] != ORC_NA) {
          n[(size_t)(x[i] - 1) * lly + (cfg[i] - 1)]++;
          nj[cfg[i] - 1]++;
        }
      for (int i = 0; i < llx; i++)
        for (int j = 0; j < lly; j++) {
          int c = n[(size_t)i * lly + j];
          if (c != 0) res += (double)c * log((double)c / nj[j]);
        }
      npar = (double)(llx - 1) * (double)lly;
      if (score == S_FNML) {                        /* cfnml(), normalized.maximum.likelihood.c:44-63 */
        for (int j = 0; j < lly; j++) if (nj[j] != 0) res -= orc_regret(nj[j], llx);
      }
      if (score == S_QNML) {                        /* cqnml() over cdlik_diff(), :10-24, 113-130 */
        long long nobs = 0;
        res = 0;
        for (int j = 0; j < lly; j++) {
          for (int i = 0; i < llx; i++) {
            int c = n[(size_t)i * lly + j];
            if (c != 0) res += (double)c * log((double)c);
          }
          if (nj[j] != 0) res -= (double)nj[j] * log((double)nj[j]);
          nobs += nj[j];
        }
        res -= orc_regret((double)nobs, (double)llx * lly);
        res += orc_regret((double)nobs, lly);
      }
      if (score == S_NAL || score == S_PNAL) {      /* src/scores/discrete.nal.c:39-68 */
        long long nobs = 0;
        for (int j = 0; j < lly; j++) nobs += nj[j];
        res = nobs == 0 ? -INFINITY : res / nobs - (score == S_PNAL ? arg : 0.0) / num * npar;
      }
    }
    free(n); free(nj); free(cfg);
  }
  if (score == S_AIC || score == S_BIC) res -= arg * npar;
  if (nparams) *nparams = npar;
  return res;
}

/* a8: robust score difference (src/learning/score/score.delta.c:234-258). */
double orc_robust_difference(double old_score, double new_score) {
  double delta = (fabs(new_score - old_score) < ORC_TOL) ? 0 : new_score - old_score;
  if (old_score == -INFINITY)
    delta = (new_score != -INFINITY) ? new_score : -INFINITY;
  return delta;
}

/* ------------------------------------------------------------------------
 * Gaussian side. a14: moments and covariance with extended-precision sums
 * (src/core/moments.c:7-45, src/core/covariance.matrix.c:65-97). */
static double orc_mean(const double *x, int n) {
  long double s = 0;
  for (int i = 0; i < n; i++) s += x[i];
  s /= n;
  return (double)s;
}

static double orc_sse(const double *x, double m, int n) {
  long double s = 0;
  for (int i = 0; i < n; i++) s += (x[i] - m) * (x[i] - m);
  return (double)s;
}

void orc_covmat(const double *const *cols, int n, int p, double *mean, double *cov) {
  for (int j = 0; j < p; j++) mean[j] = orc_mean(cols[j], n);
  if (n <= 1) { memset(cov, 0, (size_t)p * p * sizeof(double)); return; }
  for (int i = 0; i < p; i++)
    for (int j = i; j < p; j++) {
      long double t = 0;
      for (int k = 0; k < n; k++)
        t += (cols[i][k] - mean[i]) * (cols[j][k] - mean[j]);
      cov[i + j * p] = cov[j + i * p] = (double)(t / (n - 1));
    }
}

/* cyclic Jacobi eigen-decomposition of a symmetric p x p matrix (column-major,
 * destroyed); eigenvalues in d, eigenvectors in the columns of v. Stands in for
 * the LAPACK dgesdd the reference calls (src/math/linear.algebra.c:58-84): for
 * a symmetric matrix the singular values are |eigenvalues| and u_i * vt_i is
 * sign(lambda_i) q_i q_i'. */
static void orc_jacobi(double *a, int p, double *d, double *v) {
  for (int i = 0; i < p; i++)
    for (int j = 0; j < p; j++) v[i + j * p] = (i == j);
  for (int sweep = 0; sweep < 100; sweep++) {
    double off = 0, diag = 0;
    for (int i = 0; i < p; i++) {
      diag += a[i + i * p] * a[i + i * p];
      for (int j = i + 1; j < p; j++) off += a[i + j * p] * a[i + j * p];
    }
    if (off <= 1e-40 * diag || off == 0) break;
    for (int r = 0; r < p - 1; r++)
      for (int c = r + 1; c < p; c++) {
        double apq = a[r + c * p];
        if (apq == 0) continue;
        double theta = (a[c + c * p] - a[r + r * p]) / (2 * apq);
        double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1));
        double cs = 1 / sqrt(t * t + 1), sn = t * cs;
        for (int k = 0; k < p; k++) {
          double akr = a[k + r * p], akc = a[k + c * p];
          a[k + r * p] = cs * akr - sn * akc;
          a[k + c * p] = sn * akr + cs * akc;
        }
        for (int k = 0; k < p; k++) {
          double ark = a[r + k * p], ack = a[c + k * p];
          a[r + k * p] = cs * ark - sn * ack;
          a[c + k * p] = sn * ark + cs * ack;
        }
        for (int k = 0; k < p; k++) {
          double vkr = v[k + r * p], vkc = v[k + c * p];
          v[k + r * p] = cs * vkr - sn * vkc;
          v[k + c * p] = sn * vkr + cs * vkc;
        }
      }
  }
  for (int i = 0; i < p; i++) d[i] = a[i + i * p];
}

/* a15: partial correlation of variables 0 and 1 given the rest, from the
 * pseudo-inverse of the covariance matrix with singular values below
 * dim * d_max * eps dropped (src/core/correlation.c:43-92). */
double orc_pcor(const double *cov, int p, int *clamped) {
  double *a = malloc((size_t)p * p * sizeof(double));
  double *v = malloc((size_t)p * p * sizeof(double));
  double *d = malloc((size_t)p * sizeof(double));
  memcpy(a, cov, (size_t)p * p * sizeof(double));
  orc_jacobi(a, p, d, v);
  double dmax = 0;
  for (int i = 0; i < p; i++) if (fabs(d[i]) > dmax) dmax = fabs(d[i]);
  double sv_tol = p * dmax * ORC_TOL * ORC_TOL;
  double k11 = 0, k12 = 0, k22 = 0;
  for (int i = 0; i < p; i++)
    if (fabs(d[i]) > sv_tol) {
      k11 += v[0 + i * p] * v[0 + i * p] / d[i];
      k12 += v[0 + i * p] * v[1 + i * p] / d[i];
      k22 += v[1 + i * p] * v[1 + i * p] / d[i];
    }
  free(a); free(v); free(d);
  double res = (k11 < ORC_TOL || k22 < ORC_TOL) ? 0 : -k12 / sqrt(k11 * k22);
  if (clamped) *clamped = 0;
  if (res > 1) { res = 1; if (clamped) *clamped = 1; }
  else if (res < -1) { res = -1; if (clamped) *clamped = 1; }
  return res;
}

/* a16-a17: one Gaussian CI test of cols[0] vs cols[1] given cols[2..p)
 * (src/tests/patterns/utest.c:65-129, ctest.c:58-142;
 *  src/tests/gaussian/gaussian.tests.c:4-22, df.adjust.c:4-28).
 * Returns 1 if the zero-degrees-of-freedom branch was taken (an R warning in
 * the reference), else 0. `stat` is what the htest object reports: the
 * (partial) correlation for cor, the transformed statistic for zf and mi-g. */
static int orc_ci_gaussian_complete(int test, const double *const *cols, int p, int n,
    double *stat, double *df, double *pvalue);

/* rows on which all the p columns are observed, copied out contiguously: what
 * c_covmat_with_missing (src/core/covariance.matrix.c:100-184) and
 * c_cor_with_missing (src/core/correlation.c:96-155) restrict themselves to.
 * Returns the number of complete rows; *out is NULL when nothing is missing. */
static int orc_complete_rows(const double *const *cols, int p, int n, double ***out) {
  int nc = 0;
  char *keep = malloc((size_t)(n > 0 ? n : 1));
  for (int i = 0; i < n; i++) {
    keep[i] = 1;
    for (int j = 0; j < p; j++) if (isnan(cols[j][i])) { keep[i] = 0; break; }
    nc += keep[i];
  }
  *out = NULL;
  if (nc < n) {
    double **c = malloc((size_t)p * sizeof(double *));
    for (int j = 0; j < p; j++) {
      c[j] = malloc((size_t)(nc > 0 ? nc : 1) * sizeof(double));
      for (int i = 0, k = 0; i < n; i++) if (keep[i]) c[j][k++] = cols[j][i];
    }
    *out = c;
  }
  free(keep);
  return nc;
}

int orc_ci_gaussian(int test, const double *const *cols, int p, int n,
    double *stat, double *df, double *pvalue) {
  double **sub = NULL;
  int nc = orc_complete_rows(cols, p, n, &sub);
  if (!sub) return orc_ci_gaussian_complete(test, cols, p, n, stat, df, pvalue);
  int rc = orc_ci_gaussian_complete(test, (const double *const *)sub, p, nc, stat, df, pvalue);
  for (int j = 0; j < p; j++) free(sub[j]);
  free(sub);
  return rc;
}

/* James-Stein shrinkage intensity of mi-g-sh (src/tests/gaussian/shrinkage.c:15-57
 * for a marginal test, :60-106 for a conditional one): sum over pairs i < j and
 * rows of (d_i d_j - s_ij)^2 against the sum of squared covariances, on the log
 * scale; s_ij has divisor n for the matrix form and n - 1 for the pair form. */
static double orc_shrink_lambda(const double *const *cols, int p, int n, const double *mean,
    const double *cov) {
  long double num = 0, den = 0;
  double scale = p == 2 ? 1.0 : (double)(n - 1) / n;
  for (int i = 0; i < p; i++)
    for (int j = i + 1; j < p; j++) {
      double c = cov[i + j * p];
      for (int r = 0; r < n; r++) {
        long double t = (cols[i][r] - mean[i]) * (cols[j][r] - mean[j]) - c * scale;
        num += t * t;
      }
      den += (long double)c * c;
    }
  if (!(num > ORC_TOL)) return 0;
  long double l = expl(logl(num) - logl(den) + logl((long double)n) - 3 * logl((long double)(n - 1)));
  if (l > 1) l = 1;
  if (l < 0) l = 0;
  return (double)l;
}

static int orc_ci_gaussian_complete(int test, const double *const *cols, int p, int n,
    double *stat, double *df, double *pvalue) {
  int nz = p - 2;
  double d = (test == T_COR) ? n - nz - 2 : (test == T_ZF) ? n - nz - 3 : 1;
  if (d < 1 || n == 0) { *df = 0; *stat = 0; *pvalue = 1; return 1; }
  *df = d;
  double r;
  double *mean = malloc((size_t)p * sizeof(double));
  double *cov = malloc((size_t)p * p * sizeof(double));
  orc_covmat(cols, n, p, mean, cov);
  double lambda = test == T_MI_G_SH ? orc_shrink_lambda(cols, p, n, mean, cov) : 0;
  if (nz == 0) {
    double xm = orc_mean(cols[0], n), ym = orc_mean(cols[1], n);
    long double xsd = orc_sse(cols[0], xm, n), ysd = orc_sse(cols[1], ym, n), sum = 0;
    for (int i = 0; i < n; i++) sum += (cols[0][i] - xm) * (cols[1][i] - ym);
    sum = (xsd < ORC_TOL || ysd < ORC_TOL) ? 0 : sum / sqrtl(xsd * ysd);
    if (sum > 1) sum = 1; else if (sum < -1) sum = -1;
    r = (double)sum * (1 - lambda);
  } else {
    for (int i = 0; i < p; i++)
      for (int j = 0; j < p; j++) if (i != j) cov[i + j * p] *= 1 - lambda;
    r = orc_pcor(cov, p, NULL);
  }
  free(mean); free(cov);
  if (test == T_COR) {
    double t = r * sqrt(d) / sqrt(1 - r * r);
    *stat = r;
    *pvalue = 2 * Rf_pt(fabs(t), d, 0, 0);
  } else if (test == T_ZF) {
    *stat = log((1 + r) / (1 - r)) / 2 * sqrt(d);
    *pvalue = 2 * Rf_pnorm5(fabs(*stat), 0, 1, 0, 0);
  } else {
    *stat = 2 * n * (-0.5 * log(1 - r * r));
    *pvalue = Rf_pchisq(*stat, d, 0, 0);
  }
  return 0;
}

/* ------------------------------------------------------------------------
 * a18: Gaussian node score (src/scores/gaussian.loglikelihood.c:8-83 over
 * src/math/least.squares.c:8-352 and src/math/linear.algebra.c:136-274):
 * OLS of x on its parents (closed forms for 0/1/2 parents, R's pivoted
 * Householder QR otherwise), residual sd with n - p - 1 degrees of freedom,
 * loglik = sum_i dnorm(x_i, fitted_i, sd, log); -Inf when sd < sqrt(eps).
 * `k` is the aic-g / bic-g penalty coefficient (per.node.score.c:137-153). */
static double orc_score_gaussian_complete(int score, const double *x, const double *const *pa,
    int np, int n, double k, double *nparams);

/* nal-g / pnal-g (src/scores/gaussian.nal.c:8-78): the log-likelihood of the
 * locally complete rows, averaged over them, with the penalty scaled by the full
 * sample size; -Inf for singular fits and for nodes without complete rows. The
 * other scores go straight to the complete-data code. */
double orc_score_gaussian(int score, const double *x, const double *const *pa, int np,
    int n, double k, double *nparams) {
  if (score != S_NAL_G && score != S_PNAL_G)
    return orc_score_gaussian_complete(score, x, pa, np, n, k, nparams);
  const double **cols = malloc((size_t)(np + 1) * sizeof(double *));
  cols[0] = x;
  for (int j = 0; j < np; j++) cols[1 + j] = pa[j];
  double **sub = NULL;
  int nc = orc_complete_rows(cols, np + 1, n, &sub);
  const double *const *use = sub ? (const double *const *)sub : cols;
  double ll = nc == 0 ? -INFINITY
            : orc_score_gaussian_complete(S_LOGLIK_G, use[0], use + 1, np, nc, 0, nparams);
  if (nparams) *nparams = np + 2;
  if (sub) { for (int j = 0; j <= np; j++) free(sub[j]); free(sub); }
  free(cols);
  if (ll == -INFINITY) return ll;
  return ll / nc - (score == S_PNAL_G ? k : 0.0) / n * (np + 2);
}

static double orc_score_gaussian_complete(int score, const double *x, const double *const *pa,
    int np, int n, double k, double *nparams) {
  double *fitted = malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  double sd = 0, res = 0;

  if (np == 0) {
    /* glik(): the mean is a plain double sum here (:15-17). */
    double xm = 0;
    for (int i = 0; i < n; i++) xm += x[i];
    xm /= n;
    long double s = 0;
    if (n == 0) s = NAN;
    else if (n == 1) s = 0;
    else { for (int i = 0; i < n; i++) s += (x[i] - xm) * (x[i] - xm); s = sqrtl(s / (n - 1)); }
    sd = (double)s;
    for (int i = 0; i < n; i++) fitted[i] = xm;
    if (s < ORC_TOL) { free(fitted); if (nparams) *nparams = 2; res = -INFINITY; goto penalty; }
  } else if (np == 1) {
    const double *m[2] = { x, pa[0] };
    double mean[2], cov[4];
    orc_covmat(m, n, 2, mean, cov);
    int singular = fabs(cov[3]) < ORC_TOL;
    double b = singular ? NAN : cov[1] / cov[3];
    double a = singular ? mean[0] : mean[0] - mean[1] * b;
    for (int i = 0; i < n; i++) fitted[i] = singular ? a : a + b * pa[0][i];
    if (n == 0) sd = NAN;
    else if (n <= 2) sd = 0;
    else {
      long double sse = 0;
      for (int i = 0; i < n; i++) sse += (x[i] - fitted[i]) * (x[i] - fitted[i]);
      sd = sqrt(sse / (n - 2));
    }
  } else if (np == 2) {
    const double *m[3] = { x, pa[0], pa[1] };
    double mean[3], c[9], a, b1 = 0, b2 = 0;
    orc_covmat(m, n, 3, mean, c);
    int s1 = fabs(c[4]) < ORC_TOL;
    int s2 = (fabs(c[8]) < ORC_TOL) || (fabs(c[5]) / sqrt(c[4] * c[8]) > 1 - ORC_TOL);
    if (s1 && !s2) { b2 = c[2] / c[8]; a = mean[0] - b2 * mean[2]; }
    else if (!s1 && s2) { b1 = c[1] / c[4]; a = mean[0] - b1 * mean[1]; }
    else if (s1 && s2) { a = mean[0]; }
    else {
      double den = c[4] * c[8] - c[5] * c[5];
      b1 = (c[8] * c[1] - c[5] * c[2]) / den;
      b2 = (c[4] * c[2] - c[5] * c[1]) / den;
      a = mean[0] - b1 * mean[1] - b2 * mean[2];
    }
    for (int i = 0; i < n; i++) {
      if (s1 && !s2) fitted[i] = a + b2 * pa[1][i];
      else if (!s1 && s2) fitted[i] = a + b1 * pa[0][i];
      else if (s1 && s2) fitted[i] = a;
      else fitted[i] = a + b1 * pa[0][i] + b2 * pa[1][i];
    }
    if (n == 0) sd = NAN;
    else if (n <= 3) sd = 0;
    else {
      long double sse = 0;
      for (int i = 0; i < n; i++) sse += (x[i] - fitted[i]) * (x[i] - fitted[i]);
      sd = sqrt(sse / (n - 3));
    }
  } else {
    /* c_olsp() -> c_qr(): design matrix [1, parents], dqrdc2 with tol sqrt(eps),
     * fitted values and residuals from dqrsl, sd over n - ncol. */
    int ncol = np + 1, rank = 0, info = 0, job = 11;
    if (n == 1) { fitted[0] = x[0]; sd = 0; }
    else {
      double *qr = malloc((size_t)n * ncol * sizeof(double));
      double *qraux = calloc((size_t)ncol, sizeof(double));
      double *work = calloc((size_t)2 * ncol, sizeof(double));
      double *rsd = calloc((size_t)n, sizeof(double));
      int *pivot = malloc((size_t)ncol * sizeof(int));
      double tol = ORC_TOL;
      for (int i = 0; i < n; i++) qr[i] = 1.0;
      for (int j = 0; j < np; j++) memcpy(qr + (size_t)(j + 1) * n, pa[j], (size_t)n * sizeof(double));
      for (int j = 0; j < ncol; j++) pivot[j] = j + 1;
      minir_f77_dqrdc2(qr, &n, &n, &ncol, &tol, &rank, qraux, pivot, work);
      memcpy(fitted, x, (size_t)n * sizeof(double));
      minir_f77_dqrsl(qr, &n, &n, &rank, qraux, fitted, NULL, fitted, NULL, rsd, fitted, &job, &info);
      if (info != 0) { sd = 0; memcpy(fitted, x, (size_t)n * sizeof(double)); }
      else if (n <= ncol) sd = 0;
      else sd = sqrt(orc_sse(rsd, 0, n) / (n - ncol));
      free(qr); free(qraux); free(work); free(rsd); free(pivot);
    }
  }

  if (sd < ORC_TOL) res = -INFINITY;
  else for (int i = 0; i < n; i++) res += Rf_dnorm4(x[i], fitted[i], sd, 1);
  free(fitted);
  if (nparams) *nparams = np + 2;
penalty:
  if (score == S_AIC_G || score == S_BIC_G) res -= k * (np + 2);
  return res;
}

/* ------------------------------------------------------------------------
 * BGe node score (wpost / cwpost, src/scores/wishart.posterior.c:10-106), uniform
 * graph prior. cols[0] is the node, cols[1..np] its parents, nu[] their prior
 * means in the same order, data_ncols the width of the whole data frame. The
 * reference takes the two log-determinants from LAPACK dgetrf
 * (src/math/linear.algebra.c:9-56); plain Gaussian elimination with partial
 * pivoting in long double here. */
static long double orc_logdet(long double *m, int k) {
  long double ld = 0;
  for (int c = 0; c < k; c++) {
    int piv = c;
    for (int r = c + 1; r < k; r++) if (fabsl(m[r + c * k]) > fabsl(m[piv + c * k])) piv = r;
    if (m[piv + c * k] == 0) return -INFINITY;
    if (piv != c)
      for (int j = 0; j < k; j++) { long double t = m[c + j * k]; m[c + j * k] = m[piv + j * k]; m[piv + j * k] = t; }
    ld += logl(fabsl(m[c + c * k]));
    for (int r = c + 1; r < k; r++) {
      long double f = m[r + c * k] / m[c + c * k];
      for (int j = c; j < k; j++) m[r + j * k] -= f * m[c + j * k];
    }
  }
  return ld;
}

double orc_score_bge(const double *const *cols, int np, int n, int data_ncols, double iss_mu,
    double iss_w, const double *nu) {
  int k = np + 1;
  double N = n, t = iss_mu * (iss_w - data_ncols - 1) / (iss_mu + 1);
  long double lp = 0.5 * (log(iss_mu) - log(N + iss_mu));
  lp += Rf_lgammafn(0.5 * (N + iss_w - data_ncols + np + 1)) - Rf_lgammafn(0.5 * (iss_w - data_ncols + np + 1));
  lp -= 0.5 * N * log(M_PI);
  double *mean = malloc((size_t)k * sizeof(double));
  for (int j = 0; j < k; j++) mean[j] = orc_mean(cols[j], n);
  if (np == 0) {
    long double sse = orc_sse(cols[0], mean[0], n);
    long double r = t + sse + N * iss_mu / (N + iss_mu) * (mean[0] - nu[0]) * (mean[0] - nu[0]);
    lp += 0.5 * (iss_w - data_ncols + 1) * log(t) - 0.5 * (N + iss_w - data_ncols + 1) * logl(r);
  } else {
    lp += 0.5 * (iss_w - data_ncols + np + 1) * (np + 1) * log(t) - 0.5 * (iss_w - data_ncols + np) * np * log(t);
    long double *R = malloc((size_t)k * k * sizeof(long double));
    long double *Rt = malloc((size_t)np * np * sizeof(long double));
    for (int i = 0; i < k; i++)
      for (int j = 0; j < k; j++) {
        long double s = 0;
        for (int r = 0; r < n; r++) s += (cols[i][r] - mean[i]) * (cols[j][r] - mean[j]);
        R[i + j * k] = s + (i == j ? t : 0) + N * iss_mu / (N + iss_mu) * (mean[i] - nu[i]) * (mean[j] - nu[j]);
      }
    for (int i = 1; i < k; i++)
      for (int j = 1; j < k; j++) Rt[(i - 1) + (j - 1) * np] = R[i + j * k];
    lp += 0.5 * (N + iss_w - data_ncols + np) * orc_logdet(Rt, np);
    lp -= 0.5 * (N + iss_w - data_ncols + np + 1) * orc_logdet(R, k);
    free(R); free(Rt);
  }
  free(mean);
  return (double)lp;
}

/* ------------------------------------------------------------------------
 * Conditional Gaussian log-likelihood of a numeric node given numeric parents
 * gp[0..ngp) within the configurations cfg[] (1-based, nconfig of them) of its
 * discrete parents: c_fast_ccgloglik (src/scores/cg.loglikelihood.c:54-117) over
 * c_cls (src/math/conditional.least.squares.c): per configuration the closed
 * forms for 0 / 1 / 2 regressors -- with the collinearity guards applied to the
 * sums of squares, :109, 262-265 -- or R's pivoted QR for more (c_clsp -> c_qr);
 * sd = 0 when a configuration has no more rows than coefficients; any singular
 * configuration makes the node -Inf. nparams = nconfig * (ngp + 2). */
double orc_cg_loglik(const double *x, const double *const *gp, int ngp, const int *cfg, int nconfig,
    int n, double *nparams) {
  if (nparams) *nparams = (double)nconfig * (ngp + 2);
  double res = 0;
  int *rows = malloc((size_t)(n > 0 ? n : 1) * sizeof(int));
  double *y = malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  double **g = malloc((size_t)(ngp > 0 ? ngp : 1) * sizeof(double *));
  for (int j = 0; j < ngp; j++) g[j] = malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  double *fitted = malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  for (int s = 1; s <= nconfig && res != -INFINITY; s++) {
    int ns = 0;
    for (int i = 0; i < n; i++) if (cfg[i] == s) rows[ns++] = i;
    if (ns == 0) continue;
    for (int r = 0; r < ns; r++) {
      y[r] = x[rows[r]];
      for (int j = 0; j < ngp; j++) g[j][r] = gp[j][rows[r]];
    }
    double sd = 0;
    if (ns <= ngp + 1) { res = -INFINITY; break; }
    long double my = 0, m1 = 0, m2 = 0;
    for (int r = 0; r < ns; r++) my += y[r];
    my /= ns;
    if (ngp == 0) {
      for (int r = 0; r < ns; r++) fitted[r] = (double)my;
    } else if (ngp == 1) {
      for (int r = 0; r < ns; r++) m1 += g[0][r];
      m1 /= ns;
      long double v = 0, c = 0;
      for (int r = 0; r < ns; r++) { v += (g[0][r] - m1) * (g[0][r] - m1); c += (y[r] - my) * (g[0][r] - m1); }
      double b = fabsl(v) < ORC_TOL ? 0 : (double)(c / v);
      double a = (double)(my - m1 * b);
      for (int r = 0; r < ns; r++) fitted[r] = a + g[0][r] * b;
    } else if (ngp == 2) {
      for (int r = 0; r < ns; r++) { m1 += g[0][r]; m2 += g[1][r]; }
      m1 /= ns; m2 /= ns;
      long double v1 = 0, v2 = 0, c01 = 0, c02 = 0, c12 = 0;
      for (int r = 0; r < ns; r++) {
        long double d0 = y[r] - my, d1 = g[0][r] - m1, d2 = g[1][r] - m2;
        v1 += d1 * d1; v2 += d2 * d2; c01 += d0 * d1; c02 += d0 * d2; c12 += d1 * d2;
      }
      int s1 = fabsl(v1) < ORC_TOL;
      int s2 = (fabsl(v2) < ORC_TOL) || (fabsl(c12) / sqrt((double)(v1 * v2)) > 1 - ORC_TOL);
      double b1 = 0, b2 = 0;
      if (s1 && !s2) b2 = (double)(c02 / v2);
      else if (!s1 && s2) b1 = (double)(c01 / v1);
      else if (!s1 && !s2) {
        long double den = v1 * v2 - c12 * c12;
        b1 = (double)((v2 * c01 - c12 * c02) / den);
        b2 = (double)((v1 * c02 - c12 * c01) / den);
      }
      double a = (double)(my - m1 * b1 - m2 * b2);
      for (int r = 0; r < ns; r++) fitted[r] = a + g[0][r] * b1 + g[1][r] * b2;
    } else {
      /* c_qr: intercept + regressors through dqrdc2 / dqrsl (job: fitted values). */
      int ncol = ngp + 1, rank = 0, info = 0, job = 11;
      double *qr = malloc((size_t)ns * ncol * sizeof(double));
      double *qraux = calloc((size_t)ncol, sizeof(double)), *work = calloc((size_t)2 * ncol, sizeof(double));
      int *pivot = malloc((size_t)ncol * sizeof(int));
      double *rsd = malloc((size_t)ns * sizeof(double)), tol = ORC_TOL;
      for (int r = 0; r < ns; r++) qr[r] = 1;
      for (int j = 0; j < ngp; j++) for (int r = 0; r < ns; r++) qr[(size_t)(j + 1) * ns + r] = g[j][r];
      for (int j = 0; j < ncol; j++) pivot[j] = j + 1;
      minir_f77_dqrdc2(qr, &ns, &ns, &ncol, &tol, &rank, qraux, pivot, work);
      memcpy(fitted, y, (size_t)ns * sizeof(double));
      minir_f77_dqrsl(qr, &ns, &ns, &rank, qraux, fitted, NULL, fitted, NULL, rsd, fitted, &job, &info);
      free(qr); free(qraux); free(work); free(pivot); free(rsd);
      if (info != 0) { res = -INFINITY; break; }
    }
    long double ssr = 0;
    for (int r = 0; r < ns; r++) ssr += (y[r] - fitted[r]) * (y[r] - fitted[r]);
    sd = sqrt((double)(ssr / (ns - ngp - 1)));
    if (sd < ORC_TOL) { res = -INFINITY; break; }
    for (int r = 0; r < ns; r++) res += Rf_dnorm4(y[r], fitted[r], sd, 1);
  }
  for (int j = 0; j < ngp; j++) free(g[j]);
  free(g); free(rows); free(y); free(fitted);
  return res;
}
