// bnl_permutation.cu -- Monte Carlo permutation, sequential and semiparametric
// tests (src/tests/permutation/discrete.monte.carlo.c, gaussian.monte.carlo.c).
//
// The row passes run on the device: the observed contingency tables of a whole
// batch come from the counting kernels (bnl_counts), and the permuted
// cross-products of the Gaussian tests from perm_cov_kernel below. What has to
// stay sequential is the random stream: the reference draws from R's generator
// one test after the other, so the caller hands over its uniform generator
// (unif_rand in R) and the library consumes it in exactly the reference's order.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "bnl_internal.h"
#include "bnl_special.cuh"

namespace {

BnlDataSet *find_data_p(bnl_ctx *ctx, bnl_data h) {
  auto it = ctx->data.find(h);
  return it == ctx->data.end() ? nullptr : &it->second;
}

bool is_sequential(int t) { return (t >= BNL_SMC_MI && t < BNL_MC_COR) || (t >= BNL_SMC_COR && t <= BNL_SMC_ZF); }

// One stratum of a table n[x][y] (row-major, y fastest) with its margins.
struct Stratum {
  int *n;                   // llx * lly cells, rewritten by every random table
  std::vector<int> ni, nj;  // margins: invariant under the resampling
  int nk;
};

// A random llx x lly table with the given margins: Patefield's algorithm AS 159,
// as the reference's c_rcont2 runs it (src/core/contingency.tables.c:288-405) --
// cell by cell, the conditional mode first and then outwards, alternating up and
// down, until the cumulated probability passes the uniform draw; a draw that is
// not reached is rescaled by the cumulated mass and redrawn. The uniforms are
// consumed in the same order, and every floating-point expression keeps the
// reference's association, so the same stream gives the same tables. Products of
// margins are formed in 64 bits (the reference casts them to int, which overflows
// past 46 340 x 46 340).
struct TableSampler {
  const double *fact;       // fact[k] = log(k!)
  bnl_unif_fn unif;
  void *ustate;
  std::vector<int> work;
  bool failed = false;

  void draw(Stratum &s, int llx, int lly) {
    work.assign(s.nj.begin(), s.nj.end());
    int left = s.nk;        // observations not yet given to a row
    int ib = 0;
    for (int l = 0; l + 1 < llx; l++) {
      int ia = s.ni[l];     // still to place in this row
      int ic = left;
      left -= ia;
      int m = 0;
      for (; m + 1 < lly; m++) {
        const int id = work[m];
        const int ie = ic;
        ic -= id;
        ib = ie - ia;
        const int ii = ib - id;
        if (ie == 0) {      // nothing left: the rest of the row is empty
          for (int j = m; j + 1 < lly; j++) s.n[l * lly + j] = 0;
          ia = 0;
          break;
        }
        double u = unif(ustate);
        int cell = 0;
        for (;;) {
          int hi = (int)(ia * (id / (double)ie) + 0.5);          // conditional expectation, rounded
          double px = exp(fact[ia] + fact[ib] + fact[ic] + fact[id] - fact[ie] - fact[hi] - fact[id - hi] -
                          fact[ia - hi] - fact[ii + hi]);
          if (px >= u) { cell = hi; break; }
          if (px == 0.0) { failed = true; return; }
          double mass = px, py = px;
          int lo = hi;
          bool found = false;
          for (;;) {
            const long long ju = (long long)((id - hi) * (double)(ia - hi));
            const bool up = ju != 0;
            if (up) {
              ++hi;
              px = px * (double)ju / ((double)hi * (double)(ii + hi));
              mass += px;
              if (mass >= u) { cell = hi; found = true; break; }
            }
            bool down = true;
            while (down) {
              const long long jd = (long long)(lo * (double)(ii + lo));
              down = jd != 0;
              if (!down) break;
              --lo;
              py = py * (double)jd / ((double)(id - lo) * (double)(ia - lo));
              mass += py;
              if (mass >= u) { cell = lo; found = true; break; }
              if (up) break;
            }
            if (found || !up) break;
          }
          if (found) break;
          u = mass * unif(ustate);
        }
        s.n[l * lly + m] = cell;
        ia -= cell;
        work[m] -= cell;
      }
      s.n[l * lly + lly - 1] = ia;
    }
    for (int m = 0; m + 1 < lly; m++) s.n[(llx - 1) * lly + m] = work[m];
    s.n[(llx - 1) * lly + lly - 1] = ib - (lly >= 2 ? s.n[(llx - 1) * lly + lly - 2] : 0);
  }
};

// ---- statistics of a (stratified) table, summed as the reference sums them:
// terms in double, accumulator in long double (src/tests/discrete/discrete.tests.c).
double mi_of(const std::vector<Stratum> &st, int llx, int lly) {
  long double res = 0;
  for (const Stratum &s : st)
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) {
        const int c = s.n[i * lly + j];
        if (c != 0)
          res += (double)c * log((double)c * (double)s.nk / ((double)s.ni[i] * (double)s.nj[j]));
      }
  return (double)res;
}

double x2_of(const std::vector<Stratum> &st, int llx, int lly) {
  long double res = 0;
  for (const Stratum &s : st) {
    if (s.nk == 0) continue;
    for (int i = 0; i < llx; i++)
      for (int j = 0; j < lly; j++) {
        const long double e = s.ni[i] * (double)s.nj[j] / s.nk;
        if (e != 0) res += (s.n[i * lly + j] - e) * (s.n[i * lly + j] - e) / e;
      }
  }
  return (double)res;
}

// Jonckheere-Terpstra (src/tests/discrete/jonckheere.c:77-157): statistic, null
// mean and variance of one stratum. Squares are formed in double (the reference's
// int products wrap past 46 340 observations).
double jt_stat_of(const Stratum &s, int llx, int lly) {
  double res = 0;
  for (int i = 1; i < llx; i++) {
    const double half = (double)s.ni[i] * ((double)s.ni[i] + 1) / 2;
    for (int j = 0; j < i; j++) {
      double w = 0;
      for (int a = 0; a < lly; a++) {
        double below = 0;
        for (int t = 0; t < a; t++) below += s.n[i * lly + t] + s.n[j * lly + t];
        w += (below + ((double)s.n[i * lly + a] + (double)s.n[j * lly + a] + 1) / 2) * (double)s.n[i * lly + a];
      }
      res += w - half;
    }
  }
  return res;
}

double jt_mean_of(const Stratum &s, int llx) {
  double res = (double)s.nk * (double)s.nk;
  for (int i = 0; i < llx; i++) res -= (double)s.ni[i] * (double)s.ni[i];
  return res / 4;
}

double jt_var_of(const Stratum &s, int llx, int lly) {
  const double n = (double)s.nk;
  double u1 = n * (n - 1) * (2 * n + 5);
  for (int i = 0; i < llx; i++) u1 -= (double)s.ni[i] * (double)(s.ni[i] - 1) * (double)(2 * s.ni[i] + 5);
  for (int j = 0; j < lly; j++) u1 -= (double)s.nj[j] * (double)(s.nj[j] - 1) * (double)(2 * s.nj[j] + 5);
  double u2a = 0, u2b = 0, u3a = 0, u3b = 0;
  for (int i = 0; i < llx; i++) u2a += (double)s.ni[i] * (double)(s.ni[i] - 1) * (double)(s.ni[i] - 2);
  for (int j = 0; j < lly; j++) u2b += (double)s.nj[j] * (double)(s.nj[j] - 1) * (double)(s.nj[j] - 2);
  for (int i = 0; i < llx; i++) u3a += (double)s.ni[i] * (double)(s.ni[i] - 1);
  for (int j = 0; j < lly; j++) u3b += (double)s.nj[j] * (double)(s.nj[j] - 1);
  const double t2 = 36 * n * (n - 1) * (n - 2), t3 = 8 * n * (n - 1);
  return u1 / 72 + (u2a * u2b) / t2 + (u3a * u3b) / t3;
}

// centred statistic: one table as is, strata summed in long double and empty strata skipped.
double jt_centered_of(const std::vector<Stratum> &st, int llx, int lly, bool conditional) {
  if (!conditional) return jt_stat_of(st[0], llx, lly) - jt_mean_of(st[0], llx);
  long double res = 0;
  for (const Stratum &s : st) {
    if (s.nk == 0) continue;
    res += jt_stat_of(s, llx, lly) - jt_mean_of(s, llx);
  }
  return (double)res;
}

double jt_variance_of(const std::vector<Stratum> &st, int llx, int lly, bool conditional) {
  if (!conditional) return jt_var_of(st[0], llx, lly);
  long double res = 0;
  for (const Stratum &s : st) {
    const double v = jt_var_of(s, llx, lly);
    if (!std::isnan(v)) res += v;
  }
  return (double)res;
}

}  // namespace

// splitmix64 -> (0, 1): a uniform stream for callers that have none of their own
// (and the one the parity tests hand to both sides). state: uint64_t *.
extern "C" double bnl_unif_splitmix64(void *state) {
  uint64_t *s = static_cast<uint64_t *>(state);
  for (;;) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    const double u = (double)(z >> 11) * 0x1.0p-53;
    if (u > 0.0) return u;
  }
}

namespace {

bool is_discrete_mc(int t) {
  return t == BNL_MC_MI || t == BNL_MC_X2 || t == BNL_SP_MI || t == BNL_SP_X2 || t == BNL_MC_JT ||
         t == BNL_SMC_MI || t == BNL_SMC_X2 || t == BNL_SMC_JT;
}

int mc_fail(bnl_ctx *ctx, int code, const char *msg, int u) {
  return ctx ? bnl_fail(ctx, code, "%s (test %d)", msg, u) : code;
}

// The permutation stage of c_mcarlo / c_cmcarlo over tables that are already
// counted: tables[off[u]...] is n[z][x][y] of test u.
int mc_from_tables(bnl_ctx *ctx, std::vector<double> &fact, int test_e, int B, double alpha, int n,
                   const int *llxs, const int *llys, const int *conditional_flag, int32_t *tables,
                   const int64_t *off, bnl_unif_fn unif, void *unif_state, double *stat, double *df,
                   double *pvalue) {
  // the reference's callers pass alpha = 1 for the plain Monte Carlo tests (no early stop).
  if (!is_sequential(test_e)) alpha = 1.0;
  const int enough = (int)ceil(alpha * B) + 1;
  // log-factorials up to the largest stratum total (allocfact() in the reference
  // fills them up to the row count for every test).
  {
    long long need = 0;
    for (int u = 0; u < n; u++) {
      long long tot = 0;
      for (int64_t c = off[u]; c < off[u + 1]; c++) tot += tables[(size_t)c];
      need = std::max(need, tot);
    }
    if (fact.empty()) fact.push_back(0.0);
    for (long long k = (long long)fact.size(); k <= need; k++) fact.push_back(lgamma((double)(k + 1)));
  }
  TableSampler sampler;
  sampler.fact = fact.data();
  sampler.unif = unif;
  sampler.ustate = unif_state;

  for (int u = 0; u < n; u++) {
    const int llx = llxs[u], lly = llys[u];
    const bool conditional = conditional_flag[u] != 0;
    const int llz = (int)((off[u + 1] - off[u]) / ((long long)llx * lly));
    std::vector<Stratum> st(llz);
    bool constx = true, consty = true;
    for (int k = 0; k < llz; k++) {
      Stratum &s = st[k];
      s.n = tables + off[u] + (size_t)k * llx * lly;
      s.ni.assign(llx, 0);
      s.nj.assign(lly, 0);
      s.nk = 0;
      for (int i = 0; i < llx; i++)
        for (int j = 0; j < lly; j++) {
          const int c = s.n[i * lly + j];
          s.ni[i] += c; s.nj[j] += c; s.nk += c;
        }
      for (int i = 0; i < llx; i++) constx = constx && (s.ni[i] == 0 || s.ni[i] == s.nk);
      for (int j = 0; j < lly; j++) consty = consty && (s.nj[j] == 0 || s.nj[j] == s.nk);
    }
    if (df) df[u] = NAN;
    // a constant variable (or no complete observation) is independent of anything.
    if (constx || consty) { stat[u] = 0.0; pvalue[u] = 1.0; continue; }
    const bool is_mi = test_e == BNL_MC_MI || test_e == BNL_SMC_MI || test_e == BNL_SP_MI;
    const bool is_x2 = test_e == BNL_MC_X2 || test_e == BNL_SMC_X2 || test_e == BNL_SP_X2;
    const bool semi = test_e == BNL_SP_MI || test_e == BNL_SP_X2;
    auto value = [&]() {
      return is_mi ? mi_of(st, llx, lly) : (is_x2 ? x2_of(st, llx, lly) : jt_centered_of(st, llx, lly, conditional));
    };
    double observed = value();
    double count = 0, edf = 0;
    for (int b = 0; b < B; b++) {
      for (Stratum &s : st) {
        sampler.draw(s, llx, lly);
        if (sampler.failed) return mc_fail(ctx, BNL_EUNSUPPORTED, "random table generation underflowed", u);
      }
      const double v = value();
      if (semi) { edf += v; continue; }
      const bool hit = (is_mi || is_x2) ? v >= observed : fabs(v) >= fabs(observed);
      if (hit) {
        count++;
        if (count >= enough) { count = B; break; }
      }
    }
    if (is_mi) observed = 2 * observed;
    if (!is_mi && !is_x2) observed /= sqrt(jt_variance_of(st, llx, lly, conditional));
    stat[u] = observed;
    if (semi) {
      // degrees of freedom estimated as the mean of the statistic under the null.
      edf = is_mi ? edf * 2 / B : edf / B;
      if (df) df[u] = edf;
      pvalue[u] = bnl::pchisq_upper(observed, edf);
    } else {
      pvalue[u] = count / B;
    }
  }
  return BNL_OK;
}

}  // namespace

// The same over tables the caller holds (rows sharded over GPUs: bnl_counts per
// shard, all-reduce, then this on one rank). Host only: ctx may be NULL.
extern "C" int bnl_mc_from_tables(bnl_ctx *ctx, int test_e, int B, double alpha, int n, const int *llx,
                                  const int *lly, const int *nz, const int32_t *tables,
                                  const int64_t *table_off, bnl_unif_fn unif, void *unif_state,
                                  double *stat, double *df, double *pvalue) {
  if (!is_discrete_mc(test_e)) return mc_fail(ctx, BNL_EUNSUPPORTED, "not a discrete permutation test", -1);
  if (n < 0 || B < 1 || !unif || (n > 0 && (!llx || !lly || !nz || !tables || !table_off || !stat || !pvalue)))
    return mc_fail(ctx, BNL_EINVAL, "bnl_mc_from_tables: bad arguments", -1);
  if (n == 0) return BNL_OK;
  for (int u = 0; u < n; u++)
    if (llx[u] < 1 || lly[u] < 1 || table_off[u + 1] < table_off[u] ||
        (table_off[u + 1] - table_off[u]) % ((int64_t)llx[u] * lly[u]) != 0)
      return mc_fail(ctx, BNL_EINVAL, "bnl_mc_from_tables: bad table shape", u);
  // the random tables are written in place: work on a copy
  std::vector<int32_t> work(tables + table_off[0], tables + table_off[n]);
  std::vector<int64_t> off(n + 1);
  for (int u = 0; u <= n; u++) off[u] = table_off[u] - table_off[0];
  std::vector<double> local;
  return mc_from_tables(ctx, ctx ? ctx->mc_fact : local, test_e, B, alpha, n, llx, lly, nz, work.data(), off.data(),
                        unif, unif_state, stat, df, pvalue);
}

// c_mcarlo / c_cmcarlo for a batch of tests, in the caller's order.
extern "C" int bnl_ci_discrete_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n,
                                  const int *x, const int *y, const int *z_off, const int *z_idx,
                                  bnl_unif_fn unif, void *unif_state, double *stat, double *df,
                                  double *pvalue) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *ds = find_data_p(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (!is_discrete_mc(test_e))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a discrete permutation test", test_e);
  if (n < 0 || B < 1 || !unif || (n > 0 && (!x || !y || !z_off || !stat || !pvalue)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_discrete_mc: bad arguments");
  if (n == 0) return BNL_OK;
  // observed tables of the whole batch: one pass of the counting kernels.
  std::vector<int64_t> off(n + 1, 0);
  std::vector<int> llx(n), lly(n), cond(n);
  for (int u = 0; u < n; u++) {
    auto ok = [&](int c) { return c >= 0 && c < ds->ncols && !ds->is_numeric[c]; };
    if (!ok(x[u]) || !ok(y[u])) return bnl_fail(ctx, BNL_EINVAL, "test %d refers to a column that is not a factor", u);
    llx[u] = ds->nlevels[x[u]]; lly[u] = ds->nlevels[y[u]]; cond[u] = z_off[u + 1] > z_off[u];
    long long cells = (long long)llx[u] * lly[u];
    for (int k = z_off[u]; k < z_off[u + 1]; k++) {
      if (!ok(z_idx[k])) return bnl_fail(ctx, BNL_EINVAL, "test %d refers to a column that is not a factor", u);
      cells *= ds->nlevels[z_idx[k]];
      if (cells >= INT32_MAX) return bnl_fail(ctx, BNL_EOVERFLOW, "attempting to create a factor with more than INT_MAX levels");
    }
    off[u + 1] = off[u] + cells;
  }
  std::vector<int32_t> tables((size_t)off[n]);
  int rc = bnl_counts(ctx, data, n, x, y, z_off, z_idx, tables.data(), off.data());
  if (rc) return rc;
  return mc_from_tables(ctx, ctx->mc_fact, test_e, B, alpha, n, llx.data(), lly.data(), cond.data(), tables.data(),
                        off.data(), unif, unif_state, stat, df, pvalue);
}

// ---------------------------------------------------------------------------
// Gaussian permutation tests (src/tests/permutation/gaussian.monte.carlo.c).
// A permutation of y changes only the cross-products of y with the other
// columns of the test: cross[b][j] = sum_i (c_j[i] - m_j) (y[perm_b[i]] - m_y).
// One CTA per (permutation, column) gathers y through the permutation and
// reduces in a fixed order; everything after that is k x k work on the host.
#define BNL_PERM_THREADS 256

struct PermCols { int slot[BNL_MAXK]; };

__global__ void __launch_bounds__(BNL_PERM_THREADS)
perm_cross_kernel(const double *__restrict__ f64, size_t stride, const double *__restrict__ mean,
                  PermCols cols, int yslot, long long nobs, const int *__restrict__ perm,
                  double *__restrict__ cross, int k) {
  const int b = blockIdx.x, j = blockIdx.y;
  const double *cj = f64 + (size_t)cols.slot[j] * stride;
  const double *cy = f64 + (size_t)yslot * stride;
  const double mj = mean[cols.slot[j]], my = mean[yslot];
  const int *pb = perm + (size_t)b * nobs;
  double s = 0.0, c = 0.0;                            // compensated: long rows, long double in the reference
  for (long long i = threadIdx.x; i < nobs; i += BNL_PERM_THREADS) {
    const double t = (cj[i] - mj) * (cy[pb[i]] - my);
    const double yv = t - c, nv = s + yv;
    c = (nv - s) - yv;
    s = nv;
  }
  __shared__ double red[BNL_PERM_THREADS];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = BNL_PERM_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) cross[(size_t)b * k + j] = red[0];
}

namespace {

// Partial correlation of columns 0 and 1 given the rest, from the covariance
// matrix (column-major k x k): pseudo-inverse through the eigen-decomposition with
// the reference's cut on small singular values (src/core/correlation.c:43-92).
double host_pcor(std::vector<double> a, int k) {
  if (k == 2) {
    const double sxx = a[0], syy = a[3], sxy = a[2];
    double r = (sxx < bnl::kSqrtEps || syy < bnl::kSqrtEps) ? 0.0 : sxy / sqrt(sxx * syy);
    return r > 1.0 ? 1.0 : (r < -1.0 ? -1.0 : r);
  }
  std::vector<double> v((size_t)k * k, 0.0);
  for (int i = 0; i < k; i++) v[i + (size_t)i * k] = 1.0;
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < k; i++) {
      diag += a[i + (size_t)i * k] * a[i + (size_t)i * k];
      for (int j = i + 1; j < k; j++) off += a[i + (size_t)j * k] * a[i + (size_t)j * k];
    }
    if (off == 0.0 || off <= 1e-40 * diag) break;
    for (int r = 0; r < k - 1; r++)
      for (int c = r + 1; c < k; c++) {
        const double apq = a[r + (size_t)c * k];
        if (apq == 0.0) continue;
        const double theta = (a[c + (size_t)c * k] - a[r + (size_t)r * k]) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int m = 0; m < k; m++) {
          const double amr = a[m + (size_t)r * k], amc = a[m + (size_t)c * k];
          a[m + (size_t)r * k] = cs * amr - sn * amc;
          a[m + (size_t)c * k] = sn * amr + cs * amc;
        }
        for (int m = 0; m < k; m++) {
          const double arm = a[r + (size_t)m * k], acm = a[c + (size_t)m * k];
          a[r + (size_t)m * k] = cs * arm - sn * acm;
          a[c + (size_t)m * k] = sn * arm + cs * acm;
        }
        for (int m = 0; m < k; m++) {
          const double vmr = v[m + (size_t)r * k], vmc = v[m + (size_t)c * k];
          v[m + (size_t)r * k] = cs * vmr - sn * vmc;
          v[m + (size_t)c * k] = sn * vmr + cs * vmc;
        }
      }
  }
  double dmax = 0.0;
  for (int i = 0; i < k; i++) dmax = std::max(dmax, fabs(a[i + (size_t)i * k]));
  const double cut = (double)k * dmax * bnl::kSqrtEps * bnl::kSqrtEps;
  double k11 = 0.0, k12 = 0.0, k22 = 0.0;
  for (int i = 0; i < k; i++) {
    const double lam = a[i + (size_t)i * k];
    if (fabs(lam) > cut) {
      const double v0 = v[0 + (size_t)i * k], v1 = v[1 + (size_t)i * k];
      k11 += v0 * v0 / lam; k12 += v0 * v1 / lam; k22 += v1 * v1 / lam;
    }
  }
  double r = (k11 < bnl::kSqrtEps || k22 < bnl::kSqrtEps) ? 0.0 : -k12 / sqrt(k11 * k22);
  return r > 1.0 ? 1.0 : (r < -1.0 ? -1.0 : r);
}

struct DeviceBuf {
  void *p = nullptr;
  ~DeviceBuf() { if (p) cudaFree(p); }
};

}  // namespace

// c_gauss_mcarlo / c_gauss_cmcarlo for a batch of tests on complete data, in the
// caller's order. stat is the reference's htest statistic (bnl_ci_gaussian's).
extern "C" int bnl_ci_gaussian_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n,
                                  const int *x, const int *y, const int *z_off, const int *z_idx,
                                  bnl_unif_fn unif, void *unif_state, double *stat, double *pvalue) {
  return bnl_ci_gaussian_mc_ex(ctx, data, test_e, B, alpha, n, x, y, z_off, z_idx, unif, unif_state, stat, pvalue,
                               nullptr, 0);
}

// The same with extra numeric columns whose missing values also exclude a row
// (rrd_gperm keeps one missingness indicator for the whole call and only ever adds
// to it, src/tests/patterns/roundrobin.test.c:704-732 over tabular_incomplete_cases,
// src/core/data.table.c:405-429: a variable that has been dropped still masks its rows).
int bnl_ci_gaussian_mc_ex(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n, const int *x,
                          const int *y, const int *z_off, const int *z_idx, bnl_unif_fn unif, void *unif_state,
                          double *stat, double *pvalue, const int *mask_cols, int nmask) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *dsp = find_data_p(ctx, data);
  if (!dsp) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = *dsp;
  int base;
  switch (test_e) {
    case BNL_MC_COR: case BNL_SMC_COR: base = BNL_COR; break;
    case BNL_MC_ZF: case BNL_SMC_ZF: base = BNL_ZF; break;
    case BNL_MC_MI_G: case BNL_SMC_MI_G: base = BNL_MI_G; break;
    default: return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a Gaussian permutation test", test_e);
  }
  if (n < 0 || B < 1 || !unif || (n > 0 && (!x || !y || !z_off || !stat || !pvalue)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_gaussian_mc: bad arguments");
  if (n == 0) return BNL_OK;
  if (ds.nobs > INT32_MAX) return bnl_fail(ctx, BNL_EUNSUPPORTED, "too many rows for a permutation test");
  if (ds.gauss_nobs > 0)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "permutation tests need the rows, not pooled moments");
  if (!is_sequential(test_e)) alpha = 1.0;
  const double enough = ceil(alpha * B) + 1;

  // observed statistics of the whole batch from the resident moments.
  std::vector<double> dfv(n);
  int rc = bnl_ci_gaussian_any(ctx, data, base, n, x, y, z_off, z_idx, stat, dfv.data(), pvalue);
  if (rc) return rc;

  BnlDevice &dev = ctx->devs[0];
  BnlDeviceData &dd = ds.dev[0];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  // permutations per device batch: about 32 MB of row indices, and never fewer than two
  // (the first batch holds the identity next to the first drawn permutation)
  const long long pcap = std::max(2ll, std::min<long long>(B + 1, (8ll << 20) / std::max(1ll, ds.nobs)));
  DeviceBuf d_perm, d_cross, d_tmp;
  size_t tmp_doubles = 0;
  int *h_perm = nullptr;
  double *h_cross = nullptr;
  // sized for the worst case: a unit whose complete rows are few gets more permutations per batch
  const size_t perm_ints = (size_t)std::max<long long>(pcap * ds.nobs, (long long)(B + 1) * 64);
  BNL_CUDA(ctx, cudaMalloc(&d_perm.p, perm_ints * sizeof(int)));
  BNL_CUDA(ctx, cudaMalloc(&d_cross.p, (size_t)(B + 1) * BNL_MAXK * sizeof(double)));
  BNL_CUDA(ctx, cudaMallocHost((void **)&h_perm, perm_ints * sizeof(int)));
  BNL_CUDA(ctx, cudaMallocHost((void **)&h_cross, (size_t)(B + 1) * BNL_MAXK * sizeof(double)));
  struct PinnedGuard { int *a; double *b; ~PinnedGuard() { cudaFreeHost(a); cudaFreeHost(b); } } guard{h_perm, h_cross};
  std::vector<int> pool((size_t)ds.nobs);
  const int p = ds.nc;

  for (int u = 0; u < n; u++) {
    const int nz = z_off[u + 1] - z_off[u], k = nz + 2;
    PermCols cols;
    bool incomplete = false;
    for (int i = 0; i < k; i++) {
      const int c = i == 0 ? x[u] : (i == 1 ? y[u] : z_idx[z_off[u] + i - 2]);
      cols.slot[i] = ds.slot[c];          // validated by bnl_ci_gaussian_any above
      incomplete = incomplete || std::isnan(ds.h_mean[ds.slot[c]]);
    }
    for (int m = 0; m < nmask; m++) {
      const int c = mask_cols[m];
      if (c < 0 || c >= ds.ncols || !ds.is_numeric[c])
        return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_gaussian_mc: bad mask column %d", c);
      incomplete = incomplete || std::isnan(ds.h_mean[ds.slot[c]]);
    }
    // what the permutation loop reads: the resident columns and moments, or -- when a
    // column of the test has missing values -- copies of the test's columns restricted
    // to its complete rows, as ut_gperm / ct_gperm / ast_gperm build them
    // (src/tests/patterns/utest.c:368-424), with their own means and covariances.
    long long N = ds.nobs;
    const double *f64 = dd.f64, *dmean = dd.mean;
    size_t fstride = ds.stride_f64;
    int yslot = cols.slot[1];
    std::vector<double> cov((size_t)k * k);
    if (!incomplete) {
      for (int i = 0; i < k; i++)
        for (int j = 0; j < k; j++) cov[i + (size_t)j * k] = ds.h_cov[(size_t)cols.slot[i] + (size_t)cols.slot[j] * p];
    } else {
      std::vector<std::vector<double>> col(k, std::vector<double>((size_t)ds.nobs));
      for (int i = 0; i < k; i++)
        BNL_CUDA(ctx, cudaMemcpy(col[i].data(), dd.f64 + (size_t)cols.slot[i] * ds.stride_f64,
                                 (size_t)ds.nobs * sizeof(double), cudaMemcpyDeviceToHost));
      std::vector<char> masked((size_t)ds.nobs, 0);
      {
        std::vector<double> mc((size_t)ds.nobs);
        for (int m = 0; m < nmask; m++) {
          BNL_CUDA(ctx, cudaMemcpy(mc.data(), dd.f64 + (size_t)ds.slot[mask_cols[m]] * ds.stride_f64,
                                   (size_t)ds.nobs * sizeof(double), cudaMemcpyDeviceToHost));
          for (long long r = 0; r < ds.nobs; r++) masked[(size_t)r] |= std::isnan(mc[(size_t)r]);
        }
      }
      long long nc = 0;
      for (long long r = 0; r < ds.nobs; r++) {
        bool ok = !masked[(size_t)r];
        for (int i = 0; i < k; i++) ok = ok && !std::isnan(col[i][(size_t)r]);
        if (!ok) continue;
        for (int i = 0; i < k; i++) col[i][(size_t)nc] = col[i][(size_t)r];
        nc++;
      }
      if (nc < 2) { stat[u] = 0.0; pvalue[u] = 1.0; continue; }
      std::vector<double> mean(k);
      for (int i = 0; i < k; i++) {
        long double t = 0;
        for (long long r = 0; r < nc; r++) t += col[i][(size_t)r];
        mean[i] = (double)(t / nc);
      }
      for (int i = 0; i < k; i++)
        for (int j = i; j < k; j++) {
          long double t = 0;
          for (long long r = 0; r < nc; r++) t += (col[i][(size_t)r] - mean[i]) * (col[j][(size_t)r] - mean[j]);
          cov[i + (size_t)j * k] = cov[j + (size_t)i * k] = (double)(t / (nc - 1));
        }
      const size_t need = (size_t)k * nc + k;
      if (need > tmp_doubles) {
        if (d_tmp.p) { cudaFree(d_tmp.p); d_tmp.p = nullptr; }
        BNL_CUDA(ctx, cudaMalloc(&d_tmp.p, need * sizeof(double)));
        tmp_doubles = need;
      }
      double *t = static_cast<double *>(d_tmp.p);
      for (int i = 0; i < k; i++)
        BNL_CUDA(ctx, cudaMemcpy(t + (size_t)i * nc, col[i].data(), (size_t)nc * sizeof(double), cudaMemcpyHostToDevice));
      BNL_CUDA(ctx, cudaMemcpy(t + (size_t)k * nc, mean.data(), (size_t)k * sizeof(double), cudaMemcpyHostToDevice));
      f64 = t; fstride = (size_t)nc; dmean = t + (size_t)k * nc;
      for (int i = 0; i < k; i++) cols.slot[i] = i;
      yslot = 1;
      N = nc;
    }
    const long long pmax = std::max(2ll, std::min<long long>(B + 1, (long long)(perm_ints / (size_t)N)));
    if ((size_t)pmax * (size_t)N > perm_ints)
      return bnl_fail(ctx, BNL_ENOMEM, "bnl_ci_gaussian_mc: permutation buffer too small (%lld x %lld)", pmax, N);
    // a constant variable is independent of anything (no draw is consumed).
    if (cov[0] == 0.0 || cov[1 + (size_t)k] == 0.0) { stat[u] = 0.0; pvalue[u] = 1.0; continue; }
    const int kk = nz == 0 ? 1 : k;       // marginal: only the cross-product with x changes
    double observed = 0.0, count = 0.0;
    bool have_observed = false;
    long long done = 0;
    bool stop = false;
    while (done < B && !stop) {
      // never draw more permutations than the reference would before its early exit
      long long want = std::min<long long>(B - done, pmax - (have_observed ? 0 : 1));
      if (is_sequential(test_e)) want = std::min<long long>(want, (long long)(enough - count));
      want = std::max(1ll, want);
      long long np = 0;
      if (!have_observed) {               // the identity first: the observed value by the same arithmetic
        for (long long i = 0; i < N; i++) h_perm[i] = (int)i;
        np = 1;
      }
      for (long long b = 0; b < want; b++, np++) {
        // SampleNoReplace(n, n) (src/core/sampling.c:12-28): draw from a shrinking pool
        int *out = h_perm + (size_t)np * N;
        long long left = N;
        for (long long i = 0; i < N; i++) pool[(size_t)i] = (int)i;
        for (long long i = 0; i < N; i++) {
          const long long j = (long long)((double)left * unif(unif_state));
          out[i] = pool[(size_t)j];
          pool[(size_t)j] = pool[(size_t)--left];
        }
      }
      if ((size_t)np * (size_t)N > perm_ints)
        return bnl_fail(ctx, BNL_ENOMEM, "bnl_ci_gaussian_mc: %lld permutations of %lld rows overrun the buffer", np, N);
      BNL_CUDA(ctx, cudaMemcpyAsync(d_perm.p, h_perm, (size_t)np * N * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
      perm_cross_kernel<<<dim3((unsigned)np, (unsigned)kk), BNL_PERM_THREADS, 0, dev.stream>>>(
          f64, fstride, dmean, cols, yslot, N, (const int *)d_perm.p, (double *)d_cross.p, kk);
      BNL_CUDA(ctx, cudaGetLastError());
      BNL_CUDA(ctx, cudaMemcpyAsync(h_cross, d_cross.p, (size_t)np * kk * sizeof(double), cudaMemcpyDeviceToHost,
                                    dev.stream));
      BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
      for (long long b = 0; b < np; b++) {
        double v;
        if (nz == 0) {
          v = h_cross[b];
        } else {
          std::vector<double> cb = cov;   // row / column of y replaced (c_update_covmat)
          for (int j = 0; j < k; j++) {
            if (j == 1) continue;         // the variance of y does not change under a permutation
            const double c = h_cross[(size_t)b * kk + j] / (double)(N - 1);
            cb[1 + (size_t)j * k] = cb[j + (size_t)k] = c;
          }
          v = host_pcor(cb, k);
        }
        if (!have_observed) { observed = v; have_observed = true; continue; }
        done++;
        if (fabs(v) >= fabs(observed)) {
          count++;
          if (count >= enough) { count = B; stop = true; break; }
        }
      }
    }
    pvalue[u] = count / B;
  }
  return BNL_OK;
}
