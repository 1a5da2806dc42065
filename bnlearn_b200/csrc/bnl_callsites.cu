// bnl_callsites.cu -- host-side drop-in bodies for the reference's call sites
// (include/bnl_b200.h, "call sites"). They turn one reference call into batched
// units for the kernels and then replay, on the host and in the reference's
// order, the logic that consumes the numbers: early exit, min/max p-values,
// separating sets, robust score differences and the test counter.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <memory>
#include <vector>

#include "bnl_internal.h"
#include "bnl_special.cuh"

namespace {

bool is_discrete_test(int t) { return bnl_is_discrete_test(t); }
bool is_gaussian_test(int t) { return t == BNL_COR || t == BNL_ZF || t == BNL_MI_G || t == BNL_MI_G_SH; }
bool is_discrete_score(int s) {
  return s == BNL_LOGLIK || s == BNL_AIC || s == BNL_BIC || s == BNL_EBIC || s == BNL_BDE || s == BNL_MBDE || s == BNL_BDS || s == BNL_BDJ || s == BNL_K2 || s == BNL_NAL || s == BNL_PNAL || s == BNL_FNML ||
         s == BNL_QNML;
}

// Permutation tests reach the call sites through the *_mc entry points below, which
// park the number of permutations, the early-stopping threshold and the caller's
// uniform stream here for the duration of the call (a context is used by one host
// thread at a time; thread_local keeps contexts on other threads apart).
struct McArgs { int B; double alpha; bnl_unif_fn unif; void *state; const int *mask; int nmask; };
thread_local const McArgs *g_mc = nullptr;
bool is_mc_discrete(int t) {
  return t == BNL_MC_MI || t == BNL_MC_X2 || t == BNL_SP_MI || t == BNL_SP_X2 || t == BNL_MC_JT ||
         t == BNL_SMC_MI || t == BNL_SMC_X2 || t == BNL_SMC_JT;
}
bool is_mc_gaussian(int t) {
  return t == BNL_MC_COR || t == BNL_MC_ZF || t == BNL_MC_MI_G || t == BNL_SMC_COR || t == BNL_SMC_ZF ||
         t == BNL_SMC_MI_G;
}
bool is_mc_test(int t) { return g_mc && (is_mc_discrete(t) || is_mc_gaussian(t)); }

// one batch of CI tests through the matching kernel path.
int run_tests(bnl_ctx *ctx, bnl_data data, int test_e, const std::vector<int> &x,
              const std::vector<int> &y, const std::vector<int> &zoff, const std::vector<int> &zidx,
              std::vector<double> &st, std::vector<double> &df, std::vector<double> &pv) {
  const int n = (int)x.size();
  st.assign(n, 0.0); df.assign(n, 0.0); pv.assign(n, 0.0);
  if (n == 0) return BNL_OK;
  const int *zi = zidx.empty() ? nullptr : zidx.data();
  if (is_discrete_test(test_e))
    return bnl_ci_discrete(ctx, data, test_e, n, x.data(), y.data(), zoff.data(), zi, st.data(),
                           df.data(), pv.data());
  if (is_gaussian_test(test_e))
    return bnl_ci_gaussian(ctx, data, test_e, n, x.data(), y.data(), zoff.data(), zi, st.data(),
                           df.data(), pv.data());
  if (test_e == BNL_MI_CG)
    return bnl_ci_cg(ctx, data, n, x.data(), y.data(), zoff.data(), zi, st.data(), df.data(), pv.data());
  // permutation tests: in the order given, which is the order the stream is consumed in
  if (g_mc && is_mc_discrete(test_e))
    return bnl_ci_discrete_mc(ctx, data, test_e, g_mc->B, g_mc->alpha, n, x.data(), y.data(), zoff.data(), zi,
                              g_mc->unif, g_mc->state, st.data(), df.data(), pv.data());
  if (g_mc && is_mc_gaussian(test_e))
    return bnl_ci_gaussian_mc_ex(ctx, data, test_e, g_mc->B, g_mc->alpha, n, x.data(), y.data(), zoff.data(), zi,
                                 g_mc->unif, g_mc->state, st.data(), pv.data(), g_mc->mask, g_mc->nmask);
  return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not handled by this engine", test_e);
}

// lexicographic subsets of {0..m-1} of size k (src/core/sets.c:9-55).
bool next_subset(std::vector<int> &w, int m) {
  const int k = (int)w.size();
  int i = k - 1;
  while (i >= 0 && w[i] == m - k + i) i--;
  if (i < 0) return false;
  w[i]++;
  for (int j = i + 1; j < k; j++) w[j] = w[j - 1] + 1;
  return true;
}

}  // namespace

extern "C" int bnl_indep_test(bnl_ctx *ctx, bnl_data data, int test_e, int nx, const int *x, int y,
                              const int *sx, int nsx, double *pvalue, double *stat, double *df,
                              double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (nx < 0 || nsx < 0 || (nx > 0 && (!x || !pvalue)) || (nsx > 0 && !sx))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_indep_test: bad arguments");
  std::vector<int> xs(x, x + nx), ys(nx, y), zoff(nx + 1), zidx;
  for (int i = 0; i < nx; i++) {
    zoff[i] = (int)zidx.size();
    zidx.insert(zidx.end(), sx, sx + nsx);
  }
  zoff[nx] = (int)zidx.size();
  std::vector<double> st, d, pv;
  int rc = run_tests(ctx, data, test_e, xs, ys, zoff, zidx, st, d, pv);
  if (rc) return rc;
  for (int i = 0; i < nx; i++) pvalue[i] = pv[i];
  if (nx > 0) { if (stat) *stat = st[nx - 1]; if (df) *df = d[nx - 1]; }
  if (test_counter) *test_counter += nx;       // utest.c:95, ctest.c:134
  return BNL_OK;
}

namespace {

// state of one allsubs request while its subsets are being evaluated in rounds.
struct AstReq {
  int x, y, nf, m;                  // m = number of non-fixed candidates
  const int *sx;                    // fixed first, then candidates
  double p = 0.0, pmin = 1.0, pmax = 0.0;
  bool done = false;
  bool stage1 = false;              // finished by the test with no moving parts
  bool missing = false;             // some column of the request holds missing values
  int ndsep = -1;
  std::vector<int> dsep;
  int cursize = 0;
  std::vector<int> subset;          // next subset to evaluate (positions in the candidates)
  bool exhausted_size = true;       // need to move to the next size
  double counter = 0.0;
};

void range(AstReq &r, double p) { r.pmin = std::min(r.pmin, p); r.pmax = std::max(r.pmax, p); }

}  // namespace

extern "C" int bnl_allsubs_test(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x,
                                const int *y, const int *sx_off, const int *sx, const int *nfixed,
                                int minsize, int maxsize, double alpha, double *out,
                                const int *dsep_off, int *dsep, int *ndsep, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (n < 0 || (n > 0 && (!x || !y || !sx_off || !out || !ndsep || !dsep_off)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_allsubs_test: bad arguments");
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  const long long nobs = it->second.gauss_nobs > 0 && is_gaussian_test(test_e) ? it->second.gauss_nobs : it->second.nobs;
  const bool gauss = is_gaussian_test(test_e);
  if (!gauss && !is_discrete_test(test_e) && !is_mc_test(test_e) && test_e != BNL_MI_CG)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not handled by this engine", test_e);
  // A request whose columns are not all complete takes the reference's
  // *_with_missing pattern (rinterface/allsubs.test.c:72, 103-113): every subset is
  // tested on its own complete rows, there is NO early exit on p > alpha (disabled in
  // the source, patterns/allsubs.test.c:419) and no separating set; only a subset
  // without degrees of freedom (or without complete rows) ends the request.
  const BnlDataSet &dsr = it->second;
  auto col_incomplete = [&](int c) {
    return gauss && c >= 0 && c < dsr.ncols && dsr.is_numeric[c] && dsr.slot[c] < (int)dsr.h_mean.size() &&
           std::isnan(dsr.h_mean[dsr.slot[c]]);
  };

  std::vector<AstReq> rq(n);
  for (int r = 0; r < n; r++) {
    rq[r].x = x[r]; rq[r].y = y[r];
    rq[r].nf = nfixed ? nfixed[r] : 0;
    rq[r].sx = sx + sx_off[r];
    rq[r].m = sx_off[r + 1] - sx_off[r] - rq[r].nf;
    if (rq[r].m < 0 || rq[r].nf < 0)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_allsubs_test: request %d has more fixed than total variables", r);
    rq[r].missing = col_incomplete(x[r]) || col_incomplete(y[r]);
    for (int k = sx_off[r]; k < sx_off[r + 1]; k++) rq[r].missing = rq[r].missing || col_incomplete(sx[k]);
  }
  std::vector<int> bx, by, boff, bidx, owner;
  std::vector<double> st, df, pv;

  // (1) minsize == 0: the test with no moving parts goes through indep_test()
  // (rinterface/allsubs.test.c:23-64) and is counted twice.
  if (minsize == 0) {
    boff.assign(1, 0);
    for (int r = 0; r < n; r++) {
      bx.push_back(rq[r].x); by.push_back(rq[r].y);
      bidx.insert(bidx.end(), rq[r].sx, rq[r].sx + rq[r].nf);
      boff.push_back((int)bidx.size());
    }
    int rc = run_tests(ctx, data, test_e, bx, by, boff, bidx, st, df, pv);
    if (rc) return rc;
    for (int r = 0; r < n; r++) {
      AstReq &q = rq[r];
      q.p = pv[r];
      range(q, q.p);
      q.counter += 2.0;
      if (q.p > alpha) {
        q.done = q.stage1 = true;
        q.ndsep = q.nf;
        q.dsep.assign(q.sx, q.sx + q.nf);
      } else if (std::min(maxsize, q.m) <= 0) {
        q.done = q.stage1 = true;
      }
    }
  }
  // marginal-stage ranges are merged at the end (rinterface/allsubs.test.c:195-200);
  // keep them apart from the subset-stage ranges, which restart from (1, 0).
  std::vector<double> m_min(n, 1.0), m_max(n, 0.0), m_p(n, 0.0);
  for (int r = 0; r < n; r++) {
    m_min[r] = rq[r].pmin; m_max[r] = rq[r].pmax; m_p[r] = rq[r].p;
    if (!rq[r].done) { rq[r].pmin = 1.0; rq[r].pmax = 0.0; rq[r].p = 0.0; }
    rq[r].cursize = std::max(1, minsize) - 1;
  }

  // (2) subsets, size by size, in rounds: every live request contributes its
  // next run of subsets (bounded), the batch is evaluated, and each request
  // replays its results in order until its first p-value > alpha.
  // Permutation tests draw from the caller's stream subset after subset: nothing may be
  // evaluated past the early exit (ast_dperm / ast_gperm, patterns/allsubs.test.c:825-1026),
  // and with several requests the stream would have to pass from one to the next.
  const bool mc = is_mc_test(test_e);
  if (mc && n > 1)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "permutation tests take one allsubs request per call");
  const int kPerReqCap = mc ? 1 : (n > 64 ? 64 : 4096);
  for (;;) {
    bx.clear(); by.clear(); bidx.clear(); owner.clear();
    boff.assign(1, 0);
    std::vector<std::vector<int>> planned(n);   // flattened subsets per request, this round
    bool any = false;
    for (int r = 0; r < n; r++) {
      AstReq &q = rq[r];
      if (q.done) continue;
      if (q.exhausted_size) {
        q.cursize++;
        const int mx = std::min(maxsize, q.m);
        if (q.cursize > mx) { q.done = true; continue; }
        if (gauss && !q.missing) {
          // not enough degrees of freedom: independence, with the first
          // cursize + nf variables as separating set (patterns/allsubs.test.c:169-194).
          const int nz = q.cursize + q.nf;
          const double d = test_e == BNL_COR ? (double)(nobs - nz - 2)
                         : test_e == BNL_ZF ? (double)(nobs - nz - 3) : 1.0;
          if (d < 1.0) {
            q.p = q.pmin = q.pmax = 1.0;
            q.done = true;
            q.ndsep = nz;
            q.dsep.assign(q.sx, q.sx + nz);
            continue;
          }
        }
        q.subset.resize(q.cursize);
        for (int i = 0; i < q.cursize; i++) q.subset[i] = i;
        q.exhausted_size = false;
      }
      int planned_n = 0;
      std::vector<int> w = q.subset;
      do {
        bx.push_back(q.x); by.push_back(q.y); owner.push_back(r);
        bidx.insert(bidx.end(), q.sx, q.sx + q.nf);
        for (int i : w) bidx.push_back(q.sx[q.nf + i]);
        boff.push_back((int)bidx.size());
        planned[r].insert(planned[r].end(), w.begin(), w.end());
        planned_n++;
        if (!next_subset(w, q.m)) { q.exhausted_size = true; break; }
      } while (planned_n < kPerReqCap);
      if (!q.exhausted_size) q.subset = w;
      any = true;
    }
    if (!any) break;
    int rc = run_tests(ctx, data, test_e, bx, by, boff, bidx, st, df, pv);
    if (rc) return rc;
    // replay in order.
    size_t u = 0;
    while (u < owner.size()) {
      const int r = owner[u];
      AstReq &q = rq[r];
      size_t e = u;
      while (e < owner.size() && owner[e] == r) e++;
      for (size_t t = u; t < e; t++) {
        if (q.missing) {
          // patterns/allsubs.test.c:332-365: no degrees of freedom on this subset's
          // complete rows ends the request (before the counter moves), with the FIRST
          // cursize + nf variables reported; otherwise the subset is recorded and the
          // scan goes on whatever its p-value.
          if (df[t] == 0.0) {
            q.p = q.pmin = q.pmax = 1.0;
            q.done = true;
            q.ndsep = q.nf + q.cursize;
            q.dsep.assign(q.sx, q.sx + q.ndsep);
            break;
          }
          q.p = pv[t];
          range(q, q.p);
          q.counter += 1.0;
          continue;
        }
        q.p = pv[t];
        range(q, q.p);
        q.counter += 1.0;                        // patterns/allsubs.test.c:110, 249
        if (q.p > alpha) {
          q.done = true;
          q.ndsep = q.nf + q.cursize;
          q.dsep.assign(q.sx, q.sx + q.nf);
          const int *w = planned[r].data() + (t - u) * q.cursize;
          for (int i = 0; i < q.cursize; i++) q.dsep.push_back(q.sx[q.nf + w[i]]);
          break;
        }
      }
      u = e;
    }
  }

  double total = 0.0;
  for (int r = 0; r < n; r++) {
    AstReq &q = rq[r];
    double p = q.p, pmin = q.pmin, pmax = q.pmax;
    // the with-missing scan ends in ast_prepare_retval(p of the LAST subset, ..., NULL, 0):
    // an empty separating set when that p-value happens to exceed alpha.
    if (q.missing && !q.stage1 && q.ndsep < 0 && p > alpha) q.ndsep = 0;
    if (minsize == 0 && !q.stage1) { pmin = std::min(pmin, m_min[r]); pmax = std::max(pmax, m_max[r]); }
    out[3 * r + 0] = p; out[3 * r + 1] = pmin; out[3 * r + 2] = pmax;
    ndsep[r] = q.ndsep;
    if (q.ndsep > 0) {
      if (!dsep || dsep_off[r + 1] - dsep_off[r] < q.ndsep)
        return bnl_fail(ctx, BNL_EINVAL, "bnl_allsubs_test: dsep buffer too small for request %d", r);
      memcpy(dsep + dsep_off[r], q.dsep.data(), (size_t)q.ndsep * sizeof(int));
    }
    total += q.counter;
  }
  if (test_counter) *test_counter += total;
  return BNL_OK;
}

extern "C" int bnl_mi_matrix(bnl_ctx *ctx, bnl_data data, int estimator_e, int ncols,
                             const int *cols, int cond, double *out) {
  if (!ctx) return BNL_EINVAL;
  if (estimator_e != BNL_MI && estimator_e != BNL_MI_G)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "estimator %d is neither mi nor mi-g", estimator_e);
  if (ncols < 0 || (ncols > 0 && !cols) || (ncols > 1 && !out))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_mi_matrix: bad arguments");
  if (estimator_e == BNL_MI_G && cond >= 0)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "a conditioning variable is only defined for the discrete estimator");
  const long long np = (long long)ncols * (ncols - 1) / 2;
  if (np <= 0) return BNL_OK;
  if (np > INT32_MAX) return bnl_fail(ctx, BNL_EOVERFLOW, "too many column pairs");
  // every pair is one unit; the batch goes through the same kernels as a level of
  // pc.stable (the counting kernel groups units by table size, not by caller).
  std::vector<int> x((size_t)np), y((size_t)np), zoff((size_t)np + 1), zidx;
  if (cond >= 0) zidx.assign((size_t)np, cond);
  size_t k = 0;
  for (int i = 0; i < ncols; i++)
    for (int j = i + 1; j < ncols; j++, k++) {
      x[k] = cols[i]; y[k] = cols[j];
      zoff[k] = cond >= 0 ? (int)k : 0;
    }
  zoff[(size_t)np] = cond >= 0 ? (int)np : 0;
  std::vector<double> df((size_t)np), pv((size_t)np);
  if (estimator_e == BNL_MI)
    return bnl_ci_discrete_any(ctx, data, BNL_MI_COEF, (int)np, x.data(), y.data(), zoff.data(),
                               zidx.data(), out, df.data(), pv.data());
  return bnl_ci_gaussian_any(ctx, data, BNL_MI_G_COEF, (int)np, x.data(), y.data(), zoff.data(),
                             zidx.data(), out, df.data(), pv.data());
}

extern "C" int bnl_roundrobin_test(bnl_ctx *ctx, bnl_data data, int test_e, int x, const int *z,
                                   int nz, const int *is_fixed, double alpha, double *pvalue,
                                   double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (nz < 0 || (nz > 0 && (!z || !pvalue)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_roundrobin_test: bad arguments");
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  const long long nobs = it->second.gauss_nobs > 0 && is_gaussian_test(test_e) ? it->second.gauss_nobs : it->second.nobs;
  const bool gauss = is_gaussian_test(test_e);
  if (!gauss && !is_discrete_test(test_e) && !is_mc_test(test_e) && test_e != BNL_MI_CG)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not handled by this engine", test_e);
  // columns with missing values: rrd_gaustests_with_missing (each test on the rows
  // complete for the variables still in play; a test without degrees of freedom gives
  // p = 1 and does NOT drop its variable, patterns/roundrobin.test.c:264-276).
  const BnlDataSet &dsr = it->second;
  bool with_missing = false;
  if (gauss) {
    auto inc = [&](int c) {
      return c >= 0 && c < dsr.ncols && dsr.is_numeric[c] && dsr.slot[c] < (int)dsr.h_mean.size() &&
             std::isnan(dsr.h_mean[dsr.slot[c]]);
    };
    with_missing = inc(x);
    for (int i = 0; i < nz; i++) with_missing = with_missing || inc(z[i]);
  }

  int nfixed = 0;
  for (int i = 0; i < nz; i++) nfixed += is_fixed && is_fixed[i];
  for (int i = 0; i < nz - nfixed; i++) pvalue[i] = 0.0;   // rinterface/roundrobin.test.c:29
  std::vector<char> dropped(nz, 0);
  std::vector<int> xs(1), ys(1), zoff(2), zidx;
  std::vector<double> st, df, pv;
  int valid = nz, cur = 0;

  // sequentially dependent: each step's conditioning set depends on the
  // verdicts before it (patterns/roundrobin.test.c:92-96, 199-206).
  for (int i = 0; i < nz; i++) {
    if (valid < 2) break;
    if (is_fixed && is_fixed[i]) continue;
    zidx.clear();
    for (int j = 0; j < nz; j++)
      if (j != i && !dropped[j]) zidx.push_back(z[j]);
    if (gauss && !with_missing) {
      const int ncond = (int)zidx.size();
      const double d = test_e == BNL_COR ? (double)(nobs - ncond - 2)
                     : test_e == BNL_ZF ? (double)(nobs - ncond - 3) : 1.0;
      if (d < 1.0) { pvalue[cur++] = 1.0; continue; }   // no drop (roundrobin.test.c:144-154)
    }
    xs[0] = x; ys[0] = z[i];
    zoff[0] = 0; zoff[1] = (int)zidx.size();
    int rc = run_tests(ctx, data, test_e, xs, ys, zoff, zidx, st, df, pv);
    if (rc) return rc;
    // df == 0 comes back when the test's own complete rows leave no degrees of freedom
    if (with_missing && df[0] == 0.0) { pvalue[cur++] = 1.0; continue; }
    pvalue[cur] = pv[0];
    if (pvalue[cur++] > alpha) { valid--; dropped[i] = 1; }
  }
  if (test_counter) *test_counter += nz - nfixed;        // rinterface/roundrobin.test.c:150
  return BNL_OK;
}

extern "C" int bnl_score_cache_fill(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss,
                                    int nnodes, const int *amat, const int *blmat,
                                    const int *updated, int nupdated, int equivalence,
                                    const double *reference, double *cache,
                                    double *test_counter) {
  return bnl_score_cache_fill_np(ctx, data, score_e, k_or_iss, nnodes, amat, blmat, updated, nupdated,
                                 equivalence, reference, cache, test_counter, nullptr);
}

// The same with the parent counts supplied by a caller that tracks them (bnl_hc),
// which saves the n^2 column sums per call.
int bnl_score_cache_fill_np(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
                            const int *amat, const int *blmat, const int *updated, int nupdated,
                            int equivalence, const double *reference, double *cache,
                            double *test_counter, const int *nparents) {
  if (!ctx) return BNL_EINVAL;
  if (nnodes <= 0 || !amat || !reference || !cache || nupdated < 0 || (nupdated > 0 && !updated))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_score_cache_fill: bad arguments");
  if (nupdated == 0) return BNL_OK;
  const size_t n = (size_t)nnodes;
  auto A = [&](int i, int j) { return amat[(size_t)i + (size_t)j * n]; };
  auto B = [&](int i, int j) { return blmat ? blmat[(size_t)i + (size_t)j * n] : 0; };

  std::unique_ptr<BnlTraceScope> tr_prep(new BnlTraceScope("cache_fill: candidate list"));
  std::vector<char> upd(nnodes, 0);
  for (int k = 0; k < nupdated; k++) {
    if (updated[k] < 0 || updated[k] >= nnodes)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_score_cache_fill: updated node %d out of range", updated[k]);
    upd[updated[k]] = 1;
  }
  std::vector<int> colsum(nnodes, 0);
  if (equivalence) {
    if (nparents) colsum.assign(nparents, nparents + nnodes);
    else
      for (int j = 0; j < nnodes; j++)
        for (int i = 0; i < nnodes; i++) colsum[j] += A(i, j);
  }

  // The candidates that are really scored, target by target (j outer, i inner; the reference
  // goes i outer): amat, blmat and the cache are column-major, so a target's column is read
  // and written in order, and the candidates of one node are consecutive, which is what the
  // counting kernels' run detection wants when two nodes are updated after a reversal. A
  // delta does not depend on its place in the batch; the equivalence copies are applied
  // afterwards. The lists live in the context between calls (BnlCacheFillScratch).
  std::vector<int> &target = ctx->cf.target, &pa_off = ctx->cf.pa_off, &pa_idx = ctx->cf.pa_idx, &ci = ctx->cf.ci,
                   &cj = ctx->cf.cj;
  std::vector<std::pair<int, int>> &copies = ctx->cf.copies;
  target.clear(); pa_idx.clear(); ci.clear(); cj.clear(); copies.clear();
  pa_off.assign(1, 0);
  std::vector<std::vector<int>> parents(nnodes);
  std::vector<int> ulist;                       // updated nodes in increasing order
  for (int j = 0; j < nnodes; j++) if (upd[j]) ulist.push_back(j);
  for (int j : ulist) {
    const int *acol = amat + (size_t)j * n;
    for (int i = 0; i < nnodes; i++)
      if (acol[i]) parents[j].push_back(i);
  }
  {                                             // the first iteration of hc on 1000 nodes lists ~500 000 candidates
    size_t cap = ulist.size() * (size_t)nnodes, npar = 0;
    for (int j : ulist) npar += parents[j].size() + 1;
    target.reserve(cap); pa_off.reserve(cap + 1); ci.reserve(cap); cj.reserve(cap);
    pa_idx.reserve(npar * (size_t)nnodes);
    if (equivalence) copies.reserve(cap / 2 + 1);
  }
  for (int j : ulist) {
    const int *acol = amat + (size_t)j * n;
    const int *bcol = blmat ? blmat + (size_t)j * n : nullptr;
    const std::vector<int> &pj = parents[j];
    const bool j_bare = equivalence && colsum[j] == 0;
    for (int i = 0; i < nnodes; i++) {
      if (i == j) continue;
      if (bcol && bcol[i] == 1) continue;
      if (j_bare && i > j && colsum[i] == 0 && B(j, i) == 0) {
        copies.push_back(std::make_pair(i, j));
        continue;
      }
      target.push_back(j);
      if (acol[i] == 0) {          // "set": old parents, then the new one last
        pa_idx.insert(pa_idx.end(), pj.begin(), pj.end());
        pa_idx.push_back(i);
      } else {                     // "drop"
        for (int p : pj) if (p != i) pa_idx.push_back(p);
      }
      pa_off.push_back((int)pa_idx.size());
      ci.push_back(i); cj.push_back(j);
    }
  }
  const int m = (int)target.size();
  std::vector<double> &score = ctx->cf.score;
  if ((int)score.size() < std::max(m, 1)) score.resize(std::max(m, 1));
  tr_prep.reset();
  if (m > 0) {
    BNL_TRACE("cache_fill: score call");
    const int *pi = pa_idx.empty() ? nullptr : pa_idx.data();
    int rc = score_e >= BNL_LOGLIK_CG
      ? bnl_score_cg(ctx, data, score_e, k_or_iss, m, target.data(), pa_off.data(), pi, score.data(), nullptr)
      : is_discrete_score(score_e)
      ? bnl_score_discrete(ctx, data, score_e, k_or_iss, m, target.data(), pa_off.data(), pi, score.data(), nullptr)
      : bnl_score_gaussian(ctx, data, score_e, k_or_iss, m, target.data(), pa_off.data(), pi, score.data(), nullptr);
    if (rc) return rc;
  }
  BNL_TRACE("cache_fill: deltas into the cache");
  const double tol = bnl::kSqrtEps;
  for (int u = 0; u < m; u++) {
    // robust_score_difference(): tiny differences are zero; -Inf - -Inf is -Inf.
    const double old_s = reference[cj[u]], new_s = score[u];
    double delta = fabs(new_s - old_s) < tol ? 0.0 : new_s - old_s;
    if (old_s == -INFINITY) delta = (new_s != -INFINITY) ? new_s : -INFINITY;
    cache[(size_t)ci[u] + (size_t)cj[u] * n] = delta;
  }
  for (auto &c : copies)
    cache[(size_t)c.first + (size_t)c.second * n] = cache[(size_t)c.second + (size_t)c.first * n];
  if (test_counter) *test_counter += m;                 // score.delta.c:278
  return BNL_OK;
}

// ---------------------------------------------------------------------------
// mi-cg: the conditional Gaussian mutual information test on mixed data
// (ut_micg_complete, src/tests/patterns/utest.c:195-273; c_micg_chisq and
// ct_micg_complete, src/tests/patterns/ctest.c:232-334; c_micg / c_cmicg /
// c_cmicg_unroll, src/tests/conditional.gaussian/cg.mutual.information.c). Every
// case is a difference of conditional Gaussian log-likelihoods L(node | numeric
// regressors, discrete configuration) -- the units of bnl_score_cg -- plus, when x
// and y are both factors, the discrete mutual information. All the
// log-likelihoods of a call are evaluated in one batch, then combined here.

namespace {

struct CgTerms {
  std::vector<int> target, off, idx;       // score units (CSR parents)
  CgTerms() : off(1, 0) {}
  int add(int t, const std::vector<int> &numeric, const std::vector<int> &discrete) {
    target.push_back(t);
    idx.insert(idx.end(), numeric.begin(), numeric.end());
    idx.insert(idx.end(), discrete.begin(), discrete.end());
    off.push_back((int)idx.size());
    return (int)target.size() - 1;
  }
};

}  // namespace

extern "C" int bnl_ci_cg(bnl_ctx *ctx, bnl_data data, int n, const int *x, const int *y, const int *z_off,
                         const int *z_idx, double *stat, double *df, double *pvalue) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (n < 0 || (n > 0 && (!x || !y || !z_off || !stat)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_cg: bad arguments");
  if (ds.gauss_nobs > 0)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "mi-cg needs the rows, not pooled moments (bnl_data_set_moments)");
  // Incomplete data: ct_micg_with_missing (src/tests/patterns/ctest.c:336-406) runs the
  // complete-data test on the rows that are complete for ALL columns of the test. Such a
  // test gets a temporary data set of exactly those rows (device-side row gather,
  // bnl_data_resample) and goes through the code below on it; tests whose columns are
  // all complete stay in one batch.
  {
    auto col_incomplete = [&](int c) {
      if (c < 0 || c >= ds.ncols) return false;          // reported by the checks below
      if (ds.is_numeric[c]) return ds.slot[c] < (int)ds.h_mean.size() && std::isnan(ds.h_mean[ds.slot[c]]);
      if (ds.is_wide[c]) return !ds.wide_has_na.empty() && ds.wide_has_na[ds.slot[c]] != 0;
      return !ds.slot_has_na.empty() && ds.slot_has_na[ds.slot[c]] != 0;
    };
    std::vector<int> inc, comp;
    for (int u = 0; u < n; u++) {
      bool bad = col_incomplete(x[u]) || col_incomplete(y[u]);
      for (int j = z_off[u]; j < z_off[u + 1]; j++) bad = bad || col_incomplete(z_idx[j]);
      (bad ? inc : comp).push_back(u);
    }
    if (!inc.empty()) {
      int rc;
      if (!comp.empty()) {
        std::vector<int> cx, cy, coff(1, 0), cidx;
        for (int u : comp) {
          cx.push_back(x[u]); cy.push_back(y[u]);
          cidx.insert(cidx.end(), z_idx + z_off[u], z_idx + z_off[u + 1]);
          coff.push_back((int)cidx.size());
        }
        std::vector<double> s1(comp.size()), d1(comp.size()), p1(comp.size());
        if ((rc = bnl_ci_cg(ctx, data, (int)comp.size(), cx.data(), cy.data(), coff.data(),
                            cidx.empty() ? nullptr : cidx.data(), s1.data(), d1.data(), p1.data()))) return rc;
        for (size_t i = 0; i < comp.size(); i++) {
          stat[comp[i]] = s1[i];
          if (df) df[comp[i]] = d1[i];
          if (pvalue) pvalue[comp[i]] = p1[i];
        }
      }
      BnlDeviceData &dd = ds.dev[0];
      BNL_CUDA(ctx, cudaSetDevice(ctx->devs[0].id));
      const size_t nrows = (size_t)ds.nobs;
      std::vector<char> bad(nrows);
      std::vector<uint8_t> b8(nrows);
      std::vector<uint16_t> b16(nrows);
      std::vector<double> bf(nrows);
      std::vector<int> rows;
      for (int u : inc) {
        std::fill(bad.begin(), bad.end(), 0);
        auto mark = [&](int c) -> int {
          if (c < 0 || c >= ds.ncols) return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_cg: test %d refers to a column out of range", u);
          if (ds.is_numeric[c]) {
            BNL_CUDA(ctx, cudaMemcpy(bf.data(), dd.f64 + (size_t)ds.slot[c] * ds.stride_f64, nrows * sizeof(double),
                                     cudaMemcpyDeviceToHost));
            for (size_t r = 0; r < nrows; r++) bad[r] |= std::isnan(bf[r]);
          } else if (ds.is_wide[c]) {
            BNL_CUDA(ctx, cudaMemcpy(b16.data(), dd.u16 + (size_t)ds.slot[c] * ds.stride_u16, nrows * sizeof(uint16_t),
                                     cudaMemcpyDeviceToHost));
            for (size_t r = 0; r < nrows; r++) bad[r] |= b16[r] >= 0x8000u;
          } else {
            BNL_CUDA(ctx, cudaMemcpy(b8.data(), dd.u8 + (size_t)ds.slot[c] * ds.stride_u8, nrows, cudaMemcpyDeviceToHost));
            for (size_t r = 0; r < nrows; r++) bad[r] |= b8[r] >= 0x80u;
          }
          return BNL_OK;
        };
        if ((rc = mark(x[u])) || (rc = mark(y[u]))) return rc;
        for (int j = z_off[u]; j < z_off[u + 1]; j++) if ((rc = mark(z_idx[j]))) return rc;
        rows.clear();
        for (size_t r = 0; r < nrows; r++) if (!bad[r]) rows.push_back((int)r);
        if (rows.empty()) {                   // no complete observation: independence
          stat[u] = 0.0;
          if (df) df[u] = 0.0;
          if (pvalue) pvalue[u] = 1.0;
          continue;
        }
        bnl_data tmp = 0;
        if ((rc = bnl_data_resample(ctx, data, (int64_t)rows.size(), rows.data(), &tmp))) return rc;
        const int nzu = z_off[u + 1] - z_off[u], off2[2] = {0, nzu};
        double s1 = 0.0, d1 = 0.0, p1 = 0.0;
        rc = bnl_ci_cg(ctx, tmp, 1, &x[u], &y[u], off2, nzu ? z_idx + z_off[u] : nullptr, &s1, &d1, &p1);
        bnl_data_release(ctx, tmp);
        if (rc) return rc;
        stat[u] = s1;
        if (df) df[u] = d1;
        if (pvalue) pvalue[u] = p1;
      }
      return BNL_OK;
    }
  }
  const double N = (double)ds.nobs;

  // plan: what each test needs
  struct Plan {
    int kind;                 // 0 discrete MI, 1 Gaussian (both numeric, no z), 2 difference of two terms, 3 unroll
    int mi = -1;              // index into the discrete-MI batch
    int num = -1, den = -1;   // kind 2: score-unit indices
    std::vector<int> a, b, c, d;   // kind 3: per numeric z: lognum_i, logden_i (numerator), lognum'_i, logden'_i
    double df = 0.0;
  };
  std::vector<Plan> plan(n);
  CgTerms T;
  std::vector<int> mx, my, moff(1, 0), midx;        // discrete MI units (statistic = G^2)
  std::vector<int> gx, gy;                          // plain Gaussian mi-g units
  std::vector<int> gwho;
  for (int u = 0; u < n; u++) {
    const int xu = x[u], yu = y[u];
    if (xu < 0 || xu >= ds.ncols || yu < 0 || yu >= ds.ncols)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_cg: test %d refers to a column out of range", u);
    std::vector<int> zd, zc;
    long long llz = 0;
    for (int j = z_off[u]; j < z_off[u + 1]; j++) {
      const int c = z_idx[j];
      if (c < 0 || c >= ds.ncols) return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_cg: test %d refers to a column out of range", u);
      if (ds.is_numeric[c]) zc.push_back(c);
      else { zd.push_back(c); llz = (llz == 0 ? 1 : llz) * ds.nlevels[c]; }
    }
    const bool xd = !ds.is_numeric[xu], yd = !ds.is_numeric[yu];
    Plan &P = plan[u];
    const bool marginal = z_off[u + 1] == z_off[u];
    if (xd && yd) {
      P.mi = (int)mx.size();
      mx.push_back(xu); my.push_back(yu);
      midx.insert(midx.end(), zd.begin(), zd.end());
      moff.push_back((int)midx.size());
      if (zc.empty()) { P.kind = 0; continue; }
      // c_cmicg_unroll: chain rule over the numeric conditioning variables.
      P.kind = 3;
      std::vector<int> ydz(1, yu);
      ydz.insert(ydz.end(), zd.begin(), zd.end());
      for (size_t i = 0; i < zc.size(); i++) {
        std::vector<int> rest(zc.begin() + i + 1, zc.end());
        P.a.push_back(T.add(zc[i], rest, std::vector<int>(1, xu)));   // numerator: alternative, configurations of x
        P.b.push_back(T.add(zc[i], rest, zd));                        // numerator: null, configurations of Zd
        P.c.push_back(P.a.back());                                    // denominator: alternative (the same model)
        P.d.push_back(T.add(zc[i], rest, ydz));                       // denominator: null, configurations of (y, Zd)
      }
      const double ngp = (double)zc.size();
      // llz stays 0 in ct_micg_complete when there is no discrete conditioning variable.
      P.df = (double)llz * (ngp * (ngp + 3.0) / 2.0) * (double)(ds.nlevels[yu] - 1);
    } else if (!xd && !yd) {
      if (marginal) { P.kind = 1; gwho.push_back(u); gx.push_back(xu); gy.push_back(yu); continue; }
      P.kind = 2;
      std::vector<int> full(1, xu);
      full.insert(full.end(), zc.begin(), zc.end());
      P.den = T.add(yu, zc, zd);
      P.num = T.add(yu, full, zd);
      P.df = llz == 0 ? 1.0 : (double)llz;
    } else {
      // one factor, one numeric column: the numeric one is the response, the
      // factor joins the discrete configuration under the alternative.
      const int resp = xd ? yu : xu, fac = xd ? xu : yu;
      std::vector<int> alt(1, fac);
      alt.insert(alt.end(), zd.begin(), zd.end());
      P.kind = 2;
      P.den = T.add(resp, zc, zd);
      P.num = T.add(resp, zc, alt);
      P.df = marginal ? 2.0 * (double)(ds.nlevels[fac] - 1)
                      : (double)(ds.nlevels[fac] - 1) * (llz == 0 ? 1.0 : (double)llz) * (double)(zc.size() + 2);
    }
  }
  int rc;
  std::vector<double> L(T.target.size()), mst(mx.size()), mdf(mx.size()), mpv(mx.size());
  std::vector<double> gst(gx.size()), gdf(gx.size()), gpv(gx.size());
  if (!T.target.empty() &&
      (rc = bnl_score_cg(ctx, data, BNL_LOGLIK_CG, 0.0, (int)T.target.size(), T.target.data(), T.off.data(),
                         T.idx.empty() ? nullptr : T.idx.data(), L.data(), nullptr))) return rc;
  if (!mx.empty() &&
      (rc = bnl_ci_discrete(ctx, data, BNL_MI, (int)mx.size(), mx.data(), my.data(), moff.data(),
                            midx.empty() ? nullptr : midx.data(), mst.data(), mdf.data(), mpv.data()))) return rc;
  if (!gx.empty()) {
    std::vector<int> zo(gx.size() + 1, 0);
    if ((rc = bnl_ci_gaussian(ctx, data, BNL_MI_G, (int)gx.size(), gx.data(), gy.data(), zo.data(), nullptr,
                              gst.data(), gdf.data(), gpv.data()))) return rc;
  }
  size_t gi = 0;
  for (int u = 0; u < n; u++) {
    const Plan &P = plan[u];
    double s, d;
    if (P.kind == 0) { s = mst[P.mi]; d = mdf[P.mi]; }
    else if (P.kind == 1) { s = gst[gi]; d = gdf[gi]; gi++; }
    else if (P.kind == 2) {
      const double num = L[P.num], den = L[P.den];
      // a singular null model makes the alternative even more singular: independence.
      s = (isfinite(num) && isfinite(den)) ? 2.0 * N * ((num - den) / N) : 0.0;
      d = P.df;
    } else {
      // c_cmicg_unroll: every chain-rule term is (alternative - null) / n, or 0 when
      // either model is singular; the remainder term is the discrete MI per row.
      double lognum = mst[P.mi] / (2.0 * N), logden = 0.0;
      for (size_t i = 0; i < P.a.size(); i++) {
        const double a = L[P.a[i]], b = L[P.b[i]], c = L[P.c[i]], dd = L[P.d[i]];
        lognum += (isfinite(a) && isfinite(b)) ? (a - b) / N : 0.0;
        logden += (isfinite(c) && isfinite(dd)) ? (c - dd) / N : 0.0;
      }
      const double r = (isfinite(lognum) && isfinite(logden)) ? (lognum - logden) / N : 0.0;
      s = 2.0 * N * N * r;
      d = mdf[P.mi] + P.df;
    }
    stat[u] = s;
    if (df) df[u] = d;
    if (pvalue) pvalue[u] = bnl::pchisq_upper(s, d);
  }
  return BNL_OK;
}

// ---- the same call sites with a permutation test (mc-*, smc-*, sp-*) ----------
// B permutations, the early-stopping threshold of the sequential tests (the
// reference passes its alpha there and 1 otherwise; only smc-* use it) and the
// caller's uniform stream; everything else as in the plain entry points.

extern "C" int bnl_indep_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, bnl_unif_fn unif,
                                 void *unif_state, int nx, const int *x, int y, const int *sx, int nsx,
                                 double *pvalue, double *stat, double *df, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (!(is_mc_discrete(test_e) || is_mc_gaussian(test_e)))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a permutation test", test_e);
  const McArgs a{B, alpha, unif, unif_state, nullptr, 0};
  g_mc = &a;
  const int rc = bnl_indep_test(ctx, data, test_e, nx, x, y, sx, nsx, pvalue, stat, df, test_counter);
  g_mc = nullptr;
  return rc;
}

extern "C" int bnl_allsubs_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, bnl_unif_fn unif,
                                   void *unif_state, int x, int y, const int *sx, int nsx, int nfixed,
                                   int minsize, int maxsize, double alpha, double *out, int *dsep, int *ndsep,
                                   double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (!(is_mc_discrete(test_e) || is_mc_gaussian(test_e)))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a permutation test", test_e);
  const McArgs a{B, alpha, unif, unif_state, nullptr, 0};
  const int off[2] = {0, nsx};
  g_mc = &a;
  const int rc = bnl_allsubs_test(ctx, data, test_e, 1, &x, &y, off, sx, &nfixed, minsize, maxsize, alpha, out,
                                  off, dsep, ndsep, test_counter);
  g_mc = nullptr;
  return rc;
}

extern "C" int bnl_roundrobin_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, bnl_unif_fn unif,
                                      void *unif_state, int x, const int *z, int nz, const int *is_fixed,
                                      double alpha, double *pvalue, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (!(is_mc_discrete(test_e) || is_mc_gaussian(test_e)))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a permutation test", test_e);
  // Gaussian permutation tests: every column of the request masks its missing rows for
  // the whole call (rrd_gperm's indicator is never cleared), dropped variables included.
  std::vector<int> mask;
  if (is_mc_gaussian(test_e) && nz > 0 && z) { mask.push_back(x); mask.insert(mask.end(), z, z + nz); }
  const McArgs a{B, alpha, unif, unif_state, mask.empty() ? nullptr : mask.data(), (int)mask.size()};
  g_mc = &a;
  const int rc = bnl_roundrobin_test(ctx, data, test_e, x, z, nz, is_fixed, alpha, pvalue, test_counter);
  g_mc = nullptr;
  return rc;
}
