// bnl_search.cu -- hill climbing with the score cache kept inside the library:
// one call runs the whole greedy search, so that the per-iteration work outside
// the kernels is bookkeeping in C++ instead of an R (or Python) round trip per
// iteration ("device-resident score cache across hc
// iterations").
//
// Restates the loop of hill.climbing(optimized = TRUE, restart = 0)
// (R/hill-climbing.R:2-318) around the two C entry points it drives:
//   score_cache_fill  src/learning/score/hc.cache.lookup.c:10-128  (bnl_score_cache_fill)
//   hc_opt_step       src/learning/score/hc.cache.lookup.c:131-363 (Search::opt_step below)
// with arcs.to.be.added() -> hc_to_be_added (:373-489) and c_has_path
// (src/graphs/path.c:29-143) for the acyclicity checks.
//
// The reference scans all n^2 cache entries three times per iteration. The
// result of those scans depends on the order (a candidate replaces the incumbent
// only when it is better by more than sqrt(eps), so the first of several close
// candidates wins); the scans here visit candidates in the same order but skip
// what cannot win: rows whose best valid, acyclic addition is not above the
// incumbent (the acyclicity checks are bit tests on a reachability closure kept
// up to date as arcs are added), everything more than two tolerances below the
// best acceptable addition (see opt_step: exact, with a full scan as the way
// out), and, for removals and reversals, everything that is not an arc.
// 1000 nodes: 60 us of scans per iteration became 25.
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "bnl_internal.h"
#include "bnl_special.cuh"
#include "bnl_search_core.h"

namespace {

using bnlsearch::Move;
using bnlsearch::Search;

bool discrete_score(int s) { return s < 100; }

int node_scores(bnl_ctx *ctx, bnl_data data, int score_e, double arg, const Search &S,
                const std::vector<int> &nodes, double *out) {
  std::vector<int> off(1, 0), idx;
  for (int t : nodes) {
    for (int i = 0; i < S.n; i++) if (S.a[(size_t)i + (size_t)t * S.n]) idx.push_back(i);
    off.push_back((int)idx.size());
  }
  const int *pi = idx.empty() ? nullptr : idx.data();
  if (score_e >= BNL_LOGLIK_CG)
    return bnl_score_cg(ctx, data, score_e, arg, (int)nodes.size(), nodes.data(), off.data(), pi, out, nullptr);
  return discrete_score(score_e)
    ? bnl_score_discrete(ctx, data, score_e, arg, (int)nodes.size(), nodes.data(), off.data(), pi, out, nullptr)
    : bnl_score_gaussian(ctx, data, score_e, arg, (int)nodes.size(), nodes.data(), off.data(), pi, out, nullptr);
}

}  // namespace

extern "C" int bnl_hc(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
                      const int *start, const int *blmat, const int *wlmat, double maxp,
                      long long max_iter, int *amat, double *node_scores_out, long long *iterations,
                      double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (nnodes <= 0 || !amat) return bnl_fail(ctx, BNL_EINVAL, "bnl_hc: bad arguments");
  const int n = nnodes;
  Search S(n, maxp, blmat, wlmat);
  if (start)
    for (int j = 0; j < n; j++)
      for (int i = 0; i < n; i++)
        if (start[(size_t)i + (size_t)j * n]) S.set_arc(i, j, 1);
  // score equivalence decides whether score_cache_fill may copy the mirrored
  // delta (is.score.equivalent, R/sanitization-scores.R:74-97): loglik/aic/bic,
  // bde / bge with the uniform prior and the Gaussian forms are; nal/pnal only on
  // complete data; qnml is, fnml, bds, bdj, k2 are not.
  bool equivalence = score_e == BNL_LOGLIK || score_e == BNL_AIC || score_e == BNL_BIC ||
                     score_e == BNL_BDE || score_e == BNL_LOGLIK_G || score_e == BNL_AIC_G ||
                     score_e == BNL_BIC_G || score_e == BNL_BGE || score_e == BNL_QNML || score_e == BNL_EBIC ||
                     score_e == BNL_EBIC_G;
  if (score_e == BNL_NAL || score_e == BNL_PNAL) {
    auto it = ctx->data.find(data);
    if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
    equivalence = true;
    for (int f : it->second.slot_has_na) if (f) equivalence = false;
  }
  // the -cg scores are not score equivalent (the conservative default of
  // is.score.equivalent): equivalence stays false for them.
  std::vector<int> updated(n);
  for (int i = 0; i < n; i++) updated[i] = i;
  int rc = node_scores(ctx, data, score_e, k_or_iss, S, updated, S.reference.data());
  if (rc) return rc;

  long long it = 1;
  for (;;) {
    {
      BNL_TRACE("hc: score_cache_fill");
      rc = bnl_score_cache_fill_np(ctx, data, score_e, k_or_iss, n, S.a.data(), blmat, updated.data(),
                                   (int)updated.size(), equivalence ? 1 : 0, S.reference.data(),
                                   S.cache.data(), test_counter, S.nparents.data());
    }
    if (rc) return rc;
    Move mv;
    {
      { BNL_TRACE("hc: columns_changed"); S.columns_changed(updated); }   // only these columns of the cache were rewritten
      BNL_TRACE("hc: opt_step");
      mv = S.opt_step();
    }
    if (!mv.op) break;
    updated.clear();
    if (mv.op == 1) { S.set_arc(mv.from, mv.to, 1); updated.push_back(mv.to); }
    else if (mv.op == 2) { S.set_arc(mv.from, mv.to, 0); updated.push_back(mv.to); }
    else {
      S.set_arc(mv.from, mv.to, 0); S.set_arc(mv.to, mv.from, 1);
      updated.push_back(mv.from < mv.to ? mv.from : mv.to);
      updated.push_back(mv.from < mv.to ? mv.to : mv.from);
    }
    // a node whose reference score was -Inf (singular model) and that just got a
    // finite delta is rescored from scratch (R/hill-climbing.R:229-256).
    const int ends[2][2] = {{mv.from, mv.to}, {mv.to, mv.from}};
    for (int e = 0; e < (mv.op == 3 ? 2 : 1); e++) {
      const int f = ends[e][0], t = ends[e][1];
      if (S.reference[t] == -INFINITY && isfinite(S.C(f, t))) {
        std::vector<int> one(1, t);
        if ((rc = node_scores(ctx, data, score_e, k_or_iss, S, one, &S.reference[t]))) return rc;
      }
    }
    if (max_iter > 0 && it >= max_iter) break;
    it++;
  }
  memcpy(amat, S.a.data(), sizeof(int) * (size_t)n * n);
  if (node_scores_out) memcpy(node_scores_out, S.reference.data(), sizeof(double) * n);
  if (iterations) *iterations = it;
  return BNL_OK;
}

namespace {

// robust.score.difference() (R/utils-misc.R:225-242) on network scores.
double robust_diff(double new_s, double old_s) {
  if (new_s == -INFINITY && old_s == -INFINITY) return -INFINITY;
  if (new_s != -INFINITY && old_s == -INFINITY) return fabs(new_s);
  const double d = new_s - old_s;
  return fabs(d) < bnl::kSqrtEps ? 0.0 : d;
}

double total(const std::vector<double> &v) {
  double s = 0.0;
  for (double x : v) s += x;
  return s;
}

}  // namespace

extern "C" int bnl_tabu(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
                        const int *start, const int *blmat, const int *wlmat, double maxp,
                        long long max_iter, int tabu, int max_tabu, int *amat,
                        double *node_scores_out, long long *iterations, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (nnodes <= 0 || !amat || tabu < 1 || max_tabu < 0)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_tabu: bad arguments");
  const int n = nnodes;
  Search S(n, maxp, blmat, wlmat);
  if (start)
    for (int j = 0; j < n; j++)
      for (int i = 0; i < n; i++)
        if (start[(size_t)i + (size_t)j * n]) S.set_arc(i, j, 1);
  S.enable_tabu(tabu);
  bool equivalence = score_e == BNL_LOGLIK || score_e == BNL_AIC || score_e == BNL_BIC ||
                     score_e == BNL_BDE || score_e == BNL_LOGLIK_G || score_e == BNL_AIC_G ||
                     score_e == BNL_BIC_G || score_e == BNL_BGE || score_e == BNL_QNML || score_e == BNL_EBIC ||
                     score_e == BNL_EBIC_G;
  if (score_e == BNL_NAL || score_e == BNL_PNAL) {
    auto it = ctx->data.find(data);
    if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
    equivalence = true;
    for (int f : it->second.slot_has_na) if (f) equivalence = false;
  }
  std::vector<int> updated(n);
  for (int i = 0; i < n; i++) updated[i] = i;
  int rc = node_scores(ctx, data, score_e, k_or_iss, S, updated, S.reference.data());
  if (rc) return rc;

  // The loop of tabu.search(optimized = TRUE) (R/tabu.R:66-300): remember the
  // current network, take the best move that does not lead back into a remembered
  // one; when nothing improves, take the least bad move for up to max_tabu
  // iterations in a row, then return the best network seen. (The branch of
  // tabu.R:74-117 that ranks networks with singular nodes is not restated.)
  std::vector<int> best_net = S.a;
  double best_score = -INFINITY;
  int loss_iter = 0;
  long long it = 1;
  bool use_best = false;
  for (;;) {
    const int current = (int)((it - 1) % tabu);
    const double now = total(S.reference);
    if (robust_diff(now, best_score) > 0 || it == 1) { best_net = S.a; best_score = now; }
    S.remember(current);
    rc = bnl_score_cache_fill_np(ctx, data, score_e, k_or_iss, n, S.a.data(), blmat, updated.data(),
                                 (int)updated.size(), equivalence ? 1 : 0, S.reference.data(),
                                 S.cache.data(), test_counter, S.nparents.data());
    if (rc) return rc;
    S.columns_changed(updated);
    Move mv = S.opt_step(0.0);
    if (!mv.op) {
      if (loss_iter >= max_tabu) { use_best = true; break; }
      loss_iter++;
      mv = S.opt_step(-INFINITY);
      if (!mv.op) { use_best = loss_iter > 0; break; }
    } else if (robust_diff(total(S.reference), best_score) > 0) {
      loss_iter = 0;
    }
    updated.clear();
    if (mv.op == 1) { S.set_arc(mv.from, mv.to, 1); updated.push_back(mv.to); }
    else if (mv.op == 2) { S.set_arc(mv.from, mv.to, 0); updated.push_back(mv.to); }
    else {
      S.set_arc(mv.from, mv.to, 0); S.set_arc(mv.to, mv.from, 1);
      updated.push_back(mv.from < mv.to ? mv.from : mv.to);
      updated.push_back(mv.from < mv.to ? mv.to : mv.from);
    }
    if (max_iter > 0 && it >= max_iter) { use_best = loss_iter > 0; break; }
    it++;
  }
  memcpy(amat, use_best ? best_net.data() : S.a.data(), sizeof(int) * (size_t)n * n);
  if (node_scores_out) memcpy(node_scores_out, S.reference.data(), sizeof(double) * n);
  if (iterations) *iterations = it;
  return BNL_OK;
}

// ---------------------------------------------------------------------------
// pc.stable: the skeleton phase in one call (pc.stable.backend, R/pcalgo.R:2-148).
// At every level (size of the conditioning sets) all node pairs see the skeleton
// frozen at the previous level, so the pairs of a level go to bnl_allsubs_test as
// ONE batch: first every pair (x, y) conditioned on subsets of adj(x) \ y, then the
// pairs still connected, as (y, x) on subsets of adj(y) \ x. Neighbour order, the
// level-1 pruning of the second list, the stop rule and the counter are the
// reference's.

extern "C" int bnl_pc_stable(bnl_ctx *ctx, bnl_data data, int test_e, double alpha, int nnodes,
                             int max_sx, int *adj, double *pvalue, int *ndsep, long long *dsep_off,
                             int *dsep_idx, long long dsep_cap, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  if (nnodes < 2 || !adj) return bnl_fail(ctx, BNL_EINVAL, "bnl_pc_stable: bad arguments");
  const int n = nnodes;
  const long long np = (long long)n * (n - 1) / 2;
  if (np > INT32_MAX / 4) return bnl_fail(ctx, BNL_EOVERFLOW, "bnl_pc_stable: too many node pairs");
  auto pid = [&](int i, int j) { return (long long)i * n - (long long)i * (i + 1) / 2 + (j - i - 1); };  // i < j
  std::vector<int> pi((size_t)np), pj((size_t)np);
  for (int i = 0, k = 0; i < n; i++) for (int j = i + 1; j < n; j++, k++) { pi[k] = i; pj[k] = j; }
  std::vector<double> pval((size_t)np, 0.0);
  std::vector<char> has_p((size_t)np, 0), sep((size_t)np, 0), in_skel((size_t)np, 1);
  std::vector<int> max_adj((size_t)np, n - 1);
  std::vector<std::vector<int>> sepset((size_t)np);
  if (max_sx < 0 || max_sx > n - 2) max_sx = n - 2;
  bool gaussian_level0 = test_e == BNL_COR || test_e == BNL_ZF || test_e == BNL_MI_G;
  {
    auto it = ctx->data.find(data);
    if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
    const BnlDataSet &dsr = it->second;
    gaussian_level0 = gaussian_level0 && !dsr.any_nan && n <= dsr.ncols;
    for (int c = 0; c < n && gaussian_level0; c++) gaussian_level0 = dsr.is_numeric[c] != 0;
  }

  for (int level = 0; level <= max_sx; level++) {
    // adjacency lists in the order of R/pcalgo.R:88-91: rows where the node is the
    // second endpoint first, then those where it is the first.
    std::vector<std::vector<int>> nb(n);
    for (long long k = 0; k < np; k++) if (in_skel[k]) nb[pj[k]].push_back(pi[k]);
    for (long long k = 0; k < np; k++) if (in_skel[k]) nb[pi[k]].push_back(pj[k]);
    std::vector<long long> first, second, live;
    std::vector<std::vector<int>> n1((size_t)np), n2((size_t)np);
    std::vector<char> has_info((size_t)np, 0);
    for (long long k = 0; k < np; k++) {
      if (level > max_adj[k] || sep[k]) continue;
      for (int v : nb[pi[k]]) if (v != pj[k]) n1[k].push_back(v);
      for (int v : nb[pj[k]]) if (v != pi[k]) n2[k].push_back(v);
      if ((int)n1[k].size() < level && (int)n2[k].size() < level) { max_adj[k] = 0; continue; }
      has_info[k] = 1;
      live.push_back(k);
      if ((int)n1[k].size() >= level) first.push_back(k);
    }
    auto run = [&](const std::vector<long long> &list, bool swap, std::vector<char> &separated) -> int {
      const int m = (int)list.size();
      if (m == 0) return BNL_OK;
      if (level == 0 && gaussian_level0 && (long long)m == np) {
        // every pair marginally on complete numeric data: pairs enumerated and tested on the
        // device, only the p-value matrix comes back (bnl_ci_gaussian_allpairs). allsubs_test
        // counts such a test twice (rinterface/allsubs.test.c:23-30 on top of indep_test's own).
        std::vector<double> pm((size_t)n * n);
        std::vector<int> cols(n);
        for (int i = 0; i < n; i++) cols[i] = i;
        long long nk = 0;
        int rc = bnl_ci_gaussian_allpairs(ctx, data, test_e, alpha, n, cols.data(), pm.data(), nullptr, nullptr, 0, &nk,
                                          nullptr);
        if (rc) return rc;
        for (long long k : list) {
          const double pv = pm[(size_t)pi[k] * n + pj[k]];
          if (pv > alpha) { pval[k] = pv; has_p[k] = 1; sep[k] = 1; max_adj[k] = 0; separated[k] = 1; sepset[k].clear(); }
        }
        if (test_counter) *test_counter += 2.0 * (double)m;
        return BNL_OK;
      }
      std::vector<int> xs(m), ys(m), off(m + 1, 0), sx, nfix(m, 0), doff(m + 1, 0), nds(m, -1);
      for (int q = 0; q < m; q++) {
        const long long k = list[q];
        xs[q] = swap ? pj[k] : pi[k];
        ys[q] = swap ? pi[k] : pj[k];
        const std::vector<int> &c = swap ? n2[k] : n1[k];
        sx.insert(sx.end(), c.begin(), c.end());
        off[q + 1] = (int)sx.size();
        doff[q + 1] = off[q + 1];
      }
      std::vector<double> out((size_t)m * 3);
      std::vector<int> ds(std::max<size_t>(sx.size(), 1));
      int rc = bnl_allsubs_test(ctx, data, test_e, m, xs.data(), ys.data(), off.data(),
                                sx.empty() ? ds.data() : sx.data(), nfix.data(), level, level, alpha, out.data(),
                                doff.data(), ds.data(), nds.data(), test_counter);
      if (rc) return rc;
      for (int q = 0; q < m; q++) {
        const long long k = list[q];
        if (out[(size_t)q * 3] > alpha) {
          pval[k] = out[(size_t)q * 3]; has_p[k] = 1; sep[k] = 1; max_adj[k] = 0; separated[k] = 1;
          sepset[k].assign(ds.begin() + doff[q], ds.begin() + doff[q] + std::max(nds[q], 0));
        }
      }
      return BNL_OK;
    };
    std::vector<char> separated((size_t)np, 0);
    int rc = run(first, false, separated);
    if (rc) return rc;
    for (long long k : live) {
      if (separated[k]) continue;
      if (level == 1) {               // drop from the second list what the first already tried
        std::vector<int> keep;
        for (int v : n2[k]) if (std::find(n1[k].begin(), n1[k].end(), v) == n1[k].end()) keep.push_back(v);
        n2[k].swap(keep);
      }
      bool same = n1[k].size() == n2[k].size();
      if (same) {
        std::vector<int> a = n1[k], b = n2[k];
        std::sort(a.begin(), a.end()); std::sort(b.begin(), b.end());
        same = a == b;
      }
      if ((int)n2[k].size() >= level && level > 0 && !same) second.push_back(k);
    }
    rc = run(second, true, separated);
    if (rc) return rc;
    for (long long k : live)
      if (!separated[k]) {
        pval[k] = 0.0; has_p[k] = 1;
        max_adj[k] = (int)std::max(n1[k].size(), n2[k].size());
      }
    bool done = true;
    for (long long k = 0; k < np; k++) {
      in_skel[k] = has_p[k] && pval[k] < alpha;
      if (max_adj[k] > level) done = false;
    }
    if (done) break;
  }
  // outputs: symmetric adjacency, p-values, separating sets (CSR over the pairs i < j)
  long long used = 0;
  for (int i = 0; i < n; i++) adj[(size_t)i + (size_t)i * n] = 0;
  for (long long k = 0; k < np; k++) {
    const int i = pi[k], j = pj[k], e = in_skel[k] ? 1 : 0;
    adj[(size_t)i + (size_t)j * n] = adj[(size_t)j + (size_t)i * n] = e;
    if (pvalue) pvalue[(size_t)i + (size_t)j * n] = pvalue[(size_t)j + (size_t)i * n] = pval[k];
    if (ndsep) ndsep[k] = sep[k] ? (int)sepset[k].size() : -1;
    if (dsep_off) dsep_off[k] = used;
    if (sep[k]) {
      if (dsep_idx) {
        if (used + (long long)sepset[k].size() > dsep_cap)
          return bnl_fail(ctx, BNL_EINVAL, "bnl_pc_stable: dsep_idx too small (needs more than %lld entries)", dsep_cap);
        for (int v : sepset[k]) dsep_idx[used++] = v;
      } else used += (long long)sepset[k].size();
    }
  }
  if (dsep_off) dsep_off[np] = used;
  (void)pid;
  return BNL_OK;
}
