// bnl_search.cu -- hill climbing with the score cache kept inside the library:
// one call runs the whole greedy search, so that the per-iteration work outside
// the kernels is a few O(n^2) scans in C++ instead of an R (or Python) round
// trip per iteration ("device-resident score cache
// across hc iterations").
//
// Restates the loop of hill.climbing(optimized = TRUE, restart = 0)
// (R/hill-climbing.R:2-318) around the two C entry points it drives:
//   score_cache_fill  src/learning/score/hc.cache.lookup.c:10-128  (bnl_score_cache_fill)
//   hc_opt_step       src/learning/score/hc.cache.lookup.c:131-363 (opt_step below)
// with arcs.to.be.added() -> hc_to_be_added (:373-489) and c_has_path
// (src/graphs/path.c:29-143) for the acyclicity checks.
#include <math.h>
#include <string.h>

#include <vector>

#include "bnl_internal.h"
#include "bnl_special.cuh"

namespace {

// The adjacency matrix in both layouts: the scans below run with i outer and j
// inner (the reference's order decides ties), and they need g(i, j) and g(j, i)
// in the same inner loop -- one is contiguous in each layout.
struct Graph {
  int n;
  std::vector<int> a;                      // column-major, a[i + j n] = 1 for i -> j (the .Call layout)
  std::vector<int> r;                      // row-major copy, r[i n + j]
  int at(int i, int j) const { return r[(size_t)i * n + j]; }
  void set(int i, int j, int v) { a[(size_t)i + (size_t)j * n] = v; r[(size_t)i * n + j] = v; }
};

// directed path start ~> stop; `skip_direct` ignores the arc start -> stop itself.
bool has_path(const Graph &g, int start, int stop, bool skip_direct, std::vector<int> &stack,
              std::vector<char> &seen) {
  std::fill(seen.begin(), seen.end(), 0);
  stack.clear();
  stack.push_back(start);
  seen[start] = 1;
  while (!stack.empty()) {
    const int v = stack.back();
    stack.pop_back();
    const int *row = g.r.data() + (size_t)v * g.n;
    for (int w = 0; w < g.n; w++) {
      if (!row[w]) continue;
      if (skip_direct && v == start && w == stop) continue;
      if (w == stop) return true;
      if (!seen[w]) { seen[w] = 1; stack.push_back(w); }
    }
  }
  return false;
}

struct Move { int op; int from, to; };     // op: 0 none, 1 set, 2 drop, 3 reverse

// Best addition, then removal, then reversal; a later candidate replaces the
// incumbent only if it is better by more than sqrt(eps) (first wins on ties).
// `cache` is the column-major matrix score_cache_fill writes, `cache_rm` its
// row-major mirror (cache_rm[i n + j] = delta of the arc i -> j).
Move opt_step(const Graph &g, const std::vector<double> &cache, const std::vector<double> &cache_rm,
              std::vector<double> &reference, const int *wl, const int *bl,
              const std::vector<int> &nparents, double maxp) {
  const int n = g.n;
  const double tol = bnl::kSqrtEps;
  auto C = [&](int i, int j) { return cache_rm[(size_t)i * n + j]; };
  auto CT = [&](int i, int j) { return cache[(size_t)i + (size_t)j * n]; };      // same value, other layout
  auto GT = [&](int i, int j) { return g.a[(size_t)i + (size_t)j * n]; };
  auto WL = [&](int i, int j) { return wl ? wl[(size_t)i + (size_t)j * n] : 0; };
  auto BL = [&](int i, int j) { return bl ? bl[(size_t)i + (size_t)j * n] : 0; };
  std::vector<int> stack;
  std::vector<char> seen(n);
  double best = 0.0;
  Move mv{0, -1, -1};

  // additions: arcs absent in both directions, not blacklisted, target below maxp.
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j || g.at(i, j) || GT(j, i) || BL(i, j) == 1 || (double)nparents[j] >= maxp) continue;
      const double v = C(i, j);
      if (v - best > tol && !has_path(g, j, i, false, stack, seen)) { best = v; mv = Move{1, i, j}; }
    }
  // removals: present arcs that are not whitelisted.
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (!g.at(i, j) || WL(i, j) == 1) continue;
      const double v = C(i, j);
      if (v - best > tol) { best = v; mv = Move{2, i, j}; }
    }
  // reversals: the reverse must not be blacklisted and the new child must have room.
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (!g.at(i, j) || BL(j, i) == 1 || (double)nparents[i] >= maxp) continue;
      double v = C(i, j) + CT(j, i);
      if (fabs(v) < tol) v = 0.0;
      if (v - best > tol && !has_path(g, i, j, true, stack, seen)) { best = v; mv = Move{3, i, j}; }
    }
  if (mv.op) {
    reference[mv.to] += CT(mv.from, mv.to);
    if (mv.op == 3) reference[mv.from] += CT(mv.to, mv.from);
  }
  return mv;
}

bool discrete_score(int s) { return s < 100; }

int node_scores(bnl_ctx *ctx, bnl_data data, int score_e, double arg, const Graph &g,
                const std::vector<int> &nodes, double *out) {
  std::vector<int> off(1, 0), idx;
  for (int t : nodes) {
    for (int i = 0; i < g.n; i++) if (g.a[(size_t)i + (size_t)t * g.n]) idx.push_back(i);
    off.push_back((int)idx.size());
  }
  const int *pi = idx.empty() ? nullptr : idx.data();
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
  Graph g{n, std::vector<int>((size_t)n * n, 0), std::vector<int>((size_t)n * n, 0)};
  if (start)
    for (int j = 0; j < n; j++)
      for (int i = 0; i < n; i++) g.set(i, j, start[(size_t)i + (size_t)j * n] != 0);
  // score equivalence decides whether score_cache_fill may copy the mirrored
  // delta (is.score.equivalent, R/sanitization-scores.R:74-97): loglik/aic/bic,
  // bde with the uniform prior and their Gaussian forms are; nal/pnal only on
  // complete data; bds, bdj, k2 are not.
  bool equivalence = score_e == BNL_LOGLIK || score_e == BNL_AIC || score_e == BNL_BIC ||
                     score_e == BNL_BDE || score_e == BNL_LOGLIK_G || score_e == BNL_AIC_G ||
                     score_e == BNL_BIC_G;
  if (score_e == BNL_NAL || score_e == BNL_PNAL) {
    auto it = ctx->data.find(data);
    if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
    equivalence = true;
    for (char f : it->second.slot_has_na) if (f) equivalence = false;
  }
  std::vector<double> cache((size_t)n * n, 0.0), cache_rm((size_t)n * n, 0.0), reference(n);
  std::vector<int> updated(n), nparents(n);
  for (int i = 0; i < n; i++) updated[i] = i;
  int rc = node_scores(ctx, data, score_e, k_or_iss, g, updated, reference.data());
  if (rc) return rc;

  long long it = 1;
  for (;;) {
    for (int j = 0; j < n; j++) {
      int s = 0;
      const int *col = g.a.data() + (size_t)j * n;
      for (int i = 0; i < n; i++) s += col[i];
      nparents[j] = s;
    }
    rc = bnl_score_cache_fill(ctx, data, score_e, k_or_iss, n, g.a.data(), blmat, updated.data(),
                              (int)updated.size(), equivalence ? 1 : 0, reference.data(),
                              cache.data(), test_counter);
    if (rc) return rc;
    for (int j : updated)                      // only these columns of the cache changed
      for (int i = 0; i < n; i++) cache_rm[(size_t)i * n + j] = cache[(size_t)i + (size_t)j * n];
    const Move mv = opt_step(g, cache, cache_rm, reference, wlmat, blmat, nparents, maxp);
    if (!mv.op) break;
    updated.clear();
    if (mv.op == 1) { g.set(mv.from, mv.to, 1); updated.push_back(mv.to); }
    else if (mv.op == 2) { g.set(mv.from, mv.to, 0); updated.push_back(mv.to); }
    else {
      g.set(mv.from, mv.to, 0); g.set(mv.to, mv.from, 1);
      updated.push_back(mv.from < mv.to ? mv.from : mv.to);
      updated.push_back(mv.from < mv.to ? mv.to : mv.from);
    }
    // a node whose reference score was -Inf (singular model) and that just got a
    // finite delta is rescored from scratch (R/hill-climbing.R:229-256).
    const int ends[2][2] = {{mv.from, mv.to}, {mv.to, mv.from}};
    for (int e = 0; e < (mv.op == 3 ? 2 : 1); e++) {
      const int a = ends[e][0], b = ends[e][1];
      if (reference[b] == -INFINITY && isfinite(cache[(size_t)a + (size_t)b * n])) {
        std::vector<int> one(1, b);
        if ((rc = node_scores(ctx, data, score_e, k_or_iss, g, one, &reference[b]))) return rc;
      }
    }
    if (max_iter > 0 && it >= max_iter) break;
    it++;
  }
  memcpy(amat, g.a.data(), sizeof(int) * (size_t)n * n);
  if (node_scores_out) memcpy(node_scores_out, reference.data(), sizeof(double) * n);
  if (iterations) *iterations = it;
  return BNL_OK;
}
