// search_core_test.cpp -- Search::opt_step() (bnlearn_b200/csrc/bnl_search_core.h: row bounds over
// valid and acyclic additions, reachability closure kept up to date, the exact cut two tolerances
// below the best acceptable addition) against a plain scan in the reference's order
// (hc_opt_step, src/learning/score/hc.cache.lookup.c:131-363: additions i outer / j inner, then
// removals, then reversals; a candidate replaces the incumbent only when it is better by more
// than sqrt(eps)). Random DAG walks; caches with exact ties and with values spaced a fraction of
// the tolerance apart just below the top, so that the cut's way out (the band) is taken too.
//   g++ -O2 -std=c++17 -I bnlearn_b200/csrc -o search_core_test tests/search_core_test.cpp && ./search_core_test
#include <stdio.h>
#include <stdlib.h>

#include <random>

#include "bnl_search_core.h"

using bnlsearch::Move;
using bnlsearch::Search;
static const double tol = bnlsearch::kTol;

struct Naive {
  int n;
  double maxp;
  const int *bl, *wl;
  std::vector<int> a;        // a[i + j n] = 1: arc i -> j
  int A(int i, int j) const { return a[(size_t)i + (size_t)j * n]; }
  int BL(int i, int j) const { return bl ? bl[(size_t)i + (size_t)j * n] : 0; }
  int WL(int i, int j) const { return wl ? wl[(size_t)i + (size_t)j * n] : 0; }
  int nparents(int j) const { int c = 0; for (int i = 0; i < n; i++) c += A(i, j); return c; }
  bool path(int start, int stop, bool skip_direct) const {
    std::vector<int> st(1, start);
    std::vector<char> seen(n, 0);
    seen[start] = 1;
    while (!st.empty()) {
      const int v = st.back(); st.pop_back();
      for (int w = 0; w < n; w++) {
        if (!A(v, w)) continue;
        if (skip_direct && v == start && w == stop) continue;
        if (w == stop) return true;
        if (!seen[w]) { seen[w] = 1; st.push_back(w); }
      }
    }
    return false;
  }
  bool tabu_mode = false;
  std::vector<std::vector<int>> remembered;      // tabu search: the last networks
  bool is_tabu(int op, int i, int j) const {     // would the move lead into a remembered network?
    if (!tabu_mode) return false;
    std::vector<int> b = a;
    if (op == 1) b[(size_t)i + (size_t)j * n] = 1;
    else if (op == 2) b[(size_t)i + (size_t)j * n] = 0;
    else { b[(size_t)i + (size_t)j * n] = 0; b[(size_t)j + (size_t)i * n] = 1; }
    for (auto &m : remembered) if (!m.empty() && m == b) return true;
    return false;
  }
  Move step(const std::vector<double> &cache, double baseline) const {
    auto C = [&](int i, int j) { return cache[(size_t)i + (size_t)j * n]; };
    double best = baseline;
    Move mv{0, -1, -1};
    std::vector<int> np(n);
    for (int j = 0; j < n; j++) np[j] = nparents(j);
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++) {
        if (i == j || A(i, j) || A(j, i) || BL(i, j) == 1 || !((double)np[j] < maxp)) continue;
        if (!(C(i, j) - best > tol)) continue;
        if (path(j, i, false) || is_tabu(1, i, j)) continue;
        best = C(i, j); mv = Move{1, i, j};
      }
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++) {
        if (!A(i, j) || WL(i, j) == 1) continue;
        if (C(i, j) - best > tol && !is_tabu(2, i, j)) { best = C(i, j); mv = Move{2, i, j}; }
      }
    for (int i = 0; i < n; i++) {
      if ((double)np[i] >= maxp) continue;
      for (int j = 0; j < n; j++) {
        if (!A(i, j) || BL(j, i) == 1) continue;
        double v = C(i, j) + C(j, i);
        if (!tabu_mode && fabs(v) < tol) v = 0.0;
        if (v - best > tol && !path(i, j, true) && !is_tabu(3, i, j)) { best = v; mv = Move{3, i, j}; }
      }
    }
    return mv;
  }
};

int main() {
  std::mt19937_64 rng(20261107);
  std::normal_distribution<double> normal(0.0, 1.0);
  auto unif = [&](int k) { return (int)(rng() % (unsigned long long)k); };
  long long steps = 0, restarts = 0, cuts_possible = 0;
  const int sizes[] = {6, 17, 40, 70, 131};
  for (int trial = 0; trial < 80; trial++) {
    const int n = sizes[trial % 5];
    const double maxp = (trial % 3 == 0) ? 3.0 : INFINITY;
    std::vector<int> bl((size_t)n * n, 0), wl((size_t)n * n, 0);
    for (int k = 0; k < n; k++) bl[(size_t)unif(n) + (size_t)unif(n) * n] = 1;
    const bool use_lists = trial % 2 == 0;
    Search S(n, maxp, use_lists ? bl.data() : nullptr, use_lists ? wl.data() : nullptr);
    Naive N{n, maxp, use_lists ? bl.data() : nullptr, use_lists ? wl.data() : nullptr, std::vector<int>((size_t)n * n, 0)};
    // every fifth trial is a tabu search (tabu_step, src/learning/score/tabu.c:93-435): the last 7 networks are
    // remembered, moves into one of them are skipped, reversal deltas are taken as they are
    const int tabu_len = trial % 5 == 4 ? 7 : 0;
    if (tabu_len) { S.enable_tabu(tabu_len); N.tabu_mode = true; N.remembered.assign(tabu_len, std::vector<int>()); }
    // the kind of cache: 0 plain random, 1 many exact ties (few distinct values), 2 a ladder of values
    // 0.4 tolerances apart just below the top (the band above the cut line gets hit), 3 both
    auto fill_column = [&](int j, int kind) {
      for (int i = 0; i < n; i++) {
        double v = normal(rng) * 5.0;
        if (kind == 1 || kind == 3) v = (double)(unif(7) - 2) * 1.5;
        S.cache[(size_t)i + (size_t)j * n] = v;
      }
    };
    auto ladder = [&]() {
      double top = -INFINITY;
      for (double v : S.cache) top = std::max(top, v);
      const int rungs = 3 + unif(8);
      for (int k = 0; k < rungs; k++)
        S.cache[(size_t)unif(n) + (size_t)unif(n) * n] = top + 1.0 - (double)k * 0.4 * tol * (1 + unif(3));
    };
    std::vector<int> all(n);
    for (int j = 0; j < n; j++) all[j] = j;
    int kind = trial % 4;
    for (int j = 0; j < n; j++) fill_column(j, kind);
    if (kind >= 2) ladder();
    S.columns_changed(all);
    for (int it = 0; it < 400; it++) {
      const double baseline = (it % 11 == 10) ? -INFINITY : 0.0;
      if (tabu_len) { S.remember(it % tabu_len); N.remembered[it % tabu_len] = N.a; }
      const Move want = N.step(S.cache, baseline);
      const long long before = S.band_restarts;
      const Move got = S.opt_step(baseline);
      restarts += S.band_restarts - before;
      steps++;
      if (got.op != want.op || got.from != want.from || got.to != want.to) {
        printf("MISMATCH trial %d step %d n %d: got (%d, %d -> %d), want (%d, %d -> %d)\n", trial, it, n, got.op, got.from,
               got.to, want.op, want.from, want.to);
        return 1;
      }
      std::vector<int> updated;
      if (!got.op) {                      // converged: a new cache over the same graph
        kind = unif(4);
        for (int j = 0; j < n; j++) fill_column(j, kind);
        if (kind >= 2) ladder();
        S.columns_changed(all);
        continue;
      }
      cuts_possible++;
      auto set = [&](int i, int j, int v) { S.set_arc(i, j, v); N.a[(size_t)i + (size_t)j * n] = v; };
      if (got.op == 1) { set(got.from, got.to, 1); updated.push_back(got.to); }
      else if (got.op == 2) { set(got.from, got.to, 0); updated.push_back(got.to); }
      else { set(got.from, got.to, 0); set(got.to, got.from, 1); updated.push_back(std::min(got.from, got.to)); updated.push_back(std::max(got.from, got.to)); }
      // as after score_cache_fill: the columns of the nodes whose parents changed are rewritten
      for (int j : updated) fill_column(j, kind);
      if (kind >= 2 && unif(3) == 0) {    // a ladder confined to the rewritten columns
        double top = -INFINITY;
        for (double v : S.cache) top = std::max(top, v);
        for (int k = 0; k < 4; k++)
          S.cache[(size_t)unif(n) + (size_t)updated[unif((int)updated.size())] * n] = top + 0.5 - (double)k * 0.45 * tol;
      }
      S.columns_changed(updated);
    }
  }
  printf("ok: %lld steps agree with the plain scan (%lld moves, %lld scans redone without the cut)\n", steps, cuts_possible,
         restarts);
  if (restarts == 0) { printf("the band above the cut line was never hit: the test does not cover the way out\n"); return 2; }
  return 0;
}
