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

namespace {

struct Move { int op; int from, to; };     // op: 0 none, 1 set, 2 drop, 3 reverse

struct Search {
  int n;
  double maxp;
  const int *bl, *wl;                      // column-major 0/1 or null
  std::vector<int> a;                      // adjacency, column-major (the .Call layout): a[i + j n]
  std::vector<int> r;                      // the same, row-major: r[i n + j]
  std::vector<std::vector<int>> children;  // sorted, per node
  std::vector<int> nparents;
  std::vector<double> cache;               // column-major, written by bnl_score_cache_fill
  std::vector<double> cache_rm;            // row-major mirror: cache_rm[i n + j] = delta of i -> j
  std::vector<double> rowbound;            // max of cache_rm[i][.] over the valid additions of row i
  std::vector<int> rowarg;                 // its column, -1 if the row has no valid addition
  std::vector<char> dirty;                 // rowbound[i] must be recomputed
  std::vector<double> reference;
  std::vector<int> stack;
  std::vector<char> seen;
  // tabu search: the last `tabu` networks, as a Zobrist hash (xor of one random
  // 64-bit key per arc) plus the full matrix for the rare hash match.
  std::vector<uint64_t> zkey;              // n x n keys, empty for hill climbing
  uint64_t hash = 0;                       // of the current network
  std::vector<uint64_t> tabu_hash;
  std::vector<std::vector<int>> tabu_net;
  std::vector<char> tabu_used;

  Search(int n_, double maxp_, const int *bl_, const int *wl_)
      : n(n_), maxp(maxp_), bl(bl_), wl(wl_), a((size_t)n_ * n_, 0), r((size_t)n_ * n_, 0),
        children(n_), nparents(n_, 0), cache((size_t)n_ * n_, 0.0), cache_rm((size_t)n_ * n_, 0.0),
        rowbound(n_, 0.0), rowarg(n_, -1), dirty(n_, 1), reference(n_, 0.0), seen(n_, 0) {}

  int BL(int i, int j) const { return bl ? bl[(size_t)i + (size_t)j * n] : 0; }
  int WL(int i, int j) const { return wl ? wl[(size_t)i + (size_t)j * n] : 0; }
  double C(int i, int j) const { return cache_rm[(size_t)i * n + j]; }

  void set_arc(int i, int j, int v) {
    if (a[(size_t)i + (size_t)j * n] == v) return;
    a[(size_t)i + (size_t)j * n] = v;
    r[(size_t)i * n + j] = v;
    nparents[j] += v ? 1 : -1;
    std::vector<int> &ch = children[i];
    if (v) ch.insert(std::lower_bound(ch.begin(), ch.end(), j), j);
    else ch.erase(std::lower_bound(ch.begin(), ch.end(), j));
    dirty[i] = dirty[j] = 1;               // (i, j) and (j, i) changed validity as additions
    closure_dirty = true;
    if (!v) arc_removed = true;
    else added.push_back(std::make_pair(i, j));
    if (!zkey.empty()) hash ^= zkey[(size_t)i * n + j];
  }

  void enable_tabu(int len) {
    zkey.resize((size_t)n * n);
    uint64_t x = 0x9e3779b97f4a7c15ull;    // splitmix64
    for (auto &k : zkey) {
      x += 0x9e3779b97f4a7c15ull;
      uint64_t z = x;
      z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
      z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
      k = z ^ (z >> 31);
    }
    hash = 0;
    for (int i = 0; i < n; i++) for (int j : children[i]) hash ^= zkey[(size_t)i * n + j];
    tabu_hash.assign(len, 0);
    tabu_net.assign(len, std::vector<int>());
    tabu_used.assign(len, 0);
  }
  void remember(int slot) { tabu_hash[slot] = hash; tabu_net[slot] = a; tabu_used[slot] = 1; }

  // would the network after `op` on i -> j be one of the remembered ones?
  bool is_tabu(int op, int i, int j) {
    if (zkey.empty()) return false;
    uint64_t h = hash ^ zkey[(size_t)i * n + j];
    if (op == 3) h ^= zkey[(size_t)j * n + i];
    for (size_t t = 0; t < tabu_hash.size(); t++) {
      if (!tabu_used[t] || tabu_hash[t] != h) continue;
      const std::vector<int> &m = tabu_net[t];
      // same hash: compare for real (apply the move, compare, undo).
      const size_t ij = (size_t)i + (size_t)j * n, ji = (size_t)j + (size_t)i * n;
      const int old_ij = a[ij], old_ji = a[ji];
      if (op == 1) a[ij] = 1; else if (op == 2) a[ij] = 0; else { a[ij] = 0; a[ji] = 1; }
      const bool same = (m == a);
      a[ij] = old_ij; a[ji] = old_ji;
      if (same) return true;
    }
    return false;
  }

  // may i -> j be proposed as an addition? (hc_to_be_added)
  bool valid_add(int i, int j) const {
    return i != j && !r[(size_t)i * n + j] && !a[(size_t)j + (size_t)i * n] && BL(i, j) != 1 &&
           (double)nparents[j] < maxp;
  }

  void recompute_row(int i) {
    double b = -INFINITY;
    int arg = -1;
    const double *row = cache_rm.data() + (size_t)i * n;
    // the value first: a running maximum is beaten O(log n) times along a row, and only
    // those entries need the validity and reachability lookups
    for (int j = 0; j < n; j++)
      if ((arg < 0 || row[j] > b) && valid_add(i, j) && acyclic_add(i, j)) { b = row[j]; arg = j; }
    rowbound[i] = b; rowarg[i] = arg; dirty[i] = 0;
  }

  // columns `cols` of the cache were rewritten (and their validity may have
  // changed with nparents): refresh the row-major mirror and the row bounds.
  void columns_changed(const std::vector<int> &cols) {
    if (closure_dirty) rebuild_closure();
    for (int j : cols)
      for (int i = 0; i < n; i++) {
        const double v = cache[(size_t)i + (size_t)j * n];
        cache_rm[(size_t)i * n + j] = v;
        if (dirty[i]) continue;
        if (rowarg[i] == j) dirty[i] = 1;
        else if ((rowarg[i] < 0 || v > rowbound[i]) && valid_add(i, j) && acyclic_add(i, j)) { rowbound[i] = v; rowarg[i] = j; }
      }
  }

  // Reachability of the current DAG as bitsets (reach[u] has v: a directed path u ~> v of at least
  // one arc), rebuilt after the arc set changed -- children before parents in a topological order,
  // m x n / 64 word operations, ~10 us for 1000 nodes -- so that the acyclicity checks of a scan
  // (c_has_path, one depth-first search per improving candidate in the reference) are bit tests.
  // A start network with a cycle has no topological order: then the searches below are used.
  std::vector<uint64_t> reach;
  int words = 0;
  bool closure_dirty = true, closure_ok = false, arc_removed = true;
  std::vector<std::pair<int, int>> added;          // arcs added since the closure was last current
  void rebuild_closure() {
    closure_dirty = false;
    if (closure_ok && !arc_removed && !reach.empty()) {
      // arcs were only added: a -> t lets a and everything that reached a reach t and whatever t reaches
      for (auto &e : added) {
        const int a0 = e.first, t0 = e.second;
        const uint64_t *rt = reach.data() + (size_t)t0 * words;
        for (int u = 0; u < n; u++) {
          if (u != a0 && !reaches(u, a0)) continue;
          uint64_t *ru = reach.data() + (size_t)u * words;
          for (int k = 0; k < words; k++) ru[k] |= rt[k];
          ru[t0 >> 6] |= 1ull << (t0 & 63);
        }
      }
      added.clear();
      for (int i = 0; i < n; i++)
        if (!dirty[i] && rowarg[i] >= 0 && reaches(rowarg[i], i)) dirty[i] = 1;
      return;
    }
    added.clear();
    words = (n + 63) / 64;
    reach.assign((size_t)n * words, 0);
    std::vector<int> indeg(nparents), order;
    order.reserve(n);
    for (int v = 0; v < n; v++) if (indeg[v] == 0) order.push_back(v);
    for (size_t h = 0; h < order.size(); h++)
      for (int w : children[order[h]]) if (--indeg[w] == 0) order.push_back(w);
    const bool was_ok = closure_ok;
    closure_ok = (int)order.size() == n;
    if (closure_ok)
      for (int h = n - 1; h >= 0; h--) {
        const int u = order[h];
        uint64_t *ru = reach.data() + (size_t)u * words;
        for (int c : children[u]) {
          const uint64_t *rc = reach.data() + (size_t)c * words;
          for (int k = 0; k < words; k++) ru[k] |= rc[k];
          ru[c >> 6] |= 1ull << (c & 63);
        }
      }
    // The row bounds cover the additions that are valid AND close no cycle. Arcs that were
    // only added can only have made more additions cyclic: the rows whose best one is among
    // them are recomputed; after a removal (or when the closure comes and goes) all of them.
    if (arc_removed || !closure_ok || !was_ok) std::fill(dirty.begin(), dirty.end(), 1);
    else
      for (int i = 0; i < n; i++)
        if (!dirty[i] && rowarg[i] >= 0 && reaches(rowarg[i], i)) dirty[i] = 1;
    arc_removed = false;
  }
  // i -> j closes no directed cycle (always true when the closure is not available: the
  // scan's own has_path() decides then)
  bool acyclic_add(int i, int j) const { return !closure_ok || !reaches(j, i); }
  bool reaches(int u, int v) const { return (reach[(size_t)u * words + (v >> 6)] >> (v & 63)) & 1; }

  // directed path start ~> stop; `skip_direct` ignores the arc start -> stop itself.
  bool has_path(int start, int stop, bool skip_direct) {
    if (closure_dirty) rebuild_closure();
    if (closure_ok) {
      if (!skip_direct) return reaches(start, stop);
      for (int c : children[start])                    // any other way round: through another child
        if (c != stop && reaches(c, stop)) return true;
      return false;
    }
    std::fill(seen.begin(), seen.end(), 0);
    stack.clear();
    stack.push_back(start);
    seen[start] = 1;
    while (!stack.empty()) {
      const int v = stack.back();
      stack.pop_back();
      for (int w : children[v]) {
        if (skip_direct && v == start && w == stop) continue;
        if (w == stop) return true;
        if (!seen[w]) { seen[w] = 1; stack.push_back(w); }
      }
    }
    return false;
  }

  // Best addition, then removal, then reversal, in the reference's scan order.
  // Hill climbing: baseline 0, tiny reversal deltas zeroed. Tabu search
  // (tabu_step, src/learning/score/tabu.c:93-435): moves into a remembered network
  // are skipped, the baseline is 0 or -Inf (accept the least bad move) and
  // reversal deltas are taken as they are.
  Move opt_step(double baseline = 0.0) {
    const double tol = bnl::kSqrtEps;
    const bool tabu_mode = !zkey.empty();
    double best = baseline;
    Move mv{0, -1, -1};
    if (closure_dirty) rebuild_closure();
    { BNL_TRACE("hc: recompute dirty rows"); for (int i = 0; i < n; i++) if (dirty[i]) recompute_row(i); }
    // The scan accepts a candidate when it beats the running best by more than tol, so its
    // result is the LAST candidate accepted and depends on the order. Rows far below the top
    // can still be left out exactly. Let L be the value of a candidate known to be acceptable
    // (the largest row bound, if that addition is acyclic and not tabu) and theta = L - 2 tol.
    // While the running best is <= theta, a candidate above theta + tol is accepted whatever the
    // best is, and one <= theta can only move the best around below theta; once a candidate
    // above theta has been accepted the low ones are rejected anyway. So skipping every value
    // <= theta gives the same last accepted candidate -- provided no value lies in the band
    // (theta, theta + tol], where acceptance would depend on the exact best: seeing one (it
    // takes two deltas 1.5e-8 apart just below the top) the scan starts over without the cut.
    // On 1000 nodes this takes the rows scanned per iteration from ~40 to 1-2.
    double theta = -INFINITY;
    {
      BNL_TRACE("hc: top candidate");
      // the rows with the largest bounds, best first, until one's own best addition is acceptable
      const int kTries = 16;
      int tops[kTries], nt = 0;
      for (int i = 0; i < n; i++) {
        if (rowarg[i] < 0 || !(rowbound[i] - 3.0 * tol > baseline) || !std::isfinite(rowbound[i])) continue;
        int p;
        if (nt < kTries) p = nt++;
        else if (rowbound[i] > rowbound[tops[kTries - 1]]) p = kTries - 1;
        else continue;
        while (p > 0 && rowbound[i] > rowbound[tops[p - 1]]) { tops[p] = tops[p - 1]; p--; }
        tops[p] = i;
      }
      for (int t = 0; t < nt; t++) {
        const int i = tops[t], j = rowarg[i];
        if (!has_path(j, i, false) && !is_tabu(1, i, j)) { theta = rowbound[i] - 2.0 * tol; break; }
      }
    }
    for (int attempt = 0; attempt < 2; attempt++) {
      const bool cut = attempt == 0 && theta > -INFINITY;
      BnlTraceScope tr_scan(cut ? "hc: additions, cut scan" : (attempt ? "hc: additions, full scan after band" : "hc: additions, full scan"));
      bool band = false;
      best = baseline;
      mv = Move{0, -1, -1};
      for (int i = 0; i < n && !band; i++) {
        if (rowarg[i] < 0 || !(rowbound[i] - best > tol)) continue;      // nothing in this row can win
        const bool low = cut && !(best > theta);                          // no candidate above theta accepted yet
        if (low && !(rowbound[i] > theta)) continue;
        const double *row = cache_rm.data() + (size_t)i * n;
        for (int j = 0; j < n; j++) {
          const double v = row[j];
          if (cut && !(best > theta)) {
            if (!(v > theta)) continue;
            if (!(v > theta + tol)) { band = true; break; }
          }
          if (!(v - best > tol) || !valid_add(i, j)) continue;
          if (!has_path(j, i, false) && !is_tabu(1, i, j)) { best = v; mv = Move{1, i, j}; }
        }
      }
      if (!band) break;
    }
    // removals: present arcs that are not whitelisted.
    BNL_TRACE("hc: removals and reversals");
    for (int i = 0; i < n; i++)
      for (int j : children[i]) {
        if (WL(i, j) == 1) continue;
        const double v = C(i, j);
        if (v - best > tol && !is_tabu(2, i, j)) { best = v; mv = Move{2, i, j}; }
      }
    // reversals: the reverse must not be blacklisted and the new child must have room.
    for (int i = 0; i < n; i++) {
      if ((double)nparents[i] >= maxp) continue;
      for (int j : children[i]) {
        if (BL(j, i) == 1) continue;
        double v = C(i, j) + C(j, i);
        if (!tabu_mode && fabs(v) < tol) v = 0.0;
        if (v - best > tol && !has_path(i, j, true) && !is_tabu(3, i, j)) { best = v; mv = Move{3, i, j}; }
      }
    }
    if (mv.op) {
      reference[mv.to] += C(mv.from, mv.to);
      if (mv.op == 3) reference[mv.from] += C(mv.to, mv.from);
    }
    return mv;
  }
};

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
