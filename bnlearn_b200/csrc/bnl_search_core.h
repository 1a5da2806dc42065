// bnl_search_core.h -- the bookkeeping of hill climbing / tabu search between two
// score_cache_fill calls: the arc set, the score cache with its row bounds, and
// hc_opt_step (src/learning/score/hc.cache.lookup.c:131-363; tabu_step,
// src/learning/score/tabu.c:93-435). No CUDA in here: bnl_search.cu drives it, and
// tests/search_core_test.cpp checks opt_step() against a plain scan in the reference's
// order on the CPU (random caches with engineered near-ties, random move sequences).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

#ifndef BNL_TRACE
#define BNL_TRACE(label) do { } while (0)
#define BNL_SEARCH_CORE_NO_TRACE 1
#endif

namespace bnlsearch {

const double kTol = 1.4901161193847656e-08;   // sqrt(DBL_EPSILON), MACHINE_TOL in the reference

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
  long long band_restarts = 0;                     // scans redone without the cut (see opt_step)
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
    const double tol = kTol;
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
#ifndef BNL_SEARCH_CORE_NO_TRACE
      BnlTraceScope tr_scan(cut ? "hc: additions, cut scan" : (attempt ? "hc: additions, full scan after band" : "hc: additions, full scan"));
#endif
      if (attempt) band_restarts++;
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

}  // namespace bnlsearch
