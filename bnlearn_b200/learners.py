"""Restatements of the reference's UNCHANGED search loops, so that learned
structures and test-counter totals can be checked end to end without an R
interpreter. Everything that touches the data
goes through a `backend` -- callsites.GpuBackend here, a wrapper over the
reference's own .Call entry points (oracle/_ref) in the tests -- so the two
runs differ ONLY in who evaluates the statistics.

Nodes are 0-based column indices. Citations are to the reference's R/ and src/.
What is restated: hc (R/hill-climbing.R:3-293 with hc_opt_step and
hc_to_be_added, src/learning/score/hc.cache.lookup.c:131-489), tabu
(R/tabu.R:3-314, src/learning/score/tabu.c:93-435), pc.stable's skeleton phase
(R/pcalgo.R:2-148), mmpc / si.hiton.pc / iamb local discovery
(R/maxmin-pc.R, R/hiton-pc.R, R/incremental-association.R, R/backend-indep.R:55-133,
src/learning/rinterface/bn.recovery.c) and the v-structure tests
(R/backend-indep.R:135-215). Not restated: random restarts, white/blacklists in
the constraint-based learners, cpdag.backend.
"""
import math

import numpy as np

from . import callsites as cs

TOL = math.sqrt(np.finfo(np.float64).eps)
SCORE_EQUIVALENT = ("loglik", "aic", "bic", "ebic", "bde", "qnml", "loglik-g", "aic-g", "bic-g", "ebic-g", "bge")


class GpuBackend:
    """The statistics side of the learners, on the CUDA engine."""

    def __init__(self, ds, stream=None):
        self.ds = ds
        self.counter = cs.TestCounter()
        self.stream = stream                  # cs.UniformStream for the permutation tests

    def indep_test(self, xs, y, sx, test, alpha=1.0):
        return cs.indep_test(self.ds, xs, y, sx, test, self.counter, self.stream, alpha)

    def allsubs_batch(self, x, y, sx, nfixed, test, alpha, minsize, maxsize):
        if test in cs.MC_TESTS:               # one request at a time: the stream passes from one to the next
            return [self.allsubs(x[r], y[r], sx[r][nfixed[r]:], sx[r][:nfixed[r]], test, alpha, minsize, maxsize)
                    for r in range(len(x))]
        return cs.allsubs_test_batch(self.ds, x, y, sx, nfixed, test, alpha, minsize, maxsize, self.counter)

    def allsubs(self, x, y, sx, fixed, test, alpha, minsize, maxsize):
        return cs.allsubs_test(self.ds, x, y, sx, fixed, test, alpha, minsize, maxsize, self.counter,
                               self.stream)

    def roundrobin(self, x, z, fixed, test, alpha):
        return cs.roundrobin_test(self.ds, x, z, fixed, test, alpha, self.counter, self.stream)

    def per_node_score(self, targets, parents, score, arg):
        return cs.per_node_score(self.ds, targets, parents, score, arg)

    def score_cache_fill(self, score, arg, amat, blmat, updated, equivalence, reference, cache):
        return cs.score_cache_fill(self.ds, score, arg, amat, blmat, updated, equivalence, reference,
                                   cache, self.counter)


# ---------------------------------------------------------------------------
# score-based search

def has_path(start, stop, amat, notdirect=False):
    """c_has_path(): directed path start ~> stop (src/graphs/path.c:29)."""
    n = amat.shape[0]
    seen = np.zeros(n, dtype=bool)
    stack = [start]
    seen[start] = True
    while stack:
        v = stack.pop()
        for w in np.flatnonzero(amat[v]):
            if notdirect and v == start and w == stop:
                continue
            if w == stop:
                return True
            if not seen[w]:
                seen[w] = True
                stack.append(w)
    return False


def arcs_to_be_added(amat, blmat, nparents, maxp):
    """hc_to_be_added() (src/learning/score/hc.cache.lookup.c:373-489)."""
    a = 1 - amat
    np.fill_diagonal(a, 0)
    a = a * a.T
    a[:, nparents >= maxp] = 0
    a[blmat == 1] = 0
    return a


def _scan(cands_i, cands_j, vals, best, accept):
    """The reference's first-wins scan: a candidate replaces the incumbent iff
    value - max > sqrt(eps) (and `accept` agrees); returns (index or -1, max)."""
    pos, chosen = 0, -1
    m = len(vals)
    while pos < m:
        rel = np.flatnonzero(vals[pos:] - best > TOL)
        if rel.size == 0:
            break
        k = pos + int(rel[0])
        if accept(int(cands_i[k]), int(cands_j[k])):
            chosen, best = k, vals[k]
        pos = k + 1
    return chosen, best


def hc_opt_step(amat, added, cache, reference, wlmat, blmat, nparents, maxp, baseline=0.0,
                tabu=None):
    """hc_opt_step() / tabu_step(): best add, then drop, then reverse
    (src/learning/score/hc.cache.lookup.c:131-363, tabu.c:93-435). `reference`
    is updated in place like the C code does. Returns (op, from, to) or None."""
    best = baseline
    found = None

    def not_tabu(trial):
        return tabu is None or not tabu(trial)

    ai, aj = np.nonzero(added)                       # row-major: i outer, j inner
    def ok_add(i, j):
        if has_path(j, i, amat):
            return False
        if tabu is None:
            return True
        amat[i, j] = 1
        r = not_tabu(amat)
        amat[i, j] = 0
        return r
    k, best = _scan(ai, aj, cache[ai, aj], best, ok_add)
    if k >= 0:
        found = ("set", int(ai[k]), int(aj[k]))

    di, dj = np.nonzero((amat == 1) & (wlmat != 1))
    def ok_drop(i, j):
        if tabu is None:
            return True
        amat[i, j] = 0
        r = not_tabu(amat)
        amat[i, j] = 1
        return r
    k, best = _scan(di, dj, cache[di, dj], best, ok_drop)
    if k >= 0:
        found = ("drop", int(di[k]), int(dj[k]))

    ri, rj = np.nonzero((amat == 1) & (blmat.T != 1) & (nparents[:, None] < maxp))
    vals = cache[ri, rj] + cache[rj, ri]
    if tabu is None:
        vals = np.where(np.abs(vals) < TOL, 0.0, vals)   # hc zeroes tiny reversal deltas, tabu does not
    def ok_rev(i, j):
        if has_path(i, j, amat, notdirect=True):
            return False
        if tabu is None:
            return True
        amat[i, j], amat[j, i] = 0, 1
        r = not_tabu(amat)
        amat[i, j], amat[j, i] = 1, 0
        return r
    k, best = _scan(ri, rj, vals, best, ok_rev)
    if k >= 0:
        found = ("reverse", int(ri[k]), int(rj[k]))

    if found is not None:
        op, f, t = found
        reference[t] += cache[f, t]
        if op == "reverse":
            reference[f] += cache[t, f]
    return found


def _apply(amat, op):
    kind, f, t = op
    if kind == "set":
        amat[f, t] = 1
    elif kind == "drop":
        amat[f, t] = 0
    else:
        amat[f, t], amat[t, f] = 0, 1


def _parents(amat):
    return [[int(i) for i in np.flatnonzero(amat[:, j])] for j in range(amat.shape[0])]


def _fix_singular(backend, amat, op, cache, reference, score, arg):
    """reset -Inf reference scores after a move (R/hill-climbing.R:229-256)."""
    kind, f, t = op
    pairs = [(f, t)] + ([(t, f)] if kind == "reverse" else [])
    for a, b in pairs:
        if reference[b] == -np.inf and np.isfinite(cache[a, b]):
            reference[b] = backend.per_node_score([b], [_parents(amat)[b]], score, arg)[0]


def hc(backend, nnodes, score, arg=0.0, maxp=np.inf, max_iter=np.inf, start=None):
    """hill.climbing(optimized = TRUE, restart = 0). Returns (amat, per-node scores)."""
    amat = (np.zeros((nnodes, nnodes), dtype=np.int32, order="F") if start is None
            else np.asfortranarray(start, dtype=np.int32).copy(order="F"))
    blmat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")
    wlmat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")
    cache = np.zeros((nnodes, nnodes), order="F")
    equivalence = score in SCORE_EQUIVALENT
    reference = np.array(backend.per_node_score(list(range(nnodes)), _parents(amat), score, arg), dtype=float)
    updated = list(range(nnodes))
    it = 1
    while True:
        nparents = amat.sum(axis=0)
        backend.score_cache_fill(score, arg, amat, blmat, updated, equivalence, reference, cache)
        added = arcs_to_be_added(amat, blmat, nparents, maxp)
        op = hc_opt_step(amat, added, cache, reference, wlmat, blmat, nparents, maxp)
        if op is None:
            break
        _apply(amat, op)
        updated = sorted({op[1], op[2]}) if op[0] == "reverse" else [op[2]]
        _fix_singular(backend, amat, op, cache, reference, score, arg)
        if it >= max_iter:
            break
        it += 1
    return amat, reference


def _robust(new, old):
    """robust.score.difference() (R/utils-misc.R:225-242)."""
    if new == -np.inf and old == -np.inf:
        return -np.inf
    if new != -np.inf and old == -np.inf:
        return abs(new)
    return 0.0 if abs(new - old) < TOL else new - old


def tabu(backend, nnodes, score, arg=0.0, maxp=np.inf, max_iter=np.inf, tabu=10, max_tabu=None):
    """tabu.search(optimized = TRUE). Returns (amat, per-node scores of the last network)."""
    max_tabu = tabu if max_tabu is None else max_tabu
    amat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")
    blmat = np.zeros_like(amat)
    wlmat = np.zeros_like(amat)
    cache = np.zeros((nnodes, nnodes), order="F")
    equivalence = score in SCORE_EQUIVALENT
    reference = np.array(backend.per_node_score(list(range(nnodes)), _parents(amat), score, arg), dtype=float)
    updated = list(range(nnodes))
    tabu_list = [None] * tabu
    best_net, best_score, loss_iter, it = amat.copy(), -np.inf, 0, 1

    def in_tabu(trial):
        key = trial.tobytes()
        return any(t is not None and t == key for t in tabu_list)

    while True:
        current = (it - 1) % tabu
        if _robust(reference.sum(), best_score) > 0 or it == 1:
            best_net, best_score = amat.copy(), reference.sum()
        nparents = amat.sum(axis=0)
        tabu_list[current] = amat.tobytes()
        backend.score_cache_fill(score, arg, amat, blmat, updated, equivalence, reference, cache)
        added = arcs_to_be_added(amat, blmat, nparents, maxp)
        op = hc_opt_step(amat, added, cache, reference, wlmat, blmat, nparents, maxp, 0.0, in_tabu)
        if op is None:
            if loss_iter >= max_tabu:
                amat = best_net
                break
            loss_iter += 1
            op = hc_opt_step(amat, added, cache, reference, wlmat, blmat, nparents, maxp, -np.inf, in_tabu)
            if op is None:
                if loss_iter > 0:
                    amat = best_net
                break
        elif _robust(reference.sum(), best_score) > 0:
            loss_iter = 0
        _apply(amat, op)
        updated = sorted({op[1], op[2]}) if op[0] == "reverse" else [op[2]]
        if it >= max_iter:
            if loss_iter > 0:
                amat = best_net
            break
        it += 1
    return amat, reference


# ---------------------------------------------------------------------------
# constraint-based search

def pc_stable(backend, nnodes, test, alpha=0.05, max_sx=None):
    """pc.stable.backend(): skeleton + separating sets (R/pcalgo.R:2-148). All
    pairs of a level see the skeleton frozen at the previous level, so they are
    evaluated in one batched allsubs call; results and counter totals are those
    of the pair-by-pair loop. Returns (sorted edge list, {pair: dsep set},
    {pair: p-value})."""
    max_sx = nnodes if max_sx is None else max_sx
    pairs = [(i, j) for i in range(nnodes) for j in range(i + 1, nnodes)]
    state = {p: {"p": None, "dsep": None, "max_adj": nnodes - 1} for p in pairs}
    skeleton = list(pairs)
    max_dsep = min(max_sx, nnodes - 2)
    for level in range(0, max_dsep + 1):
        def nbrs(a):
            # rows where `a` is the second endpoint come first (R/pcalgo.R:88-91)
            return [p[0] for p in skeleton if p[1] == a] + [p[1] for p in skeleton if p[0] == a]
        adj = {a: nbrs(a) for a in range(nnodes)}
        live = [p for p in pairs if level <= state[p]["max_adj"]]
        first, second = [], []
        info = {}
        for p in live:
            st = state[p]
            if st["dsep"] is not None:
                continue                                  # already separated: returned unchanged
            n1 = [v for v in adj[p[0]] if v != p[1]]
            n2 = [v for v in adj[p[1]] if v != p[0]]
            if len(n1) < level and len(n2) < level:
                st["max_adj"] = 0
                continue
            info[p] = (n1, n2)
            if len(n1) >= level:
                first.append(p)
        if test in cs.MC_TESTS:
            # permutation tests draw from one stream: keep the reference's order, pair by
            # pair and both directions of a pair before the next one (R/pcalgo.R:93-141)
            separated = set()
            for p in list(info):
                n1, n2 = info[p]
                if len(n1) >= level:
                    pv, _, _, dsep = backend.allsubs(p[0], p[1], n1, [], test, alpha, level, level)
                    if pv > alpha:
                        state[p].update(p=pv, dsep=dsep if dsep is not None else [], max_adj=0)
                        separated.add(p)
                        continue
                if level == 1:
                    n2 = [v for v in n2 if v not in n1]
                info[p] = (n1, n2)
                if len(n2) >= level and level > 0 and set(n1) != set(n2):
                    pv, _, _, dsep = backend.allsubs(p[1], p[0], n2, [], test, alpha, level, level)
                    if pv > alpha:
                        state[p].update(p=pv, dsep=dsep if dsep is not None else [], max_adj=0)
                        separated.add(p)
                        continue
                state[p].update(p=0.0, dsep=None, max_adj=max(len(n1), len(n2)))
            skeleton = [p for p in pairs if state[p]["p"] is not None and state[p]["p"] < alpha]
            if all(state[p]["max_adj"] <= level for p in pairs):
                break
            continue
        res = backend.allsubs_batch([p[0] for p in first], [p[1] for p in first],
                                    [info[p][0] for p in first], [0] * len(first), test, alpha, level, level)
        separated = set()
        for p, (pv, _, _, dsep) in zip(first, res):
            if pv > alpha:
                state[p].update(p=pv, dsep=dsep if dsep is not None else [], max_adj=0)
                separated.add(p)
        for p in info:
            if p in separated:
                continue
            n1, n2 = info[p]
            if level == 1:
                n2 = [v for v in n2 if v not in n1]
            info[p] = (n1, n2)
            if len(n2) >= level and level > 0 and set(n1) != set(n2):
                second.append(p)
        res = backend.allsubs_batch([p[1] for p in second], [p[0] for p in second],
                                    [info[p][1] for p in second], [0] * len(second), test, alpha, level, level)
        for p, (pv, _, _, dsep) in zip(second, res):
            if pv > alpha:
                state[p].update(p=pv, dsep=dsep if dsep is not None else [], max_adj=0)
                separated.add(p)
        for p in info:
            if p not in separated:
                n1, n2 = info[p]
                state[p].update(p=0.0, dsep=None, max_adj=max(len(n1), len(n2)))
        skeleton = [p for p in pairs if state[p]["p"] is not None and state[p]["p"] < alpha]
        if all(state[p]["max_adj"] <= level for p in pairs):
            break
    dsep = {p: state[p]["dsep"] for p in pairs if state[p]["dsep"] is not None}
    pvals = {p: state[p]["p"] for p in pairs}
    return skeleton, dsep, pvals


def bn_recovery(sets, nnodes):
    """bn_recovery(filter = "AND") on plain membership lists
    (src/learning/rinterface/bn.recovery.c): symmetric input is returned as is,
    otherwise every list is rebuilt, in DEcreasing node order."""
    member = [[False] * nnodes for _ in range(nnodes)]
    for i in range(nnodes):
        for j in sets[i]:
            member[i][j] = True
    if all(member[i][j] == member[j][i] for i in range(nnodes) for j in range(nnodes)):
        return [list(s) for s in sets]
    return [[j for j in range(nnodes - 1, -1, -1) if j != i and member[i][j] and member[j][i]]
            for i in range(nnodes)]


def neighbour(backend, target, mb, test, alpha, empty_dsep=True, markov=True, max_sx=None):
    """neighbour() (R/backend-indep.R:55-133) without white/blacklists."""
    cand = list(mb[target])
    if len(cand) <= 1:
        return cand
    max_sx = 10 ** 9 if max_sx is None else max_sx
    for y in list(cand):
        a = [v for v in mb[target] if v != y]
        if markov:
            b = [v for v in mb[y] if v != target]
            dsep = a if len(a) < len(b) else b
        else:
            dsep = a
        pv = backend.allsubs(target, y, dsep, [], test, alpha, 0 if empty_dsep else 1,
                             min(len(dsep), max_sx))[0]
        if pv > alpha:
            cand = [v for v in cand if v != y]
    return cand


def _first_min(assoc):
    """names(which.min(association)): the first minimum in insertion order."""
    best = None
    for k, v in assoc.items():
        if best is None or v < assoc[best]:
            best = k
    return best


def mmpc_forward(backend, target, nnodes, test, alpha, max_sx):
    """maxmin.pc.forward.phase() (R/maxmin-pc.R:33-79)."""
    nodes = [v for v in range(nnodes) if v != target]
    assoc = {v: 0.0 for v in nodes}
    cpc = []
    while True:
        if len(cpc) > max_sx:
            break
        check = [v for v, a in assoc.items() if a < alpha and v not in cpc]
        new = {}
        for x in check:
            fixed = cpc[-1:] if cpc else []
            sx = cpc[:-1]
            res = backend.allsubs(x, target, sx, fixed, test, alpha, 0, len(sx))
            new[x] = max(assoc[x], res[2])
        assoc = new
        if all(a > alpha for a in assoc.values()) or not nodes:
            break
        add = _first_min(assoc)
        if assoc[add] <= alpha:
            cpc.append(add)
            nodes = [v for v in nodes if v != add]
    return cpc


def hiton_forward(backend, target, nnodes, test, alpha, max_sx):
    """si.hiton.pc.heuristic() (R/hiton-pc.R:24-86)."""
    nodes = [v for v in range(nnodes) if v != target]
    pv = backend.indep_test(nodes, target, [], test, alpha)
    assoc = {v: float(p) for v, p in zip(nodes, pv)}
    nodes = [v for v in nodes if assoc[v] <= alpha]
    cpc = []
    if all(a > alpha for a in assoc.values()):
        return cpc
    while True:
        if all(a > alpha for a in assoc.values()) or not nodes or len(cpc) > max_sx:
            break
        add = _first_min(assoc)
        ok = True
        if cpc:
            ok = backend.allsubs(target, add, cpc, [], test, alpha, 1, len(cpc))[0] <= alpha
        if ok:
            cpc.append(add)
        nodes = [v for v in nodes if v != add]
        assoc = {k: v for k, v in assoc.items() if k != add}
    return cpc


def iamb_blanket(backend, target, nnodes, test, alpha, max_sx):
    """ia.markov.blanket() (R/incremental-association.R:19-73)."""
    nodes = [v for v in range(nnodes) if v != target]
    mb, last = [], None
    while nodes and len(mb) <= max_sx:
        pv = backend.indep_test(nodes, target, mb, test, alpha)
        if all(p > alpha for p in pv):
            break
        k = int(np.argmin(pv))
        last = nodes[k]
        if pv[k] <= alpha:
            mb.append(last)
            nodes = [v for v in nodes if v != last]
    fixed = [last] if last is not None else []
    if not mb:
        return []
    pv = backend.roundrobin(target, mb, fixed, test, alpha)
    keep = [v for v, p in zip([v for v in mb if v not in fixed], pv) if p < alpha] + fixed
    return [v for v in mb if v in keep]


def _fake_mb(nbr, target):
    out = []
    for c in nbr[target]:
        out += nbr[c]
    out += nbr[target]
    seen = []
    for v in out:
        if v != target and v not in seen:
            seen.append(v)
    return seen


def local_discovery(backend, nnodes, algo, test, alpha=0.05, max_sx=None):
    """mmpc / si.hiton.pc / iamb up to the symmetric neighbourhoods
    (R/maxmin-pc.R:2-31, R/hiton-pc.R:2-22, R/incremental-association.R:2-17).
    Returns (neighbours, markov blankets)."""
    max_sx = nnodes if max_sx is None else max_sx
    if algo == "iamb":
        mb = [iamb_blanket(backend, t, nnodes, test, alpha, max_sx) for t in range(nnodes)]
        mb = bn_recovery(mb, nnodes)
        nbr = [neighbour(backend, t, mb, test, alpha, True, True, max_sx) for t in range(nnodes)]
        return bn_recovery(nbr, nnodes), mb
    fwd = mmpc_forward if algo == "mmpc" else hiton_forward
    cpc = [fwd(backend, t, nnodes, test, alpha, max_sx) for t in range(nnodes)]
    nbr = [neighbour(backend, t, cpc, test, alpha, algo == "mmpc", False, max_sx) for t in range(nnodes)]
    mb = [_fake_mb(nbr, t) for t in range(nnodes)]
    return bn_recovery(nbr, nnodes), mb


def vstructures(backend, nnodes, nbr, mb, test, alpha, dsep=None, pvals=None, max_sx=None):
    """vstruct.detect() (R/backend-indep.R:135-215): unshielded colliders y -> x <- z.
    With pc.stable's separating sets no test is run; otherwise each candidate
    costs one allsubs call with x fixed. Returns [(max_a, y, x, z)]."""
    max_sx = nnodes if max_sx is None else max_sx
    out = []
    for x in range(nnodes):
        tos = [v for v in range(nnodes) if x in nbr[v]]         # arcs v -> x, in arc-set order
        for a in range(len(tos)):
            for b in range(a + 1, len(tos)):
                y, z = tos[a], tos[b]
                if z in nbr[y] or y in nbr[z]:
                    continue
                if dsep is not None:
                    key = (min(y, z), max(y, z))
                    if x not in (dsep.get(key) or []):
                        out.append((pvals[key], y, x, z))
                else:
                    s1 = [v for v in mb[y] if v not in (x, z)]
                    s2 = [v for v in mb[z] if v not in (x, y)]
                    sx = s1 if len(s1) < len(s2) else s2
                    res = backend.allsubs(y, z, sx, [x], test, alpha, 0, min(max_sx, len(sx)))
                    if res[0] <= alpha:
                        out.append((res[2], y, x, z))
    return out


def boot_arc_counts(engine, columns, nlevels, ids, score, arg=0.0, seed=0, maxp=np.inf, native=True,
                    workers=4, devices=None):
    """The replicates `ids` of boot_strength: -> (arc count matrix [from, to], tests).
    Replicate r resamples with the numpy PCG64 stream seed + r whatever share it is
    part of, so any split of the replicates over workers, devices or ranks sums to
    the same counts (the reference's bootstrap_reduce averages per-worker results,
    src/learning/averaging/bootstrap.c:126-186)."""
    from .callsites import TestCounter
    ids = list(ids)
    n = len(columns)
    nobs = len(columns[0])

    def rows_of(r):
        return np.random.Generator(np.random.PCG64(seed + r)).integers(0, nobs, nobs)

    if not native:
        arcs, tests = np.zeros((n, n)), 0.0
        for r in ids:
            rows = rows_of(r)
            ds = engine.register([c[rows] for c in columns], nlevels)
            be = GpuBackend(ds)
            amat, _ = hc(be, n, score, arg, maxp)
            tests += be.counter.value
            arcs += amat
            ds.release()
        return arcs, tests
    from .engine import Engine

    def run(engine_w, share):
        src = engine_w.register(columns, nlevels)
        acc, cnt = np.zeros((n, n)), 0.0
        for r in share:
            ds = src.resample(rows_of(r))
            counter = TestCounter()
            amat, _, _ = ds.hc(score, arg, maxp=maxp, counter=counter)
            cnt += counter.value
            acc += amat
            ds.release()
        src.release()
        return acc, cnt

    workers = max(1, min(int(workers), len(ids)))
    if workers == 1 and devices is None:
        return run(engine, ids)
    from concurrent.futures import ThreadPoolExecutor
    if devices is None:
        pool_engines, extra = [engine], [Engine(engine.devices) for _ in range(workers - 1)]
    else:
        pool_engines, extra = [], [Engine([int(devices[w % len(devices)])]) for w in range(workers)]
    pool_engines += extra
    try:
        with ThreadPoolExecutor(workers) as pool:
            parts = list(pool.map(lambda w: run(pool_engines[w], ids[w::workers]), range(workers)))
    finally:
        for e in extra:
            e.close()
    return sum(p[0] for p in parts), sum(p[1] for p in parts)


def strength_from_counts(arcs, replicates):
    """(strength, direction) from the arc counts of `replicates` networks
    (R/arc.strength.R:167-227 with cpdag = FALSE)."""
    prob = arcs / replicates
    strength = prob + prob.T
    with np.errstate(invalid="ignore", divide="ignore"):
        direction = np.where(strength > 0, prob / strength, 0.0)
    return strength, direction


def boot_strength(engine, columns, nlevels, replicates, score, arg=0.0, seed=0, maxp=np.inf,
                  native=True, workers=4, devices=None):
    """boot.strength(algorithm = "hc", cpdag = FALSE): arc strength / direction
    from hc() on bootstrap resamples (R/arc.strength.R:123-227). Resampling uses
    numpy PCG64 streams seed + r (R's own sample() stream needs R). Returns
    (strength, direction) matrices indexed [from, to] and the total test count.
    native: the data frame is uploaded once, each replicate is a device-side row
    gather (bnl_data_resample) searched by bnl_hc, and `workers` host threads, each
    with its own context (so its own stream), keep several replicates in flight:
    one replicate's kernels are far too small to fill the GPU. Otherwise every
    replicate is resampled on the host, registered and searched one call-site at
    a time. `devices` (CUDA ordinals) spreads the workers over several GPUs --
    replicate r goes to worker r mod workers, worker w to devices[w mod ndev]
    (the reference gives each cluster slave ceil(R/s) replicates,
    R/arc.strength.R:210) -- and `engine` is then only used when devices is None.
    The result depends neither on `workers` nor on `devices`."""
    arcs, tests = boot_arc_counts(engine, columns, nlevels, range(replicates), score, arg, seed, maxp, native,
                                  workers, devices)
    strength, direction = strength_from_counts(arcs, replicates)
    return strength, direction, tests


# ------------------------ pairwise mutual information -----------------------

def aracne(ds, nnodes, estimator="mi"):
    """ARACNE (src/learning/local/mi.matrix.c:255-412, no white/blacklist): the edge
    i - j is dropped when some k has MI(i, j) < min(MI(i, k), MI(j, k)). The
    coefficient matrix comes from one engine call. Returns sorted (i, j), i < j."""
    mi = ds.mi_matrix(estimator, list(range(nnodes)))
    keep = []
    for i in range(nnodes):
        for j in range(i + 1, nnodes):
            weaker = (mi[i, j] < mi[i]) & (mi[i, j] < mi[j])
            weaker[i] = weaker[j] = False
            if not weaker.any():
                keep.append((i, j))
    return keep


def chow_liu(ds, nnodes, estimator="mi", cols=None, cond=-1):
    """Chow-Liu tree (src/learning/local/mi.matrix.c:425-594): edges in decreasing
    order of mutual information, skipping those that would close a cycle, until
    nnodes - 1 are in. `cols`/`cond` serve TAN (features given the class). The
    reference sorts with R_qsort_I and walks the result backwards; coefficients
    that tie exactly could come out in another order here."""
    cols = list(range(nnodes)) if cols is None else list(cols)
    p = len(cols)
    mi = ds.mi_matrix(estimator, cols, cond)
    iu, ju = np.triu_indices(p, 1)
    order = np.argsort(mi[iu, ju], kind="stable")[::-1]
    comp = list(range(p))                      # union-find: a path exists iff same component

    def find(a):
        while comp[a] != a:
            comp[a] = comp[comp[a]]
            a = comp[a]
        return a
    edges = []
    for o in order:
        if len(edges) >= p - 1:
            break
        a, b = find(int(iu[o])), find(int(ju[o]))
        if a == b:
            continue
        comp[a] = b
        edges.append((int(iu[o]), int(ju[o])))
    if len(edges) != p - 1:
        raise ValueError("learned %d arcs instead of %d, this is not a tree spanning all the nodes."
                         % (len(edges), p - 1))
    return sorted(edges)


def tree_directions(edges, nnodes, root):
    """Orient a tree away from `root` (src/learning/local/mi.matrix.c:597-676)."""
    adj = {i: [] for i in range(nnodes)}
    for a, b in edges:
        adj[a].append(b)
        adj[b].append(a)
    depth = {root: 1}
    frontier = [root]
    while frontier:
        nxt = []
        for a in frontier:
            for b in adj[a]:
                if b not in depth:
                    depth[b] = depth[a] + 1
                    nxt.append(b)
        frontier = nxt
    return sorted((a, b) if depth[a] < depth[b] else (b, a) for a, b in edges)
