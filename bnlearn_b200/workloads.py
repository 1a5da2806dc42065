"""Fixed unit lists for the throughput benchmarks.

pc_stable_units() is the list of CI tests pc.stable() issues on alarm-shaped data
at levels 0..max_level when the skeleton frozen for levels >= 1 is the true
one: level 0 tests every node pair marginally (R/pcalgo.R:8-28); level L tests
each surviving pair (x, y) against every size-L subset of nbr(x) \\ {y} and then
of nbr(y) \\ {x}, in the lexicographic order of first_subset / next_subset
(src/core/sets.c:9-55; R/pcalgo.R:88-143). True edges never separate, so no
early exit shortens the list: it is an upper bound of what one call does.
"""
from itertools import combinations

from . import synth


def skeleton(names, parents):
    idx = {n: i for i, n in enumerate(names)}
    nbr = {i: set() for i in range(len(names))}
    for child, pas in parents.items():
        for p in pas:
            nbr[idx[child]].add(idx[p])
            nbr[idx[p]].add(idx[child])
    return nbr


def pc_stable_units(names, levels, max_level=2, model=None):
    """-> list of (x, y, [z...]) column-index triples."""
    parents = synth.parse_modelstring(model or synth.ALARM_MODEL)
    nbr = skeleton(names, parents)
    n = len(names)
    units = [(i, j, []) for i in range(n) for j in range(i + 1, n)]
    edges = [(i, j) for i in range(n) for j in range(i + 1, n) if j in nbr[i]]
    for level in range(1, max_level + 1):
        for (x, y) in edges:
            for a, b in ((x, y), (y, x)):
                cand = sorted(nbr[a] - {b})
                if len(cand) < level:
                    continue
                for s in combinations(cand, level):
                    units.append((a, b, list(s)))
    return units


def hc_iteration_units(nnodes, parents_idx, target):
    """The score_cache_fill batch after an arc into `target` changed
    (src/learning/score/hc.cache.lookup.c:54-117): one unit per other node i,
    parents(target) + i if i is not a parent, parents(target) - i otherwise."""
    pa = list(parents_idx)
    units = []
    for i in range(nnodes):
        if i == target:
            continue
        if i in pa:
            units.append((target, [p for p in pa if p != i]))
        else:
            units.append((target, pa + [i]))
    return units
