"""Host-side mirror of the reference's call sites for this path, over the C ABI
(include/bnl_b200.h, "call sites"; implemented in csrc/bnl_callsites.cu).

Same names and argument meaning as the reference's R-level wrappers
(R/test.R:3-58, R/scores.R, R/hill-climbing.R:68-81), with columns given by
0-based index instead of name. The test counter is an explicit object: it is
the reference's global `test_counter` (src/test.counter.c:4).
"""
import ctypes as C

import numpy as np

from . import _lib
from .engine import MC_TESTS, SCORES, TESTS, csr


class TestCounter:
    """The reference's global test counter (reset.test.counter / test.counter)."""
    __test__ = False

    def __init__(self):
        self._v = C.c_double(0.0)

    def reset(self):
        self._v.value = 0.0

    @property
    def value(self):
        return self._v.value

    def add(self, n):
        self._v.value += n

    def ref(self):
        return C.byref(self._v)


class UniformStream:
    """The uniform stream handed to the permutation tests: the library's splitmix64
    on `seed` (the state carries over from call to call, like R's .Random.seed), or a
    Python callable `unif(state) -> float in (0, 1)`. `B` permutations per test."""

    def __init__(self, lib, B, seed=1, unif=None):
        self.B = int(B)
        if unif is not None:
            self._cb = C.CFUNCTYPE(C.c_double, C.c_void_p)(unif)
            self.fn, self.state, self._seed = C.cast(self._cb, C.c_void_p), None, None
        else:
            self._seed = C.c_uint64(int(seed))
            self.fn = C.cast(lib.bnl_unif_splitmix64, C.c_void_p)
            self.state = C.cast(C.pointer(self._seed), C.c_void_p)

    @property
    def position(self):
        return None if self._seed is None else self._seed.value


def _ip(a):
    return a.ctypes.data_as(_lib.c_int_p)


def _dp(a):
    return a.ctypes.data_as(_lib.c_dbl_p)


def indep_test(ds, x, y, sx, test, counter, stream=None, alpha=1.0):
    """indep.test(x, y, sx, learning = TRUE): p-values of each x[i] vs y given sx.
    Permutation tests (mc-*, smc-*, sp-*) need a UniformStream; alpha is the
    early-stopping threshold of the smc-* tests."""
    e = ds.engine
    xs = np.ascontiguousarray([x] if np.isscalar(x) else x, dtype=np.int32)
    s = np.ascontiguousarray(list(sx) or [0], dtype=np.int32)
    pv = np.zeros(max(len(xs), 1))
    if test in MC_TESTS:
        e._check(e._lib.bnl_indep_test_mc(e._ctx, ds.handle, MC_TESTS[test], stream.B, float(alpha), stream.fn,
                                          stream.state, len(xs), _ip(xs), int(y), _ip(s), len(sx), _dp(pv),
                                          None, None, counter.ref()))
        return pv[:len(xs)]
    e._check(e._lib.bnl_indep_test(e._ctx, ds.handle, TESTS[test], len(xs), _ip(xs), int(y), _ip(s),
                                   len(sx), _dp(pv), None, None, counter.ref()))
    return pv[:len(xs)]


def ci_test(ds, x, y, sx, test, counter):
    """ci.test(): (statistic, df, p.value) of one test (the htest path)."""
    e = ds.engine
    xs = np.array([x], dtype=np.int32)
    s = np.ascontiguousarray(list(sx) or [0], dtype=np.int32)
    pv, st, df = np.zeros(1), C.c_double(), C.c_double()
    e._check(e._lib.bnl_indep_test(e._ctx, ds.handle, TESTS[test], 1, _ip(xs), int(y), _ip(s),
                                   len(sx), _dp(pv), C.byref(st), C.byref(df), counter.ref()))
    return st.value, df.value, float(pv[0])


def allsubs_test_batch(ds, x, y, sx, nfixed, test, alpha, minsize, maxsize, counter):
    """Many allsubs.test() calls at once. sx[r] lists the fixed variables first.
    Returns a list of (p.value, min.p.value, max.p.value, dsep or None)."""
    e = ds.engine
    n = len(x)
    xs = np.ascontiguousarray(x, dtype=np.int32)
    ys = np.ascontiguousarray(y, dtype=np.int32)
    off, idx = csr(sx)
    nf = np.ascontiguousarray(nfixed, dtype=np.int32)
    out = np.zeros(3 * max(n, 1))
    dsep = np.zeros(max(int(off[-1]), 1), dtype=np.int32)
    nds = np.zeros(max(n, 1), dtype=np.int32)
    e._check(e._lib.bnl_allsubs_test(e._ctx, ds.handle, TESTS[test], n, _ip(xs), _ip(ys), _ip(off),
                                     _ip(idx), _ip(nf), int(minsize), int(maxsize), float(alpha),
                                     _dp(out), _ip(off), _ip(dsep), _ip(nds), counter.ref()))
    res = []
    for r in range(n):
        d = None if nds[r] < 0 else [int(v) for v in dsep[off[r]:off[r] + nds[r]]]
        res.append((out[3 * r], out[3 * r + 1], out[3 * r + 2], d))
    return res


def allsubs_test(ds, x, y, sx, fixed, test, alpha, minsize, maxsize, counter, stream=None):
    """allsubs.test(x, y, sx, fixed, ...) (R/test.R:22-39): sx excludes `fixed`;
    max is capped at length(sx) as the R wrapper does."""
    full = list(fixed) + list(sx)
    if test in MC_TESTS:
        e = ds.engine
        s = np.ascontiguousarray(full or [0], dtype=np.int32)
        out, dsep, nds = np.zeros(3), np.zeros(max(len(full), 1), dtype=np.int32), np.zeros(1, dtype=np.int32)
        e._check(e._lib.bnl_allsubs_test_mc(e._ctx, ds.handle, MC_TESTS[test], stream.B, stream.fn, stream.state,
                                            int(x), int(y), _ip(s), len(full), len(fixed), int(minsize),
                                            int(min(maxsize, len(sx))), float(alpha), _dp(out), _ip(dsep),
                                            _ip(nds), counter.ref()))
        return out[0], out[1], out[2], None if nds[0] < 0 else [int(v) for v in dsep[:nds[0]]]
    return allsubs_test_batch(ds, [x], [y], [full], [len(fixed)], test, alpha, minsize,
                              min(maxsize, len(sx)), counter)[0]


def roundrobin_test(ds, x, z, fixed, test, alpha, counter, stream=None):
    """roundrobin.test(x, z, fixed, ...): p-values in the order of the non-fixed z."""
    e = ds.engine
    if len(z) == 0:
        return np.zeros(0)
    zz = np.ascontiguousarray(z, dtype=np.int32)
    fx = np.array([1 if v in set(fixed) else 0 for v in z], dtype=np.int32)
    pv = np.zeros(max(len(z) - int(fx.sum()), 1))
    if test in MC_TESTS:
        e._check(e._lib.bnl_roundrobin_test_mc(e._ctx, ds.handle, MC_TESTS[test], stream.B, stream.fn,
                                               stream.state, int(x), _ip(zz), len(zz), _ip(fx), float(alpha),
                                               _dp(pv), counter.ref()))
        return pv[:len(z) - int(fx.sum())]
    e._check(e._lib.bnl_roundrobin_test(e._ctx, ds.handle, TESTS[test], int(x), _ip(zz), len(zz),
                                        _ip(fx), float(alpha), _dp(pv), counter.ref()))
    return pv[:len(z) - int(fx.sum())]


def per_node_score(ds, targets, parents, score, arg=0.0):
    """per.node.score(): no test-counter side effect (src/scores/per.node.score.c:31-48)."""
    return ds.score(score, targets, parents, arg)[0]


def _f_int32(a):
    return a if (a.dtype == np.int32 and a.flags.f_contiguous) else np.asfortranarray(a, dtype=np.int32)


def score_cache_fill(ds, score, arg, amat, blmat, updated, equivalence, reference, cache, counter):
    """score_cache_fill(): amat / blmat / cache are n x n arrays indexed [from, to];
    cache is updated in place and returned. Fortran-ordered int32 / float64 arrays
    are passed straight through (the reference's matrices are column-major too);
    anything else is converted, which costs a copy per call."""
    e = ds.engine
    n = amat.shape[0]
    a = _f_int32(amat)
    b = _f_int32(blmat)
    direct = cache.dtype == np.float64 and cache.flags.f_contiguous
    c = cache if direct else np.asfortranarray(cache, dtype=np.float64)
    upd = np.ascontiguousarray(updated, dtype=np.int32)
    ref = np.ascontiguousarray(reference, dtype=np.float64)
    e._check(e._lib.bnl_score_cache_fill(
        e._ctx, ds.handle, SCORES[score], float(arg), n, a.ctypes.data_as(_lib.c_int_p),
        b.ctypes.data_as(_lib.c_int_p), _ip(upd) if len(upd) else None, len(upd), int(bool(equivalence)),
        _dp(ref), c.ctypes.data_as(_lib.c_dbl_p), counter.ref()))
    if not direct:
        cache[...] = c
    return cache
