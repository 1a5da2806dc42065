"""High-level helpers over oracle/refbind.py: run the reference's own .Call
entry points (compiled in oracle/_ref/libbnref.so) on numpy data.
TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py CPU legs).

A RefData wraps a data set as the "R data frame" the reference receives, with
int32 1-based factor codes (INT_MIN = NA) or float64 columns.
"""
import ctypes as C

import numpy as np

from . import refbind as rb


class RefData:
    def __init__(self, columns, nlevels=None, names=None):
        self.ncols = len(columns)
        self.names = names or ["V%d" % i for i in range(self.ncols)]
        self.nlevels = list(nlevels) if nlevels is not None else [0] * self.ncols
        self.columns = [np.ascontiguousarray(c, dtype=np.int32 if self.nlevels[i] > 0 else np.float64)
                        for i, c in enumerate(columns)]
        self.nobs = len(self.columns[0]) if self.ncols else 0
        self._df = None
        self._complete_flags = None

    def df(self):
        if self._df is None:
            self._df = rb.data_frame(self.names, self.columns, self.nlevels)
        return self._df

    def complete(self):
        if self._complete_flags is None:       # one pass over the data, like check.data() in R
            flags = []
            for i, c in enumerate(self.columns):
                if self.nlevels[i] > 0:
                    flags.append(int(not np.any(c == np.iinfo(np.int32).min)))
                else:
                    flags.append(int(not np.any(np.isnan(c))))
            self._complete_flags = flags
        flags = self._complete_flags
        s = rb.lglvec(np.array(flags, dtype=np.int32))
        rb.set_names(s, self.names)
        return s

    def invalidate(self):
        self._df = None


def _extra(score=None, k=None, iss=None, B=None, bge=None, gamma=None, exp=None):
    names, items = [], []
    L = rb.lib()
    if exp is not None:                       # {node name: sorted 1-based rows that were interventions on it}
        names.append("exp")
        items.append(rb.rlist([rb.intvec(np.asarray(v, dtype=np.int32), view=False) for v in exp.values()],
                              list(exp.keys())))
    if gamma is not None:
        names.append("gamma"); items.append(L.Rf_ScalarReal(float(gamma)))
    if bge is not None:                       # (iss.mu, iss.w, nu vector, node names)
        iss_mu, iss_w, nu, nodes = bge
        nu_s = rb.realvec(np.asarray(nu, dtype=np.float64))
        rb.set_names(nu_s, list(nodes))
        names += ["iss.mu", "iss.w", "nu", "prior"]
        items += [L.Rf_ScalarReal(float(iss_mu)), L.Rf_ScalarReal(float(iss_w)), nu_s,
                  L.Rf_mkString(b"uniform")]
    if B is not None:                         # permutations of the mc-* / smc-* / sp-* tests
        names.append("B"); items.append(L.Rf_ScalarInteger(int(B)))
    if k is not None:
        names.append("k"); items.append(L.Rf_ScalarReal(float(k)))
    if iss is not None:
        names.append("iss"); items.append(L.Rf_ScalarReal(float(iss)))
        names.append("prior"); items.append(L.Rf_mkString(b"uniform"))
    return rb.rlist(items, names)


def per_node_score(data, parents, targets, score, k=None, iss=None, bge=None, gamma=None, exp=None):
    """reference per_node_score (src/scores/per.node.score.c:31-48).
    parents: {node name: [parent names, in order]}; bge = (iss.mu, iss.w, nu)."""
    L = rb.lib()
    net = rb.network(data.names, parents)
    if bge is not None:
        bge = (bge[0], bge[1], bge[2], data.names)
    res = L.per_node_score(net, data.df(), L.Rf_mkString(score.encode()),
                           rb.strvec(list(targets)), _extra(k=k, iss=iss, bge=bge, gamma=gamma, exp=exp),
                           L.Rf_ScalarLogical(0))
    return rb.to_numpy(res)


def indep_test(data, x, y, sx, test, alpha=0.05, B=None):
    """reference indep_test(..., learning = TRUE) (src/tests/rinterface/indep.test.c:5-19):
    x may be several names; returns the p-value vector. Bumps test_counter."""
    L = rb.lib()
    xs = [x] if isinstance(x, str) else list(x)
    res = L.indep_test(rb.strvec(xs), rb.strvec([y]), rb.strvec(list(sx)), data.df(),
                       L.Rf_mkString(test.encode()), L.Rf_ScalarReal(alpha),
                       _extra(B=B), L.Rf_ScalarLogical(1), data.complete())
    return rb.to_numpy(res)


def ci_test(data, x, y, sx, test):
    """reference indep_test(..., learning = FALSE): returns (statistic, df, p.value)
    from the htest object (src/tests/rinterface/htest.c:6-89)."""
    L = rb.lib()
    res = L.indep_test(rb.strvec([x]), rb.strvec([y]), rb.strvec(list(sx)), data.df(),
                       L.Rf_mkString(test.encode()), L.Rf_ScalarReal(0.05),
                       rb.rlist([], []), L.Rf_ScalarLogical(0), data.complete())
    n = L.Rf_length(res)
    names = rb.to_numpy(rb.get_attr(res, "names"))
    out = {}
    for i in range(n):
        el = L.VECTOR_ELT(res, i)
        if L.TYPEOF(el) == rb.REALSXP:
            out[names[i]] = float(rb.to_numpy(el)[0])
    return out.get("statistic"), out.get("parameter"), out.get("p.value")


def allsubs_test(data, x, y, sx, fixed, test, alpha, minsize, maxsize, B=None):
    """reference allsubs_test (src/tests/rinterface/allsubs.test.c:10-210).
    Returns (p.value, min.p.value, max.p.value, dsep.set or None)."""
    L = rb.lib()
    res = L.allsubs_test(rb.strvec([x]), rb.strvec([y]), rb.strvec(list(sx)),
                         rb.strvec(list(fixed)), data.df(), L.Rf_mkString(test.encode()),
                         L.Rf_ScalarReal(alpha), _extra(B=B),
                         L.Rf_ScalarInteger(minsize), L.Rf_ScalarInteger(maxsize),
                         data.complete(), L.Rf_ScalarLogical(0))
    p = rb.to_numpy(res)
    ds = rb.to_numpy(rb.get_attr(res, "dsep.set"))
    return float(p[0]), float(p[1]), float(p[2]), ds


def roundrobin_test(data, x, z, fixed, test, alpha, B=None):
    """reference roundrobin_test (src/tests/rinterface/roundrobin.test.c:10-154)."""
    L = rb.lib()
    res = L.roundrobin_test(rb.strvec([x]), rb.strvec(list(z)), rb.strvec(list(fixed)),
                            data.df(), L.Rf_mkString(test.encode()), L.Rf_ScalarReal(alpha),
                            _extra(B=B), data.complete(), L.Rf_ScalarLogical(0))
    return rb.to_numpy(res)


def score_cache_fill(data, parents, score, reference, equivalence, updated, amat,
                     cache, blmat, k=None, iss=None):
    """reference score_cache_fill (src/learning/score/hc.cache.lookup.c:10-128).
    amat/cache/blmat: column-major n x n numpy arrays (cache modified in place);
    updated: 0-based node indices; reference: per-node scores."""
    L = rb.lib()
    n = data.ncols
    net = rb.network(data.names, parents)
    ref = rb.realvec(np.asarray(reference, dtype=np.float64))
    rb.set_names(ref, data.names)
    cache_f = np.asfortranarray(cache, dtype=np.float64)
    cache_s = rb.lib().minir_vector_view(rb.REALSXP, cache_f.ctypes.data, n * n)
    amat_s = rb.intvec(np.asfortranarray(amat, dtype=np.int32).ravel(order="F"))
    bl_s = rb.intvec(np.asfortranarray(blmat, dtype=np.int32).ravel(order="F"))
    L.score_cache_fill(rb.strvec(data.names), data.df(), net,
                       L.Rf_mkString(score.encode()), _extra(k=k, iss=iss), ref,
                       L.Rf_ScalarLogical(int(equivalence)), L.Rf_ScalarLogical(1),
                       rb.intvec(np.asarray(updated, dtype=np.int32)), amat_s, cache_s,
                       bl_s, L.Rf_ScalarLogical(0))
    cache[...] = cache_f
    return cache


def hc_opt_step(names, amat, added, cache, reference, wlmat, blmat, nparents, maxp):
    """reference hc_opt_step (src/learning/score/hc.cache.lookup.c:131-363).
    Returns (op or None, from, to); `reference` is updated in place."""
    L = rb.lib()
    n = len(names)
    f = lambda m: rb.intvec(np.asfortranarray(m, dtype=np.int32).ravel(order="F"))
    ref_a = np.ascontiguousarray(reference, dtype=np.float64)
    ref_s = L.minir_vector_view(rb.REALSXP, ref_a.ctypes.data, n)
    cache_f = np.asfortranarray(cache, dtype=np.float64)
    cache_s = L.minir_vector_view(rb.REALSXP, cache_f.ctypes.data, n * n)
    res = L.hc_opt_step(f(amat), rb.strvec(names), f(added), cache_s, ref_s, f(wlmat),
                        f(blmat), rb.realvec(np.asarray(nparents, dtype=np.float64)),
                        rb.realvec(np.array([maxp], dtype=np.float64)), L.Rf_ScalarLogical(0))
    reference[...] = ref_a
    op = L.VECTOR_ELT(res, 0)
    if L.TYPEOF(op) != rb.STRSXP:
        return None, None, None
    return (rb.to_numpy(op)[0], rb.to_numpy(L.VECTOR_ELT(res, 1))[0],
            rb.to_numpy(L.VECTOR_ELT(res, 2))[0])


def _arcs(res, names):
    """(2m x 2) character matrix of both directions -> sorted list of (i, j), i < j."""
    flat = rb.to_numpy(res)
    m = len(flat) // 2
    pairs = {tuple(sorted((names.index(flat[i]), names.index(flat[i + m])))) for i in range(m)}
    return sorted(pairs)


def aracne(data, estimator="mi"):
    """reference aracne (src/learning/local/mi.matrix.c:255-412), no white/blacklist."""
    L = rb.lib()
    nil = rb.nil()
    res = L.aracne(data.df(), L.Rf_mkString(estimator.encode()), nil, nil, data.complete(),
                   L.Rf_ScalarLogical(0))
    return _arcs(res, data.names)


def chow_liu(data, estimator="mi", conditional=None, clevels=0):
    """reference chow_liu (src/learning/local/mi.matrix.c:425-594); `conditional` is
    the class variable of TAN (int32 1-based codes) or None."""
    L = rb.lib()
    nil = rb.nil()
    cond = nil if conditional is None else rb.factor(np.ascontiguousarray(conditional, dtype=np.int32), clevels)
    res = L.chow_liu(data.df(), rb.strvec(data.names), L.Rf_mkString(estimator.encode()), nil, nil,
                     data.complete(), cond, L.Rf_ScalarLogical(0))
    return _arcs(res, data.names)


# ------------------------------ kernel tier ------------------------------

def _ip(a):
    return a.ctypes.data_as(rb.c_int_p)


def _dp(a):
    return a.ctypes.data_as(rb.c_dbl_p)


def fast_config(columns, levels, offset=0):
    """reference c_fast_config (src/core/sets.c:162-218)."""
    L = rb.lib()
    cols = [np.ascontiguousarray(c, dtype=np.int32) for c in columns]
    n = len(cols[0])
    ptrs = (rb.c_int_p * len(cols))(*[_ip(c) for c in cols])
    lv = np.asarray(levels, dtype=np.int32)
    out = np.empty(n, dtype=np.int32)
    nl = C.c_int(0)
    L.c_fast_config(ptrs, n, len(cols), _ip(lv), _ip(out), C.byref(nl), offset)
    return out, nl.value


def chisqtest(x, llx, y, lly, test_e, z=None, llz=0):
    """reference c_chisqtest / c_cchisqtest (src/tests/discrete/discrete.tests.c:11-94):
    returns (statistic, df) with the MI statistic scaled to G^2."""
    L = rb.lib()
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    df = C.c_double(0)
    scale = test_e in (1, 2)
    if z is None:
        st = L.c_chisqtest(_ip(x), llx, _ip(y), lly, len(x), C.byref(df), test_e, scale)
    else:
        z = np.ascontiguousarray(z, dtype=np.int32)
        st = L.c_cchisqtest(_ip(x), llx, _ip(y), lly, _ip(z), llz, len(x), C.byref(df),
                            test_e, scale)
    return st, df.value


def tables3d(x, llx, y, lly, z=None, llz=1):
    """reference new_3d_table + fill_3d_table (src/core/contingency.tables.c:170-223):
    the joint counts n[z][x][y] as an int32 array of that shape, plus the number of
    complete rows. z: 1-based configuration vector (c_fast_config with offset 1);
    None counts the marginal table in a single stratum."""
    L = rb.lib()
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    z = np.ones(len(x), dtype=np.int32) if z is None else np.ascontiguousarray(z, dtype=np.int32)
    tab = L.new_3d_table(int(llx), int(lly), int(llz))
    L.fill_3d_table(_ip(x), _ip(y), _ip(z), C.byref(tab), len(x))
    out = np.empty((llz, llx, lly), dtype=np.int32)
    for k in range(llz):
        for i in range(llx):
            out[k, i, :] = np.ctypeslib.as_array(tab.n[k][i], shape=(lly,))
    nobs = int(tab.nobs)
    L.Free3DTAB(tab)
    return out, nobs


# test_e values of the permutation tests (src/tests/tests.h:21-34)
MC_TESTS = {"mc-mi": 50, "mc-x2": 51, "sp-mi": 52, "sp-x2": 53, "mc-jt": 54,
            "smc-mi": 60, "smc-x2": 61, "smc-jt": 62,
            "mc-cor": 70, "mc-mi-g": 71, "mc-zf": 72, "smc-cor": 80, "smc-mi-g": 81, "smc-zf": 82}


def mc_discrete(test, x, llx, y, lly, zcols=(), zlev=(), B=500, alpha=0.05, seed=1):
    """reference c_mcarlo / c_cmcarlo (src/tests/permutation/discrete.monte.carlo.c:16-159,
    162-308) on the shim's uniform stream seeded with `seed`.
    Returns (statistic, df or nan, p-value)."""
    L = rb.lib()
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    obs, pv, df = (np.full(1, np.nan) for _ in range(3))
    if seed is not None:                      # None: carry on with the stream of the previous test
        L.minir_rng_seed(int(seed))
    if len(zcols) == 0:
        L.c_mcarlo(_ip(x), llx, _ip(y), lly, len(x), int(B), _dp(obs), _dp(pv), float(alpha),
                   MC_TESTS[test], _dp(df))
    else:
        z, llz = fast_config(zcols, zlev, 1)
        L.c_cmcarlo(_ip(x), llx, _ip(y), lly, _ip(z), llz, len(x), int(B), _dp(obs), _dp(pv),
                    float(alpha), MC_TESTS[test], _dp(df))
    return float(obs[0]), float(df[0]), float(pv[0])


def mc_gaussian(test, cols, B=500, alpha=0.05, seed=1):
    """reference c_gauss_mcarlo / c_gauss_cmcarlo (src/tests/permutation/gaussian.monte.carlo.c):
    cols = [x, y, z1, ...]; returns (statistic, p-value)."""
    L = rb.lib()
    cs = [np.array(c, dtype=np.float64) for c in cols]
    n = len(cs[0])
    obs, pv = np.full(1, np.nan), np.full(1, np.nan)
    if seed is not None:
        L.minir_rng_seed(int(seed))
    if len(cs) == 2:
        L.c_gauss_mcarlo(_dp(cs[0]), _dp(cs[1]), n, int(B), _dp(pv), float(alpha), MC_TESTS[test], _dp(obs))
    else:
        ptrs = (rb.c_dbl_p * len(cs))(*[_dp(c) for c in cs])
        L.c_gauss_cmcarlo(ptrs, len(cs), n, 0, 1, int(B), _dp(obs), _dp(pv), float(alpha), MC_TESTS[test])
    return float(obs[0]), float(pv[0])
