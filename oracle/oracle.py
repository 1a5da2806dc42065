"""ctypes front end of oracle/_build/liboracle.so (the C restatement in
oracle/bnl_oracle.c). TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs, never by the product.

Columns are passed as lists of numpy arrays in the reference's conventions:
int32 1-based factor codes (INT_MIN = NA) or float64.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
NA_INT = np.iinfo(np.int32).min
c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)

TESTS = {"mi": 1, "mi-adf": 2, "x2": 3, "x2-adf": 4, "jt": 10, "mi-sh": 40, "cor": 20, "zf": 21, "mi-g": 22, "mi-cg": 30, "mi-g-sh": 41}
SCORES = {"loglik": 1, "aic": 3, "bic": 4, "bde": 6, "bds": 7, "bdj": 8, "k2": 9, "fnml": 11, "qnml": 12, "nal": 13, "pnal": 14,
          "loglik-g": 100, "aic-g": 102, "bic-g": 103, "nal-g": 106, "pnal-g": 107}


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.path.join(_HERE, "_build", "liboracle.so")
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    pp_i = C.POINTER(c_int_p)
    pp_d = C.POINTER(c_dbl_p)
    L.orc_config.restype = C.c_longlong
    L.orc_config.argtypes = [pp_i, C.c_int, C.c_int, c_int_p, c_int_p, C.c_int]
    L.orc_counts.restype = None
    L.orc_counts.argtypes = [c_int_p, C.c_int, c_int_p, C.c_int, c_int_p, C.c_int, C.c_int, c_int_p]
    L.orc_discrete_stat.restype = C.c_double
    L.orc_discrete_stat.argtypes = [c_int_p, C.c_int, C.c_int, C.c_int, C.c_int, c_dbl_p]
    L.orc_ci_discrete.restype = C.c_int
    L.orc_ci_discrete.argtypes = [C.c_int, c_int_p, C.c_int, c_int_p, C.c_int, pp_i, c_int_p,
                                  C.c_int, C.c_int, c_dbl_p, c_dbl_p, c_dbl_p]
    L.orc_score_discrete.restype = C.c_double
    L.orc_score_discrete.argtypes = [C.c_int, c_int_p, C.c_int, pp_i, c_int_p, C.c_int, C.c_int,
                                     C.c_double, c_dbl_p]
    L.orc_robust_difference.restype = C.c_double
    L.orc_robust_difference.argtypes = [C.c_double, C.c_double]
    L.orc_ci_gaussian.restype = C.c_int
    L.orc_ci_gaussian.argtypes = [C.c_int, pp_d, C.c_int, C.c_int, c_dbl_p, c_dbl_p, c_dbl_p]
    L.orc_score_gaussian.restype = C.c_double
    L.orc_score_gaussian.argtypes = [C.c_int, c_dbl_p, pp_d, C.c_int, C.c_int, C.c_double, c_dbl_p]
    L.orc_score_bge.restype = C.c_double
    L.orc_score_bge.argtypes = [pp_d, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, c_dbl_p]
    L.orc_cg_loglik.restype = C.c_double
    L.orc_cg_loglik.argtypes = [c_dbl_p, pp_d, C.c_int, c_int_p, C.c_int, C.c_int, c_dbl_p]
    L.orc_pcor.restype = C.c_double
    L.orc_pcor.argtypes = [c_dbl_p, C.c_int, c_int_p]
    for name, nargs in (("Rf_lgammafn", 1),):
        getattr(L, name).restype = C.c_double
        getattr(L, name).argtypes = [C.c_double] * nargs
    L.Rf_pchisq.restype = C.c_double
    L.Rf_pchisq.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
    L.Rf_pt.restype = C.c_double
    L.Rf_pt.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
    L.Rf_pnorm5.restype = C.c_double
    L.Rf_pnorm5.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
    _LIB = L
    return L


def _ip(a):
    return a.ctypes.data_as(c_int_p)


def _dp(a):
    return a.ctypes.data_as(c_dbl_p)


def _icols(cols):
    cs = [np.ascontiguousarray(c, dtype=np.int32) for c in cols]
    return cs, (c_int_p * max(len(cs), 1))(*[_ip(c) for c in cs])


def _dcols(cols):
    cs = [np.ascontiguousarray(c, dtype=np.float64) for c in cols]
    return cs, (c_dbl_p * max(len(cs), 1))(*[_dp(c) for c in cs])


def config(cols, levels, offset=0):
    cs, ptrs = _icols(cols)
    n = len(cs[0])
    out = np.empty(n, dtype=np.int32)
    lv = np.asarray(levels, dtype=np.int32)
    nl = lib().orc_config(ptrs, n, len(cs), _ip(lv), _ip(out), offset)
    return out, nl


def counts(x, llx, y=None, lly=1, zcols=(), zlev=()):
    """n[z][x][y] as a (llz, llx, lly) int32 array."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    n = len(x)
    llz, cfg = 1, None
    if len(zcols):
        cfg, llz = config(zcols, zlev, 1)
    yy = np.ascontiguousarray(y, dtype=np.int32) if y is not None else None
    tab = np.zeros(llz * llx * (lly if y is not None else 1), dtype=np.int32)
    lib().orc_counts(_ip(x), llx, _ip(yy) if yy is not None else None, lly,
                     _ip(cfg) if cfg is not None else None, llz, n, _ip(tab))
    return tab.reshape(llz, llx, lly if y is not None else 1)


def ci_discrete(test, x, llx, y, lly, zcols=(), zlev=()):
    """(statistic, df, p.value) of one discrete CI test."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    cs, ptrs = _icols(zcols)
    lv = np.asarray(list(zlev) or [1], dtype=np.int32)
    st, df, p = C.c_double(), C.c_double(), C.c_double()
    rc = lib().orc_ci_discrete(TESTS[test], _ip(x), llx, _ip(y), lly, ptrs, _ip(lv), len(cs),
                               len(x), C.byref(st), C.byref(df), C.byref(p))
    if rc != 0:
        raise OverflowError("configuration count does not fit an int")
    return st.value, df.value, p.value


def score_discrete(score, x, llx, pcols=(), plev=(), arg=0.0):
    """(score, nparams) of one discrete node given its parents (in order)."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    cs, ptrs = _icols(pcols)
    lv = np.asarray(list(plev) or [1], dtype=np.int32)
    npar = C.c_double()
    v = lib().orc_score_discrete(SCORES[score], _ip(x), llx, ptrs, _ip(lv), len(cs), len(x),
                                 float(arg), C.byref(npar))
    return v, npar.value


def ci_gaussian(test, cols):
    """cols = [x, y, z1, ...]; returns (statistic, df, p.value, zero_df_warning)."""
    cs, ptrs = _dcols(cols)
    st, df, p = C.c_double(), C.c_double(), C.c_double()
    w = lib().orc_ci_gaussian(TESTS[test], ptrs, len(cs), len(cs[0]), C.byref(st), C.byref(df),
                              C.byref(p))
    return st.value, df.value, p.value, bool(w)


def score_gaussian(score, x, pcols=(), k=0.0):
    x = np.ascontiguousarray(x, dtype=np.float64)
    cs, ptrs = _dcols(pcols)
    npar = C.c_double()
    v = lib().orc_score_gaussian(SCORES[score], _dp(x), ptrs, len(cs), len(x), float(k),
                                 C.byref(npar))
    return v, npar.value


def score_bge(x, pcols, data_ncols, iss_mu, iss_w, nu):
    """BGe score of one node; nu = prior means of [x] + parents."""
    cs, ptrs = _dcols([x] + list(pcols))
    nu = np.ascontiguousarray(nu, dtype=np.float64)
    return lib().orc_score_bge(ptrs, len(cs) - 1, len(cs[0]), int(data_ncols), float(iss_mu),
                               float(iss_w), _dp(nu))


def score_cg(score, x, xlev, pcols, plev, k=0.0):
    """loglik-cg / aic-cg / bic-cg of one node (src/scores/cg.loglikelihood.c:119-199):
    xlev / plev are the numbers of levels, 0 for numeric columns. Returns (score, nparams)."""
    base = score[:-3]                                   # loglik / aic / bic
    disc = [i for i, l in enumerate(plev) if l > 0]
    cont = [i for i, l in enumerate(plev) if l == 0]
    if xlev > 0:
        if cont:
            return -np.inf, 0.0
        return score_discrete(base, x, xlev, pcols, plev, k)
    if not disc:
        return score_gaussian(base + "-g", x, pcols, k)
    cfg, nconfig = config([pcols[i] for i in disc], [plev[i] for i in disc], 1)
    xs, _ = _dcols([x])
    gs, gptrs = _dcols([pcols[i] for i in cont])
    npar = C.c_double()
    v = lib().orc_cg_loglik(_dp(xs[0]), gptrs, len(cont), _ip(cfg), int(nconfig), len(xs[0]), C.byref(npar))
    if base != "loglik":
        v -= k * npar.value
    return v, npar.value


def ci_cg(x, xlev, y, ylev, zcols=(), zlev=()):
    """mi-cg on mixed data (src/tests/patterns/utest.c:195-273, ctest.c:232-334 over
    src/tests/conditional.gaussian/cg.mutual.information.c): (statistic, df, p.value).
    Levels are 0 for numeric columns."""
    n = len(x)
    zd = [i for i, l in enumerate(zlev) if l > 0]
    zc = [i for i, l in enumerate(zlev) if l == 0]
    llz = int(np.prod([zlev[i] for i in zd])) if zd else 0
    dcols, dlev = [zcols[i] for i in zd], [zlev[i] for i in zd]
    ccols = [zcols[i] for i in zc]

    def L(node, regs, dc, dl):
        return score_cg("loglik-cg", node, 0, list(regs) + list(dc), [0] * len(regs) + list(dl))[0]

    def diff(num, den):
        return (num - den) / n if (np.isfinite(num) and np.isfinite(den)) else 0.0

    if xlev > 0 and ylev > 0:
        st, df, _ = ci_discrete("mi", x, xlev, y, ylev, dcols, dlev)
        if zc:
            lognum, logden = st / (2.0 * n), 0.0
            for i in range(len(ccols)):
                rest = ccols[i + 1:]
                alt = L(ccols[i], rest, [x], [xlev])
                lognum += diff(alt, L(ccols[i], rest, dcols, dlev))
                logden += diff(alt, L(ccols[i], rest, [y] + dcols, [ylev] + dlev))
            r = (lognum - logden) / n if (np.isfinite(lognum) and np.isfinite(logden)) else 0.0
            st = 2.0 * n * n * r
            df = df + llz * (len(zc) * (len(zc) + 3) / 2.0) * (ylev - 1)
    elif xlev == 0 and ylev == 0:
        if not len(zcols):
            st, df, _, _ = ci_gaussian("mi-g", [x, y])
        else:
            st = 2.0 * n * diff(L(y, [x] + ccols, dcols, dlev), L(y, ccols, dcols, dlev))
            df = 1.0 if llz == 0 else float(llz)
    else:
        resp, fac, flev = (y, x, xlev) if xlev > 0 else (x, y, ylev)
        st = 2.0 * n * diff(L(resp, ccols, [fac] + dcols, [flev] + dlev), L(resp, ccols, dcols, dlev))
        df = 2.0 * (flev - 1) if not len(zcols) else (flev - 1) * (1 if llz == 0 else llz) * (len(zc) + 2.0)
    return st, df, lib().Rf_pchisq(st, df, 0, 0)


def robust_difference(old, new):
    return lib().orc_robust_difference(float(old), float(new))
