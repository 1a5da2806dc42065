"""ctypes bindings to oracle/_ref/libbnref.so -- the reference's own C files
compiled against the header shim (oracle/Makefile). TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs,
never by bnlearn_b200/.

Two tiers:
  * kernel tier: plain-pointer functions (c_fast_config, c_chisqtest,
    c_cchisqtest, fill_*d_table, c_covmat, c_fast_pcor, ...);
  * call-site tier: the .Call entry points themselves (per_node_score,
    score_cache_fill, hc_opt_step, indep_test, allsubs_test, roundrobin_test)
    driven with SEXPs built through the mini-R shim, including test_counter.
"""
import ctypes as C
import glob
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

NILSXP, LGLSXP, INTSXP, REALSXP, STRSXP, VECSXP = 0, 10, 13, 14, 16, 19
c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)


def available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libbnref.so"))


def _load_openblas():
    import scipy
    pat = os.path.join(os.path.dirname(scipy.__file__), "..", "scipy.libs",
                       "libscipy_openblas*.so")
    hits = glob.glob(pat)
    if hits:
        C.CDLL(hits[0], mode=C.RTLD_GLOBAL)


class Counts3D(C.Structure):
    """counts3d of the reference (src/core/contingency.tables.h:27-38)."""
    _fields_ = [("llx", C.c_int), ("lly", C.c_int), ("llz", C.c_int), ("nobs", C.c_int),
                ("n", C.POINTER(C.POINTER(C.POINTER(C.c_int)))), ("ni", C.POINTER(C.POINTER(C.c_int))),
                ("nj", C.POINTER(C.POINTER(C.c_int))), ("nk", C.POINTER(C.c_int))]


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.path.join(_HERE, "_ref", "libbnref.so")
    if not os.path.exists(path):
        raise RuntimeError("oracle/_ref/libbnref.so missing: run `make -C oracle ref` "
                           "(needs /root/reference) or __graft_entry__.build()")
    _load_openblas()
    L = C.CDLL(path)
    vp = C.c_void_p
    sig = {
        "Rf_allocVector": (vp, [C.c_int, C.c_ssize_t]),
        "minir_vector_view": (vp, [C.c_int, vp, C.c_int]),
        "Rf_mkChar": (vp, [C.c_char_p]),
        "Rf_mkString": (vp, [C.c_char_p]),
        "Rf_install": (vp, [C.c_char_p]),
        "Rf_setAttrib": (vp, [vp, vp, vp]),
        "Rf_getAttrib": (vp, [vp, vp]),
        "SET_STRING_ELT": (None, [vp, C.c_ssize_t, vp]),
        "SET_VECTOR_ELT": (vp, [vp, C.c_ssize_t, vp]),
        "VECTOR_ELT": (vp, [vp, C.c_ssize_t]),
        "STRING_ELT": (vp, [vp, C.c_ssize_t]),
        "CHAR": (C.c_char_p, [vp]),
        "Rf_length": (C.c_int, [vp]),
        "TYPEOF": (C.c_int, [vp]),
        "INTEGER": (c_int_p, [vp]),
        "REAL": (c_dbl_p, [vp]),
        "Rf_ScalarReal": (vp, [C.c_double]),
        "Rf_ScalarInteger": (vp, [C.c_int]),
        "Rf_ScalarLogical": (vp, [C.c_int]),
        "minir_arena_reset": (None, []),
        "minir_warning_count": (C.c_int, []),
        "minir_warning_reset": (None, []),
        # kernel tier
        "c_fast_config": (None, [C.POINTER(c_int_p), C.c_int, C.c_int, c_int_p,
                                 c_int_p, c_int_p, C.c_int]),
        "c_chisqtest": (C.c_double, [c_int_p, C.c_int, c_int_p, C.c_int, C.c_int,
                                     c_dbl_p, C.c_int, C.c_bool]),
        "c_cchisqtest": (C.c_double, [c_int_p, C.c_int, c_int_p, C.c_int, c_int_p,
                                      C.c_int, C.c_int, c_dbl_p, C.c_int, C.c_bool]),
        "new_3d_table": (Counts3D, [C.c_int, C.c_int, C.c_int]),
        "fill_3d_table": (None, [c_int_p, c_int_p, c_int_p, C.POINTER(Counts3D), C.c_int]),
        "Free3DTAB": (None, [Counts3D]),
        "Rf_pchisq": (C.c_double, [C.c_double, C.c_double, C.c_int, C.c_int]),
        "Rf_pt": (C.c_double, [C.c_double, C.c_double, C.c_int, C.c_int]),
        "Rf_pnorm5": (C.c_double, [C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]),
        "Rf_lgammafn": (C.c_double, [C.c_double]),
        "Rf_dnorm4": (C.c_double, [C.c_double, C.c_double, C.c_double, C.c_int]),
        "c_mean": (C.c_double, [c_dbl_p, C.c_int]),
        "c_sse": (C.c_double, [c_dbl_p, C.c_double, C.c_int]),
        "c_fast_cor": (C.c_double, [c_dbl_p, c_dbl_p, C.c_int, C.c_double, C.c_double,
                                    C.c_longdouble, C.c_longdouble]),
        "c_ols": (None, [C.POINTER(c_dbl_p), c_dbl_p, C.c_int, C.c_int, c_dbl_p,
                         c_dbl_p, c_dbl_p, c_dbl_p, c_int_p, C.c_bool]),
        # permutation tests (src/tests/permutation/*.c); the uniform stream is the shim's
        "minir_rng_seed": (None, [C.c_ulonglong]),
        "minir_rng_peek": (C.c_ulonglong, []),
        "c_mcarlo": (None, [c_int_p, C.c_int, c_int_p, C.c_int, C.c_int, C.c_int, c_dbl_p, c_dbl_p,
                            C.c_double, C.c_int, c_dbl_p]),
        "c_cmcarlo": (None, [c_int_p, C.c_int, c_int_p, C.c_int, c_int_p, C.c_int, C.c_int, C.c_int,
                             c_dbl_p, c_dbl_p, C.c_double, C.c_int, c_dbl_p]),
        "c_gauss_mcarlo": (None, [c_dbl_p, c_dbl_p, C.c_int, C.c_int, c_dbl_p, C.c_double, C.c_int, c_dbl_p]),
        "c_gauss_cmcarlo": (None, [C.POINTER(c_dbl_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   c_dbl_p, c_dbl_p, C.c_double, C.c_int]),
        # call-site tier
        "per_node_score": (vp, [vp] * 6),
        "score_cache_fill": (vp, [vp] * 13),
        "hc_opt_step": (vp, [vp] * 10),
        "score_delta": (vp, [vp] * 9),
        "indep_test": (vp, [vp] * 9),
        "allsubs_test": (vp, [vp] * 12),
        "roundrobin_test": (vp, [vp] * 9),
        "aracne": (vp, [vp] * 6),
        "chow_liu": (vp, [vp] * 8),
        "reset_test_counter": (vp, []),
        "get_test_counter": (vp, []),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _LIB = L
    return L


def _sym(name):
    return C.c_void_p.in_dll(lib(), name).value


def nil():
    return _sym("R_NilValue")


# --------------------------- SEXP construction ---------------------------

_KEEPALIVE = []


def reset():
    """Release every SEXP built so far (and the numpy arrays they view)."""
    lib().minir_arena_reset()
    del _KEEPALIVE[:]


def strvec(items):
    L = lib()
    s = L.Rf_allocVector(STRSXP, len(items))
    for i, it in enumerate(items):
        L.SET_STRING_ELT(s, i, L.Rf_mkChar(it.encode()))
    return s


def set_names(x, names):
    lib().Rf_setAttrib(x, _sym("R_NamesSymbol"), strvec(names))
    return x


def rlist(items, names=None):
    L = lib()
    s = L.Rf_allocVector(VECSXP, len(items))
    for i, it in enumerate(items):
        L.SET_VECTOR_ELT(s, i, it)
    if names is not None:
        set_names(s, names)
    return s


def intvec(arr, view=True):
    a = np.ascontiguousarray(arr, dtype=np.int32)
    _KEEPALIVE.append(a)
    return lib().minir_vector_view(INTSXP, a.ctypes.data, a.size)


def realvec(arr):
    a = np.ascontiguousarray(arr, dtype=np.float64)
    _KEEPALIVE.append(a)
    return lib().minir_vector_view(REALSXP, a.ctypes.data, a.size)


def lglvec(arr):
    a = np.ascontiguousarray(arr, dtype=np.int32)
    _KEEPALIVE.append(a)
    return lib().minir_vector_view(LGLSXP, a.ctypes.data, a.size)


def factor(codes, nlevels):
    """int32 1-based codes (INT_MIN = NA) -> an R factor viewing the array."""
    L = lib()
    s = intvec(codes)
    L.Rf_setAttrib(s, _sym("R_LevelsSymbol"), strvec(["l%d" % i for i in range(nlevels)]))
    L.Rf_setAttrib(s, _sym("R_ClassSymbol"), L.Rf_mkString(b"factor"))
    return s


def data_frame(names, columns, nlevels=None):
    """columns: list of int32 (factor, with nlevels[i] > 0) or float64 arrays."""
    cols = []
    for i, c in enumerate(columns):
        if nlevels is not None and nlevels[i] > 0:
            cols.append(factor(c, int(nlevels[i])))
        else:
            cols.append(realvec(c))
    df = rlist(cols, names)
    lib().Rf_setAttrib(df, _sym("R_ClassSymbol"), lib().Rf_mkString(b"data.frame"))
    return df


def network(names, parents):
    """A `bn`-like list carrying what the score code reads: net$nodes[[n]]$parents
    (and $children). parents: {node: [parent names in order]}."""
    children = {n: [] for n in names}
    for n in names:
        for p in parents.get(n, []):
            children[p].append(n)
    nodes = [rlist([strvec(parents.get(n, [])), strvec(children[n])],
                   ["parents", "children"]) for n in names]
    return rlist([rlist(nodes, names)], ["nodes"])


def to_numpy(s):
    L = lib()
    n = L.Rf_length(s)
    t = L.TYPEOF(s)
    if t == REALSXP:
        return np.ctypeslib.as_array(L.REAL(s), shape=(n,)).copy() if n else np.zeros(0)
    if t in (INTSXP, LGLSXP):
        return np.ctypeslib.as_array(L.INTEGER(s), shape=(n,)).copy() if n else np.zeros(0, np.int32)
    if t == STRSXP:
        return [L.CHAR(L.STRING_ELT(s, i)).decode() for i in range(n)]
    if t == NILSXP:
        return None
    raise TypeError("cannot convert SEXP type %d" % t)


def get_attr(s, name):
    return lib().Rf_getAttrib(s, lib().Rf_install(name.encode()))


def test_counter():
    return float(to_numpy(lib().get_test_counter())[0])


def reset_test_counter():
    lib().reset_test_counter()
