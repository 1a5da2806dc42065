"""Loader for libbnl_b200.so (the CUDA engine behind include/bnl_b200.h).

There is no CPU fallback: if the library is missing, or no CUDA device is
present when a context is created, this raises instead of degrading.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbnl_b200.so")
SPECIAL_HOST_PATH = os.path.join(_HERE, "libbnl_special_host.so")

c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)
c_i64_p = C.POINTER(C.c_int64)
c_i32_p = C.POINTER(C.c_int32)

# every symbol include/bnl_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "bnl_init": (C.c_int, [C.POINTER(C.c_void_p), c_int_p, C.c_int]),
    "bnl_shutdown": (None, [C.c_void_p]),
    "bnl_last_error": (C.c_char_p, [C.c_void_p]),
    "bnl_last_call_profile": (C.c_int, [C.c_void_p, c_dbl_p, c_dbl_p, c_dbl_p, c_int_p]),
    "bnl_data_register": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), c_int_p, c_int_p,
                                    C.c_int64, C.POINTER(C.c_uint64)]),
    "bnl_data_release": (C.c_int, [C.c_void_p, C.c_uint64]),
    "bnl_data_flush": (C.c_int, [C.c_void_p]),
    "bnl_counts": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p,
                             c_i32_p, c_i64_p]),
    "bnl_score_discrete": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p,
                                     c_int_p, c_int_p, c_dbl_p, c_dbl_p]),
    "bnl_data_set_interventions": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int]),
    "bnl_ci_discrete": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, c_int_p,
                                  c_int_p, c_int_p, c_dbl_p, c_dbl_p, c_dbl_p]),
    "bnl_ci_gaussian": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, c_int_p,
                                  c_int_p, c_int_p, c_dbl_p, c_dbl_p, c_dbl_p]),
    "bnl_unif_splitmix64": (C.c_double, [C.c_void_p]),
    "bnl_mc_from_tables": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, c_int_p, c_int_p, c_int_p,
                                     c_i32_p, c_i64_p, C.c_void_p, C.c_void_p, c_dbl_p, c_dbl_p, c_dbl_p]),
    "bnl_ci_gaussian_mc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_double, C.c_int, c_int_p,
                                     c_int_p, c_int_p, c_int_p, C.c_void_p, C.c_void_p, c_dbl_p, c_dbl_p]),
    "bnl_indep_test_mc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p,
                                    C.c_int, c_int_p, C.c_int, c_int_p, C.c_int, c_dbl_p, c_dbl_p, c_dbl_p, c_dbl_p]),
    "bnl_allsubs_test_mc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_int, c_int_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, c_dbl_p,
                                      c_int_p, c_int_p, c_dbl_p]),
    "bnl_roundrobin_test_mc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                         c_int_p, C.c_int, c_int_p, C.c_double, c_dbl_p, c_dbl_p]),
    "bnl_ci_discrete_mc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_double, C.c_int, c_int_p,
                                     c_int_p, c_int_p, c_int_p, C.c_void_p, C.c_void_p, c_dbl_p, c_dbl_p,
                                     c_dbl_p]),
    "bnl_ci_gaussian_allpairs": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p, c_dbl_p,
                                           c_int_p, c_int_p, C.c_longlong, C.POINTER(C.c_longlong), c_dbl_p]),
    "bnl_score_gaussian": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p,
                                     c_int_p, c_int_p, c_dbl_p, c_dbl_p]),
    "bnl_gaussian_moments": (C.c_int, [C.c_void_p, C.c_uint64, c_dbl_p, c_dbl_p]),
    "bnl_data_set_moments": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int64, c_dbl_p, c_dbl_p]),
    "bnl_indep_test": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, C.c_int, c_int_p,
                                 C.c_int, c_dbl_p, c_dbl_p, c_dbl_p, c_dbl_p]),
    "bnl_allsubs_test": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p,
                                   c_int_p, c_int_p, C.c_int, C.c_int, C.c_double, c_dbl_p, c_int_p,
                                   c_int_p, c_int_p, c_dbl_p]),
    "bnl_roundrobin_test": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, C.c_int,
                                      c_int_p, C.c_double, c_dbl_p, c_dbl_p]),
    "bnl_data_alloc": (C.c_int, [C.c_void_p, C.c_int, c_int_p, c_int_p, C.c_int64, C.POINTER(C.c_uint64)]),
    "bnl_data_device_view": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_size_t), C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "bnl_data_commit": (C.c_int, [C.c_void_p, C.c_uint64]),
    "bnl_data_resample": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int64, c_int_p, C.POINTER(C.c_uint64)]),
    "bnl_stat_from_tables": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int64, C.c_int, c_int_p,
                                       c_int_p, c_int_p, c_int_p, c_int_p, C.POINTER(C.c_int64), c_dbl_p,
                                       c_dbl_p, c_dbl_p]),
    "bnl_ci_cg": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p, c_dbl_p, c_dbl_p,
                            c_dbl_p]),
    "bnl_score_cg": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p, c_int_p, c_int_p,
                               c_dbl_p, c_dbl_p]),
    "bnl_pc_stable": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, C.c_int, c_int_p, c_dbl_p,
                                c_int_p, C.POINTER(C.c_longlong), c_int_p, C.c_longlong, c_dbl_p]),
    "bnl_tabu": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p, c_int_p, c_int_p,
                           C.c_double, C.c_longlong, C.c_int, C.c_int, c_int_p, c_dbl_p,
                           C.POINTER(C.c_longlong), c_dbl_p]),
    "bnl_set_ebic_gamma": (C.c_int, [C.c_void_p, C.c_uint64, C.c_double]),
    "bnl_set_bge_prior": (C.c_int, [C.c_void_p, C.c_uint64, C.c_double, C.c_double, c_dbl_p]),
    "bnl_hc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p, c_int_p, c_int_p,
                         C.c_double, C.c_longlong, c_int_p, c_dbl_p, C.POINTER(C.c_longlong), c_dbl_p]),
    "bnl_mi_matrix": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, c_int_p, C.c_int, c_dbl_p]),
    "bnl_ci_discrete_noalloc": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_void_p), c_int_p, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]),
    "bnl_score_cache_fill": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_int, c_int_p,
                                       c_int_p, c_int_p, C.c_int, C.c_int, c_dbl_p, c_dbl_p, c_dbl_p]),
}

_LIB = None


def build(verbose=False):
    """Compile the CUDA library for sm_100a (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc")], stdout=out)


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "bnlearn_b200: %s is missing. Build it with `make -C bnlearn_b200/csrc` "
            "(or __graft_entry__.build()); there is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if the ABI and the header disagree
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def load_special_host():
    """Host build of csrc/bnl_special.cuh (same code the kernels run)."""
    lib = C.CDLL(SPECIAL_HOST_PATH)
    for name, n in (("bnl_host_lgamma", 1), ("bnl_host_pchisq_upper", 2),
                    ("bnl_host_pt_two_sided", 2), ("bnl_host_pnorm_two_sided", 1)):
        f = getattr(lib, name)
        f.restype = C.c_double
        f.argtypes = [C.c_double] * n
    return lib
