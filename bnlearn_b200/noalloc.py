"""Python face of bnl_ci_discrete_noalloc (include/bnl_b200.h): one discrete CI
test on columns that are ALREADY on the device, with no allocation by the
library. Meant for callers whose framework owns all device memory (a PyTorch or
RMM pool, a pre-reserved serving process) and for loops that must not touch the
allocator: they pass raw device addresses of their own buffers.
"""
import ctypes as C

import numpy as np

from . import _lib
from .engine import BnlError, TESTS

MAX_CELLS = 65536


def ci_discrete_noalloc(test, col_ptrs, nlevels, nobs, out_ptr, stream=0):
    """col_ptrs: device addresses (ints) of the packed uint8 columns x, y, z...;
    out_ptr: device address of 4 float64 (statistic, df, p-value, complete rows).
    The work is queued on `stream` (a raw cudaStream_t value) and not synchronised.
    All calls share one built-in scratch table: do not let two of them overlap on
    the GPU, and make the first call on a device outside any stream capture."""
    lib = _lib.load()
    k = len(col_ptrs)
    ptrs = (C.c_void_p * k)(*[int(p) for p in col_ptrs])
    lv = np.asarray(nlevels, dtype=np.int32)
    err = C.create_string_buffer(256)
    rc = lib.bnl_ci_discrete_noalloc(TESTS[test], k, ptrs, lv.ctypes.data_as(_lib.c_int_p), int(nobs),
                                     C.c_void_p(int(out_ptr)), C.c_void_p(int(stream)), err, len(err))
    if rc != 0:
        raise BnlError(rc, err.value.decode())
