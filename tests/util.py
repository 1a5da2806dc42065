"""Shared helpers for the test-suite: fixture data sets in the reference's
conventions (int32 1-based codes, float64 columns)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_discrete(name):
    """-> (names, [int32 1-based column, ...], nlevels)"""
    z = np.load(os.path.join(GOLDEN, "data_%s.npz" % name))
    codes = z["codes"].astype(np.int32) + 1
    codes[z["codes"] == 255] = np.iinfo(np.int32).min          # NA in the fixtures with missing cells
    return [str(s) for s in z["names"]], [np.ascontiguousarray(c) for c in codes], \
        [int(v) for v in z["nlevels"]]


def load_gaussian(name):
    z = np.load(os.path.join(GOLDEN, "data_%s.npz" % name))
    return [str(s) for s in z["names"]], [np.ascontiguousarray(c) for c in z["values"]]


def load_mixed(name):
    """-> (names, columns, nlevels) of a data set with factors (int32 1-based codes)
    and numeric columns (float64); nlevels is 0 for the numeric ones."""
    z = np.load(os.path.join(GOLDEN, "data_%s.npz" % name))
    nlev = [int(v) for v in z["nlevels"]]
    cols, fi, gi = [], 0, 0
    for l in nlev:
        if l > 0:
            cols.append(np.ascontiguousarray(z["codes"][fi].astype(np.int32) + 1))
            fi += 1
        else:
            cols.append(np.ascontiguousarray(z["values"][gi]))
            gi += 1
    return [str(s) for s in z["names"]], cols, nlev


def rel_close(a, b, rtol=1e-9, atol=0.0):
    a, b = float(a), float(b)
    if a == b:
        return True
    if np.isinf(a) or np.isinf(b) or np.isnan(a) or np.isnan(b):
        return (np.isnan(a) and np.isnan(b)) or a == b
    return abs(a - b) <= max(atol, rtol * max(abs(a), abs(b)))
