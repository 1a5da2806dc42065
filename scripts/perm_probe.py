#!/usr/bin/env python
"""Permutation tests at config B's row count: a batch of mc-mi tests through
bnl_ci_discrete_mc (tables counted on the GPU, random tables on one host core)
next to the reference's c_mcarlo on the same columns (one host core: table pass +
log-factorial table + the same random tables). Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bnlearn_b200 import Engine  # noqa: E402

rows = int(os.environ.get("ROWS", "10000000"))
B = int(os.environ.get("B", "100"))
rng = np.random.default_rng(20261111)
lev = [3, 2, 4, 3, 2, 3, 4, 2]
cols = [rng.integers(1, l + 1, rows).astype(np.int32) for l in lev]
cols[1] = ((cols[0] + (rng.random(rows) < 0.7) * rng.integers(0, 2, rows)) % 2 + 1).astype(np.int32)
units = [(i, j, []) for i in range(4) for j in range(4, 8)] + [(0, 1, [2]), (3, 5, [6, 7])]
out = {"rows": rows, "B": B, "tests": len(units)}
with Engine() as e:
    d = e.register(cols, lev)
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    d.ci_discrete_mc("mc-mi", xs[:2], ys[:2], zs[:2], 10, seed=1)          # warm-up (log-factorials, buffers)
    t0 = time.perf_counter()
    st, df, pv = d.ci_discrete_mc("mc-mi", xs, ys, zs, B, seed=7)
    out["engine_s"] = time.perf_counter() - t0
    out["engine_tests_per_s"] = len(units) / out["engine_s"]
    t0 = time.perf_counter()
    d.ci_discrete("mi", xs, ys, zs)
    out["counting_only_s"] = time.perf_counter() - t0
    d.release()
try:
    from oracle import refapi as ra
    nref = int(os.environ.get("NREF", "3"))
    t0 = time.perf_counter()
    ref = []
    for k, (x, y, z) in enumerate(units[:nref]):
        ref.append(ra.mc_discrete("mc-mi", cols[x], lev[x], cols[y], lev[y], [cols[i] for i in z],
                                  [lev[i] for i in z], B=B, alpha=1.0, seed=7 if k == 0 else None))
    out["reference_s_per_test"] = (time.perf_counter() - t0) / nref
    out["reference_tests_per_s_1core"] = 1.0 / out["reference_s_per_test"]
    out["identical_to_reference"] = all(st[k] == ref[k][0] and pv[k] == ref[k][2] for k in range(nref))
except Exception as exc:                                                  # oracle/_ref not built
    out["reference"] = "unavailable: %s" % exc
print(json.dumps(out))
