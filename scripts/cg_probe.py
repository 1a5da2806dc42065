#!/usr/bin/env python
"""Conditional Gaussian path at scale: synthetic mixed data (factors + numeric
columns), a batch of cg score units, and a whole hc(bic-cg) search."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bnlearn_b200 import Engine  # noqa: E402

rows = int(os.environ.get("ROWS", 1_000_000))
nf, ng = 20, 20
rng = np.random.default_rng(11)
lev = [int(v) for v in rng.integers(2, 5, nf)] + [0] * ng
cols = [rng.integers(1, l + 1, rows).astype(np.int32) for l in lev[:nf]]
base = rng.normal(size=(rows, 4))
for j in range(ng):
    f = cols[j % nf]
    cols.append(base @ rng.normal(size=4) + 0.7 * f + rng.normal(size=rows))
with Engine() as eng:
    t0 = time.perf_counter()
    d = eng.register(cols, lev)
    t_reg = time.perf_counter() - t0
    # 400 units: numeric node, 1-3 discrete parents, 0-3 numeric parents
    tg, pa = [], []
    for u in range(400):
        node = nf + int(rng.integers(0, ng))
        ndp = int(os.environ.get("NDP", 0)) or int(rng.integers(1, 4))
        dp = [int(i) for i in rng.choice(nf, size=ndp, replace=False)]
        gp = [nf + int(i) for i in rng.choice([g for g in range(ng) if nf + g != node], size=int(rng.integers(0, 4)), replace=False)]
        tg.append(node); pa.append(gp + dp)
    d.score_cg("loglik-cg", tg, pa)
    t0 = time.perf_counter()
    sc, npar = d.score_cg("loglik-cg", tg, pa)
    t_units = time.perf_counter() - t0
    bytes_alg = sum((1 + sum(1 for c in p if lev[c] == 0)) * 8 + sum(1 for c in p if lev[c] > 0) for p in pa) * rows * 2
    t0 = time.perf_counter()
    amat, scores, iters = d.hc("bic-cg", np.log(rows) / 2, maxp=4)
    t_hc = time.perf_counter() - t0
    print(json.dumps({"rows": rows, "columns": "%d factors + %d numeric" % (nf, ng), "register_s": t_reg,
                      "cg_units": len(tg), "cg_units_s": t_units, "cg_units_per_s": len(tg) / t_units,
                      "algorithmic_GBps_two_passes": bytes_alg / t_units / 1e9,
                      "finite": int(np.isfinite(sc).sum()),
                      "hc_bic_cg_s": t_hc, "hc_iterations": int(iters), "hc_arcs": int(amat.sum())}))
