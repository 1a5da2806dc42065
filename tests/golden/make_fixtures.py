"""Generate the committed fixtures under tests/golden/.

Runs ONLY in the build container (needs /root/reference and
oracle/_ref/libbnref.so): `python tests/golden/make_fixtures.py`.

Outputs
  data_<name>.npz       the reference's bundled data frames re-encoded as numpy
                        (uint8 0-based codes + level counts, or float64), from
                        /root/reference/data/<name>.rda via rda_reader.py
  ref_outputs.json      outputs of the reference's OWN C entry points
                        (per_node_score, indep_test, allsubs_test, roundrobin_test,
                        score_cache_fill + test_counter) on those data sets
  special_functions.json  mpmath values for lgamma / chi-square, t and normal
                        upper tails on the decision-relevant ranges
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from rda_reader import data_frame_columns, read_rda  # noqa: E402

REF = "/root/reference"


def export_datasets():
    for name in ("learning.test", "alarm", "insurance", "gaussian.test"):
        obj = read_rda(os.path.join(REF, "data", name + ".rda"))[name]
        names, cols = data_frame_columns(obj)
        out = os.path.join(HERE, "data_%s.npz" % name.replace(".", "_"))
        if isinstance(cols[0], tuple):
            codes = np.stack([c[0] for c in cols]).astype(np.int64) - 1
            assert codes.min() >= 0 and codes.max() < 255
            np.savez_compressed(out, names=np.array(names), codes=codes.astype(np.uint8),
                                nlevels=np.array([len(c[1]) for c in cols], dtype=np.int32))
        else:
            np.savez_compressed(out, names=np.array(names),
                                values=np.stack(cols).astype(np.float64))
        print("wrote", out)


def load_discrete(name):
    z = np.load(os.path.join(HERE, "data_%s.npz" % name))
    return [str(s) for s in z["names"]], z["codes"].astype(np.int32) + 1, z["nlevels"]


def load_gaussian(name):
    z = np.load(os.path.join(HERE, "data_%s.npz" % name))
    return [str(s) for s in z["names"]], z["values"]


def reference_outputs():
    from oracle import refapi as ra, refbind as rb
    out = {}
    rng = np.random.default_rng(20261106)

    # ---- discrete: learning.test and alarm -------------------------------
    for ds in ("learning_test", "alarm"):
        names, codes, nlev = load_discrete(ds)
        data = ra.RefData(list(codes), list(nlev), names)
        n, p = data.nobs, len(names)
        rec = {"tests": [], "scores": [], "allsubs": [], "roundrobin": []}
        # CI tests: random (x, y, S) with |S| in 0..3
        for t in range(40 if ds == "alarm" else 30):
            k = int(rng.integers(0, 4))
            idx = rng.choice(p, size=2 + k, replace=False)
            x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
            for test in ("mi", "x2", "mi-adf", "x2-adf"):
                st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in s], test)
                rec["tests"].append({"test": test, "x": x, "y": y, "z": s,
                                     "stat": st, "df": df, "p": pv})
            rb.reset()
            data.invalidate()
        # node scores: random parent sets with 0..3 parents
        for t in range(30):
            k = int(rng.integers(0, 4))
            idx = rng.choice(p, size=1 + k, replace=False)
            node, pa = int(idx[0]), [int(i) for i in idx[1:]]
            par = {names[node]: [names[i] for i in pa]}
            for score, kw in (("loglik", {}), ("aic", {"k": 1.0}), ("bic", {"k": np.log(n) / 2}),
                              ("bde", {"iss": 1.0}), ("bde", {"iss": 10.0}), ("bds", {"iss": 1.0}),
                              ("k2", {}), ("bdj", {})):
                v = ra.per_node_score(data, par, [names[node]], score, **kw)[0]
                rec["scores"].append({"score": score, "arg": list(kw.values())[0] if kw else 0.0,
                                      "node": node, "parents": pa, "value": float(v)})
            rb.reset()
            data.invalidate()
        # allsubs (with test counter) and round-robin
        for t in range(12):
            m = int(rng.integers(2, min(6, p - 2) + 1)) if p > 6 else int(rng.integers(2, p - 1))
            idx = rng.choice(p, size=2 + m, replace=False)
            x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
            nf = int(rng.integers(0, 2))
            mn = int(rng.integers(0, 2))
            mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
            for test in ("mi", "x2"):
                rb.reset_test_counter()
                pv, pmin, pmax, ds_ = ra.allsubs_test(
                    data, names[x], names[y], [names[i] for i in s], [names[i] for i in s[:nf]],
                    test, 0.05, mn, mx)
                rec["allsubs"].append({"test": test, "x": x, "y": y, "sx": s, "nfixed": nf,
                                       "min": mn, "max": mx, "alpha": 0.05, "p": pv,
                                       "pmin": pmin, "pmax": pmax,
                                       "dsep": None if ds_ is None else [names.index(d) for d in ds_],
                                       "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
                rb.reset_test_counter()
                pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + s],
                                         [names[i] for i in s[:nf]], test, 0.05)
                rec["roundrobin"].append({"test": test, "x": x, "z": [y] + s, "nfixed": nf,
                                          "alpha": 0.05, "p": [float(v) for v in pvs],
                                          "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
            rb.reset()
            data.invalidate()
        out[ds] = rec

    # ---- score cache fill on learning.test (empty graph, then one arc) ----
    names, codes, nlev = load_discrete("learning_test")
    data = ra.RefData(list(codes), list(nlev), names)
    n, p = data.nobs, len(names)
    rec = []
    for score, kw in (("bic", {"k": np.log(n) / 2}), ("bde", {"iss": 1.0})):
        for parents in ({}, {"B": ["A"], "D": ["A", "C"]}):
            amat = np.zeros((p, p), dtype=np.int32)
            for ch, pas in parents.items():
                for pa in pas:
                    amat[names.index(pa), names.index(ch)] = 1
            ref = ra.per_node_score(data, parents, names, score, **kw)
            cache = np.zeros((p, p), dtype=np.float64, order="F")
            rb.reset_test_counter()
            ra.score_cache_fill(data, parents, score, ref, True, list(range(p)), amat, cache,
                                np.zeros((p, p), dtype=np.int32), **kw)
            rec.append({"score": score, "arg": list(kw.values())[0],
                        "parents": {k: v for k, v in parents.items()},
                        "reference": [float(v) for v in ref],
                        "cache": [[float(v) for v in row] for row in cache],
                        "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
            rb.reset()
            data.invalidate()
    out["score_cache_fill"] = rec

    # ---- Gaussian: gaussian.test ------------------------------------------
    names, vals = load_gaussian("gaussian_test")
    data = ra.RefData(list(vals), None, names)
    n, p = data.nobs, len(names)
    rec = {"tests": [], "scores": []}
    for t in range(30):
        k = int(rng.integers(0, 5))
        idx = rng.choice(p, size=2 + k, replace=False)
        x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        for test in ("cor", "zf", "mi-g", "mi-g-sh"):
            st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in s], test)
            rec["tests"].append({"test": test, "x": x, "y": y, "z": s, "stat": st, "df": df, "p": pv})
        rb.reset()
        data.invalidate()
    for t in range(30):
        k = int(rng.integers(0, 6))
        idx = rng.choice(p, size=1 + k, replace=False)
        node, pa = int(idx[0]), [int(i) for i in idx[1:]]
        par = {names[node]: [names[i] for i in pa]}
        for score, kw in (("loglik-g", {}), ("aic-g", {"k": 1.0}), ("bic-g", {"k": np.log(n) / 2})):
            v = ra.per_node_score(data, par, [names[node]], score, **kw)[0]
            rec["scores"].append({"score": score, "arg": list(kw.values())[0] if kw else 0.0,
                                  "node": node, "parents": pa, "value": float(v)})
        rb.reset()
        data.invalidate()
    out["gaussian_test"] = rec

    with open(os.path.join(HERE, "ref_outputs.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_outputs.json")


def incomplete_outputs():
    """learning.test with 8% of the cells of every column set to NA (seeded), then
    the reference's NA-aware arms: discrete CI tests on locally complete rows and
    the node-average likelihood scores. Separate file and RNG so that the
    complete-data fixtures above keep their values."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261107)
    names, codes, nlev = load_discrete("learning_test")
    codes = codes.astype(np.int64) - 1
    mask = rng.random(codes.shape) < 0.08
    packed = np.where(mask, 255, codes).astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "data_learning_test_na.npz"), names=np.array(names),
                        codes=packed, nlevels=np.asarray(nlev, dtype=np.int32))
    cols = [np.where(packed[i] == 255, np.iinfo(np.int32).min, packed[i].astype(np.int32) + 1)
            .astype(np.int32) for i in range(len(names))]
    data = ra.RefData(cols, list(nlev), names)
    n, p = data.nobs, len(names)
    rec = {"tests": [], "scores": [], "allsubs": [], "roundrobin": []}
    for t in range(24):
        k = int(rng.integers(0, 4))
        idx = rng.choice(p, size=2 + k, replace=False)
        x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        for test in ("mi", "x2", "mi-adf", "x2-adf", "jt", "mi-sh"):
            st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in s], test)
            rec["tests"].append({"test": test, "x": x, "y": y, "z": s, "stat": st,
                                 "df": 0.0 if df is None else df, "p": pv})
        rb.reset()
        data.invalidate()
    # the *_with_missing call-site patterns (allsubs.test.c:290-448, roundrobin.test.c:222-344)
    for t in range(8):
        m = int(rng.integers(2, p - 1))
        idx = rng.choice(p, size=2 + m, replace=False)
        x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        nf = int(rng.integers(0, 2))
        mn = int(rng.integers(0, 2))
        mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
        for test in ("mi", "x2"):
            rb.reset_test_counter()
            pv, pmin, pmax, ds_ = ra.allsubs_test(
                data, names[x], names[y], [names[i] for i in s], [names[i] for i in s[:nf]],
                test, 0.05, mn, mx)
            rec["allsubs"].append({"test": test, "x": x, "y": y, "sx": s, "nfixed": nf,
                                   "min": mn, "max": mx, "alpha": 0.05, "p": pv,
                                   "pmin": pmin, "pmax": pmax,
                                   "dsep": None if ds_ is None else [names.index(d) for d in ds_],
                                   "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
            rb.reset_test_counter()
            pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + s],
                                     [names[i] for i in s[:nf]], test, 0.05)
            rec["roundrobin"].append({"test": test, "x": x, "z": [y] + s, "nfixed": nf,
                                      "alpha": 0.05, "p": [float(v) for v in pvs],
                                      "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset()
        data.invalidate()
    for t in range(24):
        k = int(rng.integers(0, 4))
        idx = rng.choice(p, size=1 + k, replace=False)
        node, pa = int(idx[0]), [int(i) for i in idx[1:]]
        par = {names[node]: [names[i] for i in pa]}
        for score, kw in (("nal", {}), ("pnal", {"k": np.log(n) / 2}), ("pnal", {"k": 0.25})):
            v = ra.per_node_score(data, par, [names[node]], score, **kw)[0]
            rec["scores"].append({"score": score, "arg": list(kw.values())[0] if kw else 0.0,
                                  "node": node, "parents": pa, "value": float(v)})
        rb.reset()
        data.invalidate()
    # ---- gaussian.test (first 2000 rows) with 6% missing cells ----------------
    gnames, vals = load_gaussian("gaussian_test")
    vals = vals[:, :2000].copy()
    vals[rng.random(vals.shape) < 0.06] = np.nan
    np.savez_compressed(os.path.join(HERE, "data_gaussian_test_na.npz"), names=np.array(gnames),
                        values=vals)
    gdata = ra.RefData(list(vals), None, gnames)
    gp, gn = len(gnames), vals.shape[1]
    grec = {"tests": [], "scores": []}
    for t in range(24):
        k = int(rng.integers(0, 5))
        idx = rng.choice(gp, size=2 + k, replace=False)
        x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        for test in ("cor", "zf", "mi-g", "mi-g-sh"):
            st, df, pv = ra.ci_test(gdata, gnames[x], gnames[y], [gnames[i] for i in s], test)
            grec["tests"].append({"test": test, "x": x, "y": y, "z": s, "stat": st, "df": df, "p": pv})
        rb.reset()
        gdata.invalidate()
    for t in range(24):
        k = int(rng.integers(0, 6))
        idx = rng.choice(gp, size=1 + k, replace=False)
        node, pa = int(idx[0]), [int(i) for i in idx[1:]]
        par = {gnames[node]: [gnames[i] for i in pa]}
        for score, kw in (("nal-g", {}), ("pnal-g", {"k": np.log(gn) / 2})):
            v = ra.per_node_score(gdata, par, [gnames[node]], score, **kw)[0]
            grec["scores"].append({"score": score, "arg": list(kw.values())[0] if kw else 0.0,
                                   "node": node, "parents": pa, "value": float(v)})
        rb.reset()
        gdata.invalidate()
    with open(os.path.join(HERE, "ref_incomplete.json"), "w") as f:
        json.dump({"learning_test_na": rec, "gaussian_test_na": grec}, f)
    print("wrote ref_incomplete.json")


def jt_shrinkage_outputs():
    """Jonckheere-Terpstra and shrinkage-MI tests on complete data (own file and RNG)."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261108)
    out = {}
    for ds in ("learning_test", "alarm"):
        names, codes, nlev = load_discrete(ds)
        data = ra.RefData(list(codes), list(nlev), names)
        p = len(names)
        rec = {"data": ds, "tests": [], "scores": [], "allsubs": [], "roundrobin": []}
        for t in range(30):
            k = int(rng.integers(0, 4))
            idx = rng.choice(p, size=2 + k, replace=False)
            x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
            for test in ("jt", "mi-sh"):
                st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in s], test)
                rec["tests"].append({"test": test, "x": x, "y": y, "z": s, "stat": st,
                                     "df": 0.0 if df is None else df, "p": pv})
            rb.reset()
            data.invalidate()
        for t in range(6):
            m = int(rng.integers(2, min(5, p - 2) + 1))
            idx = rng.choice(p, size=2 + m, replace=False)
            x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
            nf = int(rng.integers(0, 2))
            mn = int(rng.integers(0, 2))
            mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
            for test in ("jt", "mi-sh"):
                rb.reset_test_counter()
                pv, pmin, pmax, ds_ = ra.allsubs_test(
                    data, names[x], names[y], [names[i] for i in s], [names[i] for i in s[:nf]],
                    test, 0.05, mn, mx)
                rec["allsubs"].append({"test": test, "x": x, "y": y, "sx": s, "nfixed": nf,
                                       "min": mn, "max": mx, "alpha": 0.05, "p": pv,
                                       "pmin": pmin, "pmax": pmax,
                                       "dsep": None if ds_ is None else [names.index(d) for d in ds_],
                                       "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
                rb.reset_test_counter()
                pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + s],
                                         [names[i] for i in s[:nf]], test, 0.05)
                rec["roundrobin"].append({"test": test, "x": x, "z": [y] + s, "nfixed": nf,
                                          "alpha": 0.05, "p": [float(v) for v in pvs],
                                          "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
            rb.reset()
            data.invalidate()
        out[ds + "_jt_sh"] = rec
    with open(os.path.join(HERE, "ref_jt_shrinkage.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_jt_shrinkage.json")


def nml_outputs():
    """fNML / qNML scores: learning.test and alarm (large-n regret approximations)
    and the first 600 rows of alarm (exact regret table, n <= 1000)."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261111)
    names, codes, nlev = load_discrete("alarm")
    np.savez_compressed(os.path.join(HERE, "data_alarm_600.npz"), names=np.array(names),
                        codes=(codes[:, :600] - 1).astype(np.uint8), nlevels=np.asarray(nlev, dtype=np.int32))
    out = {}
    for ds in ("learning_test", "alarm", "alarm_600"):
        names, codes, nlev = load_discrete(ds)
        data = ra.RefData(list(codes), list(nlev), names)
        p = len(names)
        rec = {"data": ds, "tests": [], "scores": [], "allsubs": [], "roundrobin": []}
        for t in range(30):
            k = int(rng.integers(0, 5))
            idx = rng.choice(p, size=1 + k, replace=False)
            node, pa = int(idx[0]), [int(i) for i in idx[1:]]
            par = {names[node]: [names[i] for i in pa]}
            for score in ("fnml", "qnml"):
                v = ra.per_node_score(data, par, [names[node]], score)[0]
                rec["scores"].append({"score": score, "arg": 0.0, "node": node, "parents": pa,
                                      "value": float(v)})
            rb.reset()
            data.invalidate()
        out[ds + "_nml"] = rec
    with open(os.path.join(HERE, "ref_nml.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_nml.json")


def load_mixed(name):
    """-> (names, columns, nlevels): int32 1-based codes for factors, float64 otherwise."""
    z = np.load(os.path.join(HERE, "data_%s.npz" % name))
    nlev = [int(v) for v in z["nlevels"]]
    cols, fi, gi = [], 0, 0
    for l in nlev:
        if l > 0:
            cols.append(z["codes"][fi].astype(np.int32) + 1)
            fi += 1
        else:
            cols.append(np.ascontiguousarray(z["values"][gi]))
            gi += 1
    return [str(s) for s in z["names"]], cols, nlev


def cg_outputs():
    """Conditional Gaussian scores on the bundled mixed data set clgaussian.test
    (4 factors, 4 numeric columns, 5000 rows); own file and RNG."""
    from oracle import refapi as ra, refbind as rb
    obj = read_rda(os.path.join(REF, "data", "clgaussian.test.rda"))["clgaussian.test"]
    names, cols = data_frame_columns(obj)
    nlev = [len(c[1]) if isinstance(c, tuple) else 0 for c in cols]
    codes = np.stack([np.asarray(c[0], dtype=np.int64) - 1 for c in cols if isinstance(c, tuple)])
    values = np.stack([np.asarray(c, dtype=np.float64) for c in cols if not isinstance(c, tuple)])
    np.savez_compressed(os.path.join(HERE, "data_clgaussian_test.npz"), names=np.array(names),
                        nlevels=np.asarray(nlev, dtype=np.int32), codes=codes.astype(np.uint8), values=values)
    names, columns, nlev = load_mixed("clgaussian_test")
    data = ra.RefData(columns, nlev, names)
    rng = np.random.default_rng(20261112)
    p, n = len(names), len(columns[0])
    recs = []
    for t in range(60):
        k = int(rng.integers(0, 6))
        idx = rng.choice(p, size=1 + k, replace=False)
        node, pa = int(idx[0]), [int(i) for i in idx[1:]]
        par = {names[node]: [names[i] for i in pa]}
        for score, kw in (("loglik-cg", {}), ("aic-cg", {"k": 1.0}), ("bic-cg", {"k": np.log(n) / 2})):
            v = ra.per_node_score(data, par, [names[node]], score, **kw)[0]
            recs.append({"score": score, "arg": list(kw.values())[0] if kw else 0.0, "node": node,
                         "parents": pa, "value": float(v)})
        rb.reset()
        data.invalidate()
    tests = []
    for t in range(120):
        k = int(rng.integers(0, 5))
        idx = rng.choice(p, size=2 + k, replace=False)
        x, y, s = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in s], "mi-cg")
        tests.append({"test": "mi-cg", "x": x, "y": y, "z": s, "stat": st, "df": df, "p": pv})
        if t % 10 == 9:
            rb.reset()
            data.invalidate()
    subs, rr = [], []
    for t in range(10):                      # ast_micg_complete / rrd_micg_complete
        m = int(rng.integers(2, 5))
        idx = [int(i) for i in rng.choice(p, size=2 + m, replace=False)]
        x, y, sx = idx[0], idx[1], idx[2:]
        nf = int(rng.integers(0, 2))
        mn = int(rng.integers(0, 2))
        mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
        rb.reset_test_counter()
        pv, pmin, pmax, ds_ = ra.allsubs_test(data, names[x], names[y], [names[i] for i in sx],
                                              [names[i] for i in sx[:nf]], "mi-cg", 0.05, mn, mx)
        subs.append({"x": x, "y": y, "sx": sx, "nfixed": nf, "min": mn, "max": mx, "alpha": 0.05, "p": pv,
                     "pmin": pmin, "pmax": pmax, "dsep": None if ds_ is None else [names.index(d) for d in ds_],
                     "counter": rb.test_counter()})
        rb.reset_test_counter()
        pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + sx], [names[i] for i in sx[:nf]],
                                 "mi-cg", 0.05)
        rr.append({"x": x, "z": [y] + sx, "nfixed": nf, "alpha": 0.05, "p": [float(v) for v in pvs],
                   "counter": rb.test_counter()})
        rb.reset()
        data.invalidate()
    with open(os.path.join(HERE, "ref_cg.json"), "w") as f:
        json.dump({"clgaussian_test": {"scores": recs, "tests": tests, "allsubs": subs, "roundrobin": rr}}, f)
    print("wrote ref_cg.json")


def inject_missing(cols, nlev, seed=20261113, frac=0.05):
    """The incomplete version of a mixed data set used by ref_cg_nal.json (the
    test re-creates it from the same seed)."""
    rng = np.random.default_rng(seed)
    out = []
    for c, l in zip(cols, nlev):
        c = c.copy()
        m = rng.random(len(c)) < frac
        c[m] = np.iinfo(np.int32).min if l > 0 else np.nan
        out.append(c)
    return out


def cg_nal_outputs():
    """nal-cg / pnal-cg (nal_cgnode, src/scores/cg.nal.c) on clgaussian.test with
    5 % of the cells missing; own file and RNG."""
    from oracle import refapi as ra, refbind as rb
    names, columns, nlev = load_mixed("clgaussian_test")
    columns = inject_missing(columns, nlev)
    data = ra.RefData(columns, nlev, names)
    rng = np.random.default_rng(20261114)
    p, n = len(names), len(columns[0])
    recs = []
    for t in range(60):
        k = int(rng.integers(0, 6))
        idx = rng.choice(p, size=1 + k, replace=False)
        node, pa = int(idx[0]), [int(i) for i in idx[1:]]
        par = {names[node]: [names[i] for i in pa]}
        for score, kw in (("nal-cg", {}), ("pnal-cg", {"k": np.log(n) / 2}), ("pnal-cg", {"k": 0.3})):
            v = ra.per_node_score(data, par, [names[node]], score, **kw)[0]
            recs.append({"score": score, "arg": list(kw.values())[0] if kw else 0.0, "node": node,
                         "parents": pa, "value": float(v)})
        rb.reset()
        data.invalidate()
    # mi-cg on the same incomplete data: ct_micg_with_missing tests on the rows complete for
    # all the columns of the test (src/tests/patterns/ctest.c:336-406)
    tests = []
    for t in range(60):
        k = int(rng.integers(0, 4))
        idx = rng.choice(p, size=2 + k, replace=False)
        x, y, sx = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
        st, df, pv = ra.ci_test(data, names[x], names[y], [names[i] for i in sx], "mi-cg")
        tests.append({"x": x, "y": y, "z": sx, "stat": st, "df": df, "p": pv})
        if t % 10 == 9:
            rb.reset()
            data.invalidate()
    subs, rr = [], []
    for t in range(10):
        m = int(rng.integers(2, 5))
        idx = [int(i) for i in rng.choice(p, size=2 + m, replace=False)]
        x, y, sx = idx[0], idx[1], idx[2:]
        nf = int(rng.integers(0, 2))
        mn = int(rng.integers(0, 2))
        mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
        rb.reset_test_counter()
        pv, pmin, pmax, ds_ = ra.allsubs_test(data, names[x], names[y], [names[i] for i in sx],
                                              [names[i] for i in sx[:nf]], "mi-cg", 0.05, mn, mx)
        subs.append({"x": x, "y": y, "sx": sx, "nfixed": nf, "min": mn, "max": mx, "alpha": 0.05, "p": pv,
                     "pmin": pmin, "pmax": pmax, "dsep": None if ds_ is None else [names.index(d) for d in ds_],
                     "counter": rb.test_counter()})
        rb.reset_test_counter()
        pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + sx], [names[i] for i in sx[:nf]],
                                 "mi-cg", 0.05)
        rr.append({"x": x, "z": [y] + sx, "nfixed": nf, "alpha": 0.05, "p": [float(v) for v in pvs],
                   "counter": rb.test_counter()})
        rb.reset()
        data.invalidate()
    with open(os.path.join(HERE, "ref_cg_nal.json"), "w") as f:
        json.dump({"clgaussian_test_na": {"seed": 20261113, "frac": 0.05, "scores": recs, "tests": tests,
                                          "allsubs": subs, "roundrobin": rr}}, f)
    print("wrote ref_cg_nal.json")


def ebic_outputs():
    """Extended BIC (ebic, ebic-g, ebic-cg) with the default gamma = 0.5 and with 0.3."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261113)
    out = {}
    for ds, score, loader in (("learning_test", "ebic", "d"), ("gaussian_test", "ebic-g", "g"),
                              ("clgaussian_test", "ebic-cg", "m")):
        if loader == "d":
            names, codes, nlev = load_discrete(ds)
            data = ra.RefData(list(codes), list(nlev), names)
        elif loader == "g":
            names, vals = load_gaussian(ds)
            data = ra.RefData(list(vals), None, names)
        else:
            names, cols, nlev = load_mixed(ds)
            data = ra.RefData(cols, nlev, names)
        p, n = len(names), data.nobs
        recs = []
        for t in range(20):
            k = int(rng.integers(0, 4))
            idx = rng.choice(p, size=1 + k, replace=False)
            node, pa = int(idx[0]), [int(i) for i in idx[1:]]
            for gamma in (0.5, 0.3):
                v = ra.per_node_score(data, {names[node]: [names[i] for i in pa]}, [names[node]], score,
                                      k=np.log(n) / 2, gamma=gamma)[0]
                recs.append({"score": score, "arg": float(np.log(n) / 2), "gamma": gamma, "node": node,
                             "parents": pa, "value": float(v)})
            rb.reset()
            data.invalidate()
        out[ds] = recs
    with open(os.path.join(HERE, "ref_ebic.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_ebic.json")


def bge_outputs():
    """BGe scores on gaussian.test with the front end's default hyper-parameters
    (iss.mu = 1, iss.w = ncol + 2, nu = column means) and with arbitrary ones."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261110)
    names, vals = load_gaussian("gaussian_test")
    data = ra.RefData(list(vals), None, names)
    p = len(names)
    recs = []
    settings = [(1.0, float(p + 2), [float(v.mean()) for v in vals]),
                (3.0, float(p + 5), [float(x) for x in rng.normal(size=p)])]
    for iss_mu, iss_w, nu in settings:
        for t in range(20):
            k = int(rng.integers(0, 6))
            idx = rng.choice(p, size=1 + k, replace=False)
            node, pa = int(idx[0]), [int(i) for i in idx[1:]]
            v = ra.per_node_score(data, {names[node]: [names[i] for i in pa]}, [names[node]], "bge",
                                  bge=(iss_mu, iss_w, nu))[0]
            recs.append({"iss_mu": iss_mu, "iss_w": iss_w, "nu": nu, "node": node, "parents": pa,
                         "value": float(v)})
        rb.reset()
        data.invalidate()
    with open(os.path.join(HERE, "ref_bge.json"), "w") as f:
        json.dump({"gaussian_test": recs}, f)
    print("wrote ref_bge.json")


def mi_matrix_outputs():
    """aracne / chow.liu (and the TAN flavour of chow.liu, given a class variable)
    from the reference on the bundled data sets; own file, no RNG."""
    from oracle import refapi as ra, refbind as rb
    out = {}
    for ds in ("learning_test", "alarm", "insurance"):
        names, codes, nlev = load_discrete(ds)
        data = ra.RefData(list(codes), list(nlev), names)
        rec = {"aracne": ra.aracne(data), "chow_liu": ra.chow_liu(data)}
        sub = ra.RefData(list(codes[1:]), list(nlev[1:]), names[1:])
        rec["chow_liu_given_first"] = ra.chow_liu(sub, conditional=codes[0], clevels=int(nlev[0]))
        out[ds] = rec
        rb.reset()
    names, vals = load_gaussian("gaussian_test")
    data = ra.RefData(list(vals), None, names)
    out["gaussian_test"] = {"aracne": ra.aracne(data, "mi-g"), "chow_liu": ra.chow_liu(data, "mi-g")}
    with open(os.path.join(HERE, "ref_mi_matrix.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_mi_matrix.json")


def incomplete_gaussian_patterns():
    """allsubs / roundrobin on incomplete numeric data: the reference's
    ast_gaustests_with_missing / rrd_gaustests_with_missing
    (patterns/allsubs.test.c:290-448, roundrobin.test.c:222-344). Three data sets
    built from data_gaussian_test_na.npz: every column incomplete; columns 0-2
    complete (requests on those alone take the complete-data patterns); and the
    first 9 rows only (degrees of freedom run out). Own file and RNG."""
    from oracle import refapi as ra, refbind as rb
    rng = np.random.default_rng(20261112)
    z = np.load(os.path.join(HERE, "data_gaussian_test_na.npz"))
    gnames, na = [str(v) for v in z["names"]], z["values"]
    _, full = load_gaussian("gaussian_test")
    mixed = na.copy()
    mixed[:3] = full[:3, :na.shape[1]]
    sets = {"all_incomplete": na, "mixed": mixed, "tiny": mixed[:, :9].copy()}
    out = {}
    p = len(gnames)
    for key, vals in sets.items():
        data = ra.RefData([np.ascontiguousarray(v) for v in vals], None, gnames)
        rec = {"allsubs": [], "roundrobin": []}
        for t in range(10):
            m = int(rng.integers(2, p - 1))
            idx = rng.choice(p, size=2 + m, replace=False)
            if key == "mixed" and t < 3:
                idx = rng.permutation(3)[:3]                 # complete columns only
                idx = np.concatenate([idx, rng.choice(np.arange(3, p), size=0)])
                m = 1
            x, y, sx = int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]
            nf = int(rng.integers(0, 2)) if m > 1 else 0
            mn = int(rng.integers(0, 2))
            mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
            for test in ("cor", "zf", "mi-g", "mi-g-sh"):
                rb.reset_test_counter()
                pv, pmin, pmax, ds_ = ra.allsubs_test(data, gnames[x], gnames[y], [gnames[i] for i in sx],
                                                      [gnames[i] for i in sx[:nf]], test, 0.05, mn, mx)
                rec["allsubs"].append({"test": test, "x": x, "y": y, "sx": sx, "nfixed": nf, "min": mn, "max": mx,
                                       "alpha": 0.05, "p": pv, "pmin": pmin, "pmax": pmax,
                                       "dsep": None if ds_ is None else [gnames.index(d) for d in ds_],
                                       "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
                rb.reset_test_counter()
                pvs = ra.roundrobin_test(data, gnames[x], [gnames[i] for i in [y] + sx],
                                         [gnames[i] for i in sx[:nf]], test, 0.05)
                rec["roundrobin"].append({"test": test, "x": x, "z": [y] + sx, "nfixed": nf, "alpha": 0.05,
                                          "p": [float(v) for v in pvs], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
            rb.reset()
            data.invalidate()
        out[key] = rec
    with open(os.path.join(HERE, "ref_incomplete_gaussian_patterns.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_incomplete_gaussian_patterns.json")


def permutation_outputs():
    """Permutation, sequential and semiparametric tests from the reference's
    c_mcarlo / c_cmcarlo (and the Gaussian c_gauss_mcarlo / c_gauss_cmcarlo) with the
    shim's splitmix64 stream: every batch is seeded once and its tests then share
    the stream, as the tests of one R session do. Own file."""
    from oracle import refapi as ra, refbind as rb
    out = {}
    for ds in ("learning_test", "learning_test_na"):
        names, codes, nlev = load_discrete(ds)
        z = np.load(os.path.join(HERE, "data_%s.npz" % ds))
        codes[z["codes"] == 255] = np.iinfo(np.int32).min      # missing cells
        codes = [np.ascontiguousarray(c) for c in codes]
        units = [(0, 1, []), (0, 5, []), (2, 5, []), (3, 4, []), (0, 1, [2]), (1, 4, [0, 2]), (5, 2, [3]),
                 (0, 4, [1, 2, 3])]
        batches = []
        for test, B, alpha, seed in (("mc-mi", 150, 1.0, 11), ("mc-x2", 150, 1.0, 12), ("mc-jt", 150, 1.0, 13),
                                     ("sp-mi", 100, 1.0, 14), ("sp-x2", 100, 1.0, 15), ("smc-mi", 400, 0.05, 16),
                                     ("smc-x2", 400, 0.05, 17), ("smc-jt", 400, 0.01, 18)):
            res = []
            for k, (x, y, z) in enumerate(units):
                st, df, p = ra.mc_discrete(test, codes[x], int(nlev[x]), codes[y], int(nlev[y]),
                                           [codes[i] for i in z], [int(nlev[i]) for i in z], B=B, alpha=alpha,
                                           seed=seed if k == 0 else None)
                res.append({"x": x, "y": y, "z": z, "stat": st, "df": None if np.isnan(df) else df, "p": p})
            batches.append({"test": test, "B": B, "alpha": alpha, "seed": seed, "results": res,
                            "end_state": str(rb.lib().minir_rng_peek())})      # where the stream stands afterwards
        out[ds] = batches
    names, vals = load_gaussian("gaussian_test")
    vals = [np.ascontiguousarray(v) for v in vals]
    units = [(0, 1, []), (0, 2, []), (3, 5, []), (0, 1, [2]), (1, 4, [0, 2]), (5, 6, [0, 1, 2, 3])]
    batches = []
    for test, B, alpha, seed in (("mc-cor", 120, 1.0, 21), ("mc-zf", 120, 1.0, 22), ("mc-mi-g", 120, 1.0, 23),
                                 ("smc-cor", 300, 0.05, 24), ("smc-zf", 300, 0.05, 25), ("smc-mi-g", 300, 0.02, 26)):
        res = []
        for k, (x, y, z) in enumerate(units):
            st, p = ra.mc_gaussian(test, [vals[x], vals[y]] + [vals[i] for i in z], B=B, alpha=alpha,
                                   seed=seed if k == 0 else None)
            res.append({"x": x, "y": y, "z": z, "stat": st, "p": p})
        batches.append({"test": test, "B": B, "alpha": alpha, "seed": seed, "results": res,
                        "end_state": str(rb.lib().minir_rng_peek())})
    out["gaussian_test"] = batches
    # the call sites with permutation tests (ut/ct_dperm, ast_dperm / ast_gperm, rrd_dperm /
    # rrd_gperm): one stream per record list, carried from call to call
    L = rb.lib()
    cs = {}
    names, codes, nlev = load_discrete("learning_test")
    data = ra.RefData([np.ascontiguousarray(c) for c in codes], [int(v) for v in nlev], names)
    rng = np.random.default_rng(20261116)
    recs = []
    L.minir_rng_seed(31)
    p = len(names)
    for t in range(10):
        m = int(rng.integers(2, p - 1))
        idx = [int(i) for i in rng.choice(p, size=2 + m, replace=False)]
        x, y, sx = idx[0], idx[1], idx[2:]
        nf = int(rng.integers(0, 2))
        mn = int(rng.integers(0, 2))
        mx = int(rng.integers(max(mn, 1), min(3, m - nf) + 1))
        test = ("mc-mi", "smc-x2", "sp-mi", "mc-jt", "smc-mi")[t % 5]
        B = 60
        rb.reset_test_counter()
        pv = ra.indep_test(data, [names[x], names[sx[0]]], names[y], [names[i] for i in sx[1:2]], test, 0.05, B=B)
        recs.append({"site": "indep", "test": test, "B": B, "x": [x, sx[0]], "y": y, "sx": sx[1:2], "alpha": 0.05,
                     "p": [float(v) for v in pv], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset_test_counter()
        pv, pmin, pmax, ds_ = ra.allsubs_test(data, names[x], names[y], [names[i] for i in sx],
                                              [names[i] for i in sx[:nf]], test, 0.05, mn, mx, B=B)
        recs.append({"site": "allsubs", "test": test, "B": B, "x": x, "y": y, "sx": sx, "nfixed": nf, "min": mn,
                     "max": mx, "alpha": 0.05, "p": pv, "pmin": pmin, "pmax": pmax,
                     "dsep": None if ds_ is None else [names.index(d) for d in ds_], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset_test_counter()
        pvs = ra.roundrobin_test(data, names[x], [names[i] for i in [y] + sx], [names[i] for i in sx[:nf]],
                                 test, 0.05, B=B)
        recs.append({"site": "roundrobin", "test": test, "B": B, "x": x, "z": [y] + sx, "nfixed": nf, "alpha": 0.05,
                     "p": [float(v) for v in pvs], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset()
        data.invalidate()
    cs["learning_test"] = {"seed": 31, "records": recs, "end_state": str(L.minir_rng_peek())}
    gnames, vals = load_gaussian("gaussian_test")
    gdata = ra.RefData([np.ascontiguousarray(v[:1500]) for v in vals], None, gnames)   # first 1500 rows
    recs = []
    L.minir_rng_seed(32)
    gp = len(gnames)
    for t in range(6):
        m = int(rng.integers(2, 4))
        idx = [int(i) for i in rng.choice(gp, size=2 + m, replace=False)]
        x, y, sx = idx[0], idx[1], idx[2:]
        test = ("mc-cor", "smc-zf", "mc-mi-g")[t % 3]
        B = 40
        rb.reset_test_counter()
        pv, pmin, pmax, ds_ = ra.allsubs_test(gdata, gnames[x], gnames[y], [gnames[i] for i in sx], [], test, 0.05,
                                              0, 2, B=B)
        recs.append({"site": "allsubs", "test": test, "B": B, "x": x, "y": y, "sx": sx, "nfixed": 0, "min": 0,
                     "max": 2, "alpha": 0.05, "p": pv, "pmin": pmin, "pmax": pmax,
                     "dsep": None if ds_ is None else [gnames.index(d) for d in ds_], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset_test_counter()
        pvs = ra.roundrobin_test(gdata, gnames[x], [gnames[i] for i in [y] + sx], [], test, 0.05, B=B)
        recs.append({"site": "roundrobin", "test": test, "B": B, "x": x, "z": [y] + sx, "nfixed": 0, "alpha": 0.05,
                     "p": [float(v) for v in pvs], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset()
        gdata.invalidate()
    cs["gaussian_test_1500"] = {"seed": 32, "records": recs, "end_state": str(L.minir_rng_peek())}
    # incomplete numeric data: the *_gperm patterns test on the rows complete for the test's columns
    zna = np.load(os.path.join(HERE, "data_gaussian_test_na.npz"))
    nnames, nvals = [str(v) for v in zna["names"]], zna["values"]
    ndata = ra.RefData([np.ascontiguousarray(v) for v in nvals], None, nnames)
    recs = []
    L.minir_rng_seed(33)
    for t in range(6):
        m = int(rng.integers(2, 4))
        idx = [int(i) for i in rng.choice(len(nnames), size=2 + m, replace=False)]
        x, y, sx = idx[0], idx[1], idx[2:]
        test = ("mc-cor", "smc-zf", "mc-mi-g")[t % 3]
        B = 30
        rb.reset_test_counter()
        pv = ra.indep_test(ndata, [nnames[x], nnames[sx[0]]], nnames[y], [nnames[i] for i in sx[1:2]], test, 0.05, B=B)
        recs.append({"site": "indep", "test": test, "B": B, "x": [x, sx[0]], "y": y, "sx": sx[1:2], "alpha": 0.05,
                     "p": [float(v) for v in pv], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset_test_counter()
        pv, pmin, pmax, ds_ = ra.allsubs_test(ndata, nnames[x], nnames[y], [nnames[i] for i in sx], [], test, 0.05,
                                              0, 2, B=B)
        recs.append({"site": "allsubs", "test": test, "B": B, "x": x, "y": y, "sx": sx, "nfixed": 0, "min": 0,
                     "max": 2, "alpha": 0.05, "p": pv, "pmin": pmin, "pmax": pmax,
                     "dsep": None if ds_ is None else [nnames.index(d) for d in ds_], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset_test_counter()
        pvs = ra.roundrobin_test(ndata, nnames[x], [nnames[i] for i in [y] + sx], [], test, 0.05, B=B)
        recs.append({"site": "roundrobin", "test": test, "B": B, "x": x, "z": [y] + sx, "nfixed": 0, "alpha": 0.05,
                     "p": [float(v) for v in pvs], "counter": rb.test_counter(), "state": str(rb.lib().minir_rng_peek())})
        rb.reset()
        ndata.invalidate()
    cs["gaussian_test_na"] = {"seed": 33, "records": recs, "end_state": str(L.minir_rng_peek())}
    out["callsites"] = cs
    with open(os.path.join(HERE, "ref_permutation.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_permutation.json")


def mbde_outputs():
    """mBDe node scores from the reference's per_node_score (dirichlet_node with the
    `exp` list, src/scores/dirichlet.posterior.c) on learning.test with made-up
    interventions on three nodes. Own file and RNG."""
    from oracle import refapi as ra, refbind as rb
    names, codes, nlev = load_discrete("learning_test")
    codes = [np.ascontiguousarray(c) for c in codes]
    n = len(codes[0])
    rng = np.random.default_rng(20261110)
    exp = {0: np.arange(1, n + 1, 7), 3: np.sort(rng.choice(n, 300, replace=False)) + 1,
           5: np.sort(rng.choice(n, n - 40, replace=False)) + 1}          # almost every row of F
    data = ra.RefData(codes, [int(v) for v in nlev], names)
    recs = []
    for t in (0, 3, 5):
        others = [i for i in range(len(names)) if i != t]
        for pa in ([], others[:1], others[1:3], others[:4], others[::-1]):
            par = {nm: [] for nm in names}
            par[names[t]] = [names[i] for i in pa]
            for iss in (1.0, 10.0):
                v = ra.per_node_score(data, par, [names[t]], "mbde", iss=iss,
                                      exp={names[k]: r for k, r in exp.items()})[0]
                recs.append({"node": t, "parents": pa, "iss": iss, "value": float(v)})
            rb.reset()
            data.invalidate()
    with open(os.path.join(HERE, "ref_mbde.json"), "w") as f:
        json.dump({"learning_test": {"exp": {str(k): [int(i) for i in r] for k, r in exp.items()}, "scores": recs}}, f)
    print("wrote ref_mbde.json")


def special_function_fixtures():
    import mpmath as mp
    mp.mp.dps = 50
    rec = {"lgamma": [], "pchisq_upper": [], "pt_upper2": [], "pnorm_upper2": []}
    for x in [1e-3, 0.01, 0.05, 0.1, 1 / 3, 0.5, 0.9, 1.0, 1.5, 2.0, 2.5, 3.7, 9.99, 10.0, 10.5,
              17.25, 100.125, 1234.5, 20000.0625, 1e6 + 1 / 81, 1e7 + 0.5, 3.3e8]:
        rec["lgamma"].append([x, float(mp.loggamma(mp.mpf(x)))])
    for df in [1, 2, 4, 9, 12, 36, 108, 243, 1024, 9216]:
        for mult in [0.0, 0.1, 0.5, 0.9, 1.0, 1.1, 1.5, 2.0, 4.0, 10.0]:
            x = df * mult
            if x == 0 and df > 0:
                v = 1.0
            else:
                v = float(mp.gammainc(mp.mpf(df) / 2, mp.mpf(x) / 2, mp.inf, regularized=True))
            if v > 1e-300:
                rec["pchisq_upper"].append([x, df, v])
    for df in [1, 2, 3, 10, 30, 100, 4997, 19998, 999998, 9999997]:
        for t in [0.0, 0.01, 0.5, 1.0, 1.96, 2.5, 5.0, 10.0, 30.0]:
            x = mp.mpf(df) / (mp.mpf(df) + mp.mpf(t) ** 2)
            v = float(mp.betainc(mp.mpf(df) / 2, mp.mpf(1) / 2, 0, x, regularized=True))
            if v > 1e-300:
                rec["pt_upper2"].append([t, df, v])
    for z in [0.0, 0.1, 0.5, 1.0, 1.959963984540054, 3.0, 5.0, 8.0, 15.0, 30.0]:
        rec["pnorm_upper2"].append([z, float(mp.erfc(mp.mpf(z) / mp.sqrt(2)))])
    with open(os.path.join(HERE, "special_functions.json"), "w") as f:
        json.dump(rec, f)
    print("wrote special_functions.json")


if __name__ == "__main__":
    export_datasets()
    special_function_fixtures()
    reference_outputs()
    incomplete_outputs()
    jt_shrinkage_outputs()
    nml_outputs()
    cg_outputs()
    cg_nal_outputs()
    ebic_outputs()
    bge_outputs()
    mi_matrix_outputs()
    incomplete_gaussian_patterns()
    permutation_outputs()
    mbde_outputs()
