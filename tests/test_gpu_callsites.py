"""GPU parity of the call-site layer (early exit, min/max p-values, separating
sets, score-cache deltas and the test counter) against outputs of the
reference's own allsubs_test / roundrobin_test / score_cache_fill."""
import numpy as np
import pytest

DISCRETE_SETS = ["learning_test", "alarm", "learning_test_na", "learning_test_jt_sh", "alarm_jt_sh",
                 "learning_test_nml", "alarm_nml", "alarm_600_nml"]
from util import load_discrete, load_gaussian, rel_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from bnlearn_b200 import Engine
    with Engine() as e:
        yield e


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_allsubs_against_reference(engine, golden, ds):
    from bnlearn_b200 import callsites as cs
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    d = engine.register(cols, nlev)
    counter = cs.TestCounter()
    for r in golden[ds]["allsubs"]:
        counter.reset()
        nf = r["nfixed"]
        p, pmin, pmax, dsep = cs.allsubs_test(d, r["x"], r["y"], r["sx"][nf:], r["sx"][:nf], r["test"],
                                              r["alpha"], r["min"], r["max"], counter)
        assert counter.value == r["counter"], r
        assert dsep == r["dsep"], (r, dsep)
        for got, ref in ((p, r["p"]), (pmin, r["pmin"]), (pmax, r["pmax"])):
            assert rel_close(got, ref, 1e-9, 1e-300), (r, got, ref)
    # the same requests in one batched call give the same answers and total count.
    recs = [r for r in golden[ds]["allsubs"] if r["test"] == "mi" and r["min"] == 0 and r["max"] == 2]
    if recs:
        counter.reset()
        res = cs.allsubs_test_batch(d, [r["x"] for r in recs], [r["y"] for r in recs],
                                    [r["sx"] for r in recs], [r["nfixed"] for r in recs], "mi", 0.05, 0, 2,
                                    counter)
        assert counter.value == sum(r["counter"] for r in recs)
        for got, r in zip(res, recs):
            assert got[3] == r["dsep"] and rel_close(got[0], r["p"], 1e-9, 1e-300)
    d.release()


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_roundrobin_against_reference(engine, golden, ds):
    from bnlearn_b200 import callsites as cs
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    d = engine.register(cols, nlev)
    counter = cs.TestCounter()
    for r in golden[ds]["roundrobin"]:
        counter.reset()
        z = r["z"]
        fixed = z[1:1 + r["nfixed"]]
        pv = cs.roundrobin_test(d, r["x"], z, fixed, r["test"], r["alpha"], counter)
        assert counter.value == r["counter"]
        assert len(pv) == len(r["p"])
        for got, ref in zip(pv, r["p"]):
            assert rel_close(got, ref, 1e-9, 1e-300), (r, list(pv))
    d.release()


def test_score_cache_fill_against_reference(engine, golden):
    from bnlearn_b200 import callsites as cs
    names, cols, nlev = load_discrete("learning_test")
    d = engine.register(cols, nlev)
    n = len(names)
    counter = cs.TestCounter()
    for r in golden["score_cache_fill"]:
        amat = np.zeros((n, n), dtype=np.int32)
        for child, pas in r["parents"].items():
            for p in pas:
                amat[names.index(p), names.index(child)] = 1
        parents = [[i for i in range(n) if amat[i, j]] for j in range(n)]
        ref = cs.per_node_score(d, list(range(n)), parents, r["score"], r["arg"])
        for got, want in zip(ref, r["reference"]):
            assert rel_close(got, want, 1e-9)
        cache = np.zeros((n, n))
        counter.reset()
        cs.score_cache_fill(d, r["score"], r["arg"], amat, np.zeros((n, n), np.int32), list(range(n)), True,
                            np.array(r["reference"]), cache, counter)
        assert counter.value == r["counter"]
        want = np.array(r["cache"])
        for i in range(n):
            for j in range(n):
                assert rel_close(cache[i, j], want[i, j], 1e-9, 1e-7), (r["score"], i, j, cache[i, j], want[i, j])
    d.release()


def test_indep_test_counts_and_matches_batch(engine):
    from bnlearn_b200 import callsites as cs
    names, cols, nlev = load_discrete("alarm")
    d = engine.register(cols, nlev)
    counter = cs.TestCounter()
    pv = cs.indep_test(d, [0, 1, 2, 3], 5, [7, 9], "mi", counter)
    assert counter.value == 4
    _, _, ref = d.ci_discrete("mi", [0, 1, 2, 3], [5] * 4, [[7, 9]] * 4)
    assert np.array_equal(pv, ref)
    st, df, p = cs.ci_test(d, 3, 5, [], "x2", counter)
    assert counter.value == 5 and df == (nlev[3] - 1) * (nlev[5] - 1)
    d.release()


def test_gaussian_allsubs_and_roundrobin_run(engine):
    """Gaussian call sites: checked against direct evaluation of every subset."""
    from bnlearn_b200 import callsites as cs
    names, cols = load_gaussian("gaussian_test")
    d = engine.register(cols)
    counter = cs.TestCounter()
    p, pmin, pmax, dsep = cs.allsubs_test(d, 0, 5, [1, 2, 3], [], "cor", 0.05, 0, 3, counter)
    subsets = [[], [1], [2], [3], [1, 2], [1, 3], [2, 3], [1, 2, 3]]
    _, _, pv = d.ci_gaussian("cor", [0] * 8, [5] * 8, subsets)
    first = next((i for i, v in enumerate(pv) if v > 0.05), None)
    if first is None:
        assert dsep is None and counter.value == 9 and p == pv[-1]
        assert pmin == pv.min() and pmax == pv.max()
    else:
        assert dsep == subsets[first] and counter.value == first + 2
    counter.reset()
    out = cs.roundrobin_test(d, 0, [1, 2, 3, 4], [], "zf", 0.05, counter)
    assert counter.value == 4 and len(out) == 4
    d.release()


@pytest.mark.parametrize("key", ["all_incomplete", "mixed", "tiny"])
def test_gaussian_subset_patterns_on_incomplete_data(engine, key):
    """allsubs / roundrobin with cor, zf, mi-g, mi-g-sh when some columns hold missing
    values: the reference's *_with_missing patterns (every subset on its own complete
    rows, no early exit, no separating set; a subset without degrees of freedom ends
    the request) against tests/golden/ref_incomplete_gaussian_patterns.json. In
    "mixed" the requests on complete columns alone keep the complete-data pattern."""
    import json
    import os
    from bnlearn_b200 import callsites as cs
    from util import GOLDEN
    with open(os.path.join(GOLDEN, "ref_incomplete_gaussian_patterns.json")) as f:
        rec = json.load(f)[key]
    z = np.load(os.path.join(GOLDEN, "data_gaussian_test_na.npz"))
    vals = z["values"].copy()
    if key != "all_incomplete":
        _, full = load_gaussian("gaussian_test")
        for c in range(3):
            vals[c] = full[c][:vals.shape[1]]
    if key == "tiny":
        vals = vals[:, :9]
    d = engine.register([np.ascontiguousarray(v) for v in vals])
    counter = cs.TestCounter()
    for r in rec["allsubs"]:
        counter.reset()
        nf = r["nfixed"]
        p, pmin, pmax, dsep = cs.allsubs_test(d, r["x"], r["y"], r["sx"][nf:], r["sx"][:nf], r["test"],
                                              r["alpha"], r["min"], r["max"], counter)
        assert counter.value == r["counter"], (r, counter.value)
        assert dsep == r["dsep"], (r, dsep)
        tol = 1e-6 if r["test"] == "mi-g-sh" else 1e-9
        for got, ref in ((p, r["p"]), (pmin, r["pmin"]), (pmax, r["pmax"])):
            assert rel_close(got, ref, tol, 1e-300), (r, got, ref)
    for r in rec["roundrobin"]:
        counter.reset()
        zz = r["z"]
        pv = cs.roundrobin_test(d, r["x"], zz, zz[1:1 + r["nfixed"]], r["test"], r["alpha"], counter)
        assert counter.value == r["counter"]
        assert len(pv) == len(r["p"])
        tol = 1e-6 if r["test"] == "mi-g-sh" else 1e-9
        # Up to and including the first variable that is dropped the reference is the
        # yardstick. After a drop, rrd_gaustests_with_missing refills cov.mat for one
        # column fewer but leaves cov.dim as it was (roundrobin.test.c:324-325, where the
        # complete-data pattern calls covariance_drop_variable, :202), so c_fast_pcor
        # reads the new matrix with the old leading dimension: its later p-values are
        # not the tests' (1.0, 0.0, ... in the fixture). From there on the sequence is
        # checked against direct evaluation of the intended tests.
        first_drop = next((i for i, v in enumerate(r["p"]) if v > r["alpha"]), len(r["p"]))
        for got, ref in zip(pv[:first_drop + 1], r["p"][:first_drop + 1]):
            assert rel_close(got, ref, tol, 1e-300), (r, list(pv))
        live, cur, nfx = list(zz), 0, r["nfixed"]
        for i, zi in enumerate(zz):
            if len(live) < 2:
                break
            if 1 <= i <= nfx:
                continue
            _, df1, p1 = d.ci_gaussian(r["test"], [r["x"]], [zi], [[v for v in live if v != zi]])
            expect = 1.0 if df1[0] == 0.0 else p1[0]
            assert rel_close(pv[cur], expect, 1e-12, 1e-300), (r, cur, list(pv))
            if df1[0] != 0.0 and p1[0] > r["alpha"]:
                live.remove(zi)
            cur += 1
    d.release()


def test_mi_cg_subset_patterns(engine):
    """allsubs / roundrobin with mi-cg on mixed data, complete (ast_micg_complete /
    rrd_micg_complete) and with missing cells (ast_micg_with_missing /
    rrd_micg_with_missing: every test on the rows complete for its own columns),
    early exit as usual."""
    import json
    import os
    from bnlearn_b200 import callsites as cs
    from util import GOLDEN, load_mixed
    with open(os.path.join(GOLDEN, "ref_cg_nal.json")) as f:
        fx = json.load(f)["clgaussian_test_na"]
    names, cols, nlev = load_mixed("clgaussian_test")
    rng = np.random.default_rng(fx["seed"])
    na = []
    for c, l in zip(cols, nlev):
        c = c.copy()
        m = rng.random(len(c)) < fx["frac"]
        c[m] = np.iinfo(np.int32).min if l > 0 else np.nan
        na.append(c)
    with open(os.path.join(GOLDEN, "ref_cg.json")) as f:
        complete = json.load(f)["clgaussian_test"]
    for data, fxs in ((cols, complete), (na, fx)):
        _mi_cg_patterns(engine, cs, data, nlev, fxs)


def _mi_cg_patterns(engine, cs, data, nlev, fx):
    d = engine.register(data, nlev)
    counter = cs.TestCounter()
    for r in fx["allsubs"]:
        counter.reset()
        nf = r["nfixed"]
        p, pmin, pmax, dsep = cs.allsubs_test(d, r["x"], r["y"], r["sx"][nf:], r["sx"][:nf], "mi-cg", r["alpha"],
                                              r["min"], r["max"], counter)
        assert counter.value == r["counter"], (r, counter.value)
        assert dsep == r["dsep"], (r, dsep)
        for got, ref in ((p, r["p"]), (pmin, r["pmin"]), (pmax, r["pmax"])):
            assert rel_close(got, ref, 1e-8, 1e-300), (r, got, ref)
    for r in fx["roundrobin"]:
        counter.reset()
        zz = r["z"]
        pv = cs.roundrobin_test(d, r["x"], zz, zz[1:1 + r["nfixed"]], "mi-cg", r["alpha"], counter)
        assert counter.value == r["counter"]
        for got, ref in zip(pv, r["p"]):
            assert rel_close(got, ref, 1e-8, 1e-300), (r, list(pv))
    d.release()


def test_indep_test_with_mi_cg(engine):
    """mi-cg through the indep_test call site: p-values of the batched entry point,
    counter advanced by one per test."""
    from bnlearn_b200 import callsites as cs
    from util import load_mixed
    names, cols, nlev = load_mixed("clgaussian_test")
    d = engine.register(cols, nlev)
    counter = cs.TestCounter()
    pv = cs.indep_test(d, [0, 3, 4], 6, [1, 7], "mi-cg", counter)
    _, _, ref = d.ci_cg([0, 3, 4], [6, 6, 6], [[1, 7]] * 3)
    assert counter.value == 3 and np.array_equal(pv, ref)
    st, df, p = cs.ci_test(d, 3, 0, [], "mi-cg", counter)
    assert counter.value == 4 and df == 2.0 and 0.0 <= p <= 1.0 and st > 0
    d.release()
