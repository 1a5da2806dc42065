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
