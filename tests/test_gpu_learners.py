"""End to end: the restated search loops (bnlearn_b200/learners.py) driven by the
CUDA backend and by the reference's own C entry points must produce identical
graphs, separating sets and test-counter totals (north_star: "identical learned
graphs and test-counter totals")."""
import numpy as np
import pytest

from util import load_discrete, load_gaussian

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from bnlearn_b200 import Engine
    with Engine() as e:
        yield e


def _backends(engine, cols, nlev=None):
    from bnlearn_b200.learners import GpuBackend
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libbnref.so is not built (needs the reference sources at build time)")
    from refbackend import RefBackend
    ds = engine.register(cols, nlev)
    return GpuBackend(ds), RefBackend(cols, nlev), ds


@pytest.mark.parametrize("ds,score,arg", [("learning_test", "bic", None), ("learning_test", "bde", 1.0),
                                          ("alarm", "bic", None), ("alarm", "bde", 1.0)])
def test_hc_identical_graph_and_counter(engine, ds, score, arg):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete(ds)
    n = len(cols[0])
    if ds == "alarm":
        cols = [c[:5000] for c in cols]      # keeps the reference-backend run short
        n = 5000
    arg = np.log(n) / 2 if arg is None else arg
    gpu, ref, d = _backends(engine, cols, nlev)
    a1, s1 = learners.hc(gpu, len(cols), score, arg)
    a2, s2 = learners.hc(ref, len(cols), score, arg)
    assert np.array_equal(a1, a2)
    assert gpu.counter.value == ref.counter.value
    assert np.allclose(s1, s2, rtol=1e-9, atol=0)
    d.release()


def test_tabu_identical_graph_and_counter(engine):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete("learning_test")
    gpu, ref, d = _backends(engine, cols, nlev)
    a1, _ = learners.tabu(gpu, len(cols), "bde", 1.0, tabu=5)
    a2, _ = learners.tabu(ref, len(cols), "bde", 1.0, tabu=5)
    assert np.array_equal(a1, a2)
    assert gpu.counter.value == ref.counter.value
    d.release()


def test_hc_bic_on_alarm_recovers_config_a(engine):
    """Config A (hc + BIC on the bundled alarm data, all 20000 rows) runs on the
    GPU path; the score of the learned DAG is reproduced by the oracle."""
    from bnlearn_b200 import learners
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("alarm")
    n = len(cols[0])
    ds = engine.register(cols, nlev)
    be = learners.GpuBackend(ds)
    amat, scores = learners.hc(be, len(cols), "bic", np.log(n) / 2)
    assert 40 <= amat.sum() <= 60 and be.counter.value > 1000
    total = 0.0
    for j in range(len(cols)):
        pa = [int(i) for i in np.flatnonzero(amat[:, j])]
        total += orc.score_discrete("bic", cols[j], nlev[j], [cols[i] for i in pa], [nlev[i] for i in pa],
                                    np.log(n) / 2)[0]
    assert abs(total - scores.sum()) <= 1e-9 * abs(total)
    ds.release()


@pytest.mark.parametrize("ds,test", [("learning_test", "mi"), ("learning_test", "x2"), ("alarm", "mi")])
def test_pc_stable_identical_skeleton_dsep_counter(engine, ds, test):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete(ds)
    if ds == "alarm":
        cols = [c[:3000] for c in cols]
    gpu, ref, d = _backends(engine, cols, nlev)
    k1, d1, p1 = learners.pc_stable(gpu, len(cols), test, 0.05)
    k2, d2, p2 = learners.pc_stable(ref, len(cols), test, 0.05)
    assert k1 == k2 and d1 == d2
    assert gpu.counter.value == ref.counter.value
    for key in p1:
        assert np.isclose(p1[key], p2[key], rtol=1e-9, atol=1e-300)
    v1 = learners.vstructures(gpu, len(cols), None, None, test, 0.05, d1, p1) if False else None
    d.release()


@pytest.mark.parametrize("test,B", [("mc-mi", 40), ("smc-x2", 60), ("sp-mi", 30)])
def test_pc_stable_with_permutation_tests(engine, test, B):
    """pc.stable with a permutation test: one uniform stream through the whole run,
    pair by pair as pc.heuristic visits them (R/pcalgo.R:69-148). Same skeleton,
    separating sets, p-values, counter -- and the stream ends at the same position as
    the reference's, i.e. exactly the same draws were consumed."""
    from bnlearn_b200 import callsites as cs, learners
    from bnlearn_b200.learners import GpuBackend
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libbnref.so is not built")
    from refbackend import RefBackend
    names, cols, nlev = load_discrete("learning_test")
    d = engine.register(cols, nlev)
    stream = cs.UniformStream(engine._lib, B, seed=77)
    gpu = GpuBackend(d, stream)
    k1, d1, p1 = learners.pc_stable(gpu, len(cols), test, 0.05)
    ref = RefBackend(cols, nlev, B=B, seed=77)
    k2, d2, p2 = learners.pc_stable(ref, len(cols), test, 0.05)
    assert k1 == k2 and d1 == d2
    assert gpu.counter.value == ref.counter.value
    for key in p1:
        assert np.isclose(p1[key], p2[key], rtol=1e-9, atol=1e-300), (key, p1[key], p2[key])
    assert stream.position == refbind.lib().minir_rng_peek()
    d.release()


@pytest.mark.parametrize("algo,test,B", [("mmpc", "mc-mi", 30), ("si.hiton.pc", "smc-mi", 50), ("iamb", "mc-x2", 30)])
def test_local_discovery_with_permutation_tests(engine, algo, test, B):
    """mmpc / si.hiton.pc / iamb on permutation tests: indep_test, allsubs and
    roundrobin calls interleave on one stream; neighbourhoods, Markov blankets,
    counter and final stream position as with the reference's call sites."""
    from bnlearn_b200 import callsites as cs, learners
    from bnlearn_b200.learners import GpuBackend
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libbnref.so is not built")
    from refbackend import RefBackend
    names, cols, nlev = load_discrete("learning_test")
    cols = [c[:2000] for c in cols]
    d = engine.register(cols, nlev)
    stream = cs.UniformStream(engine._lib, B, seed=91)
    gpu = GpuBackend(d, stream)
    n1, m1 = learners.local_discovery(gpu, len(cols), algo, test, 0.05)
    ref = RefBackend(cols, nlev, B=B, seed=91)
    n2, m2 = learners.local_discovery(ref, len(cols), algo, test, 0.05)
    assert n1 == n2 and m1 == m2
    assert gpu.counter.value == ref.counter.value
    assert stream.position == refbind.lib().minir_rng_peek()
    d.release()


@pytest.mark.parametrize("algo", ["mmpc", "si.hiton.pc", "iamb"])
def test_local_discovery_identical(engine, algo):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete("learning_test")
    gpu, ref, d = _backends(engine, cols, nlev)
    n1, m1 = learners.local_discovery(gpu, len(cols), algo, "mi", 0.05)
    n2, m2 = learners.local_discovery(ref, len(cols), algo, "mi", 0.05)
    assert n1 == n2 and m1 == m2
    v1 = learners.vstructures(gpu, len(cols), n1, m1, "mi", 0.05)
    v2 = learners.vstructures(ref, len(cols), n2, m2, "mi", 0.05)
    assert [v[1:] for v in v1] == [v[1:] for v in v2]
    assert gpu.counter.value == ref.counter.value
    d.release()


def test_pc_stable_gaussian_identical(engine):
    from bnlearn_b200 import learners
    names, cols = load_gaussian("gaussian_test")
    gpu, ref, d = _backends(engine, cols)
    for test in ("cor", "zf"):
        gpu.counter.reset(); ref.counter.reset()
        k1, d1, _ = learners.pc_stable(gpu, len(cols), test, 0.05)
        k2, d2, _ = learners.pc_stable(ref, len(cols), test, 0.05)
        assert k1 == k2 and d1 == d2
        assert gpu.counter.value == ref.counter.value
    a1, _ = learners.hc(gpu, len(cols), "bic-g", np.log(len(cols[0])) / 2)
    a2, _ = learners.hc(ref, len(cols), "bic-g", np.log(len(cols[0])) / 2)
    assert np.array_equal(a1, a2)
    d.release()


def test_boot_strength_runs_and_is_deterministic(engine):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete("learning_test")
    s1, d1, t1 = learners.boot_strength(engine, cols, nlev, 8, "bde", 1.0, seed=20261109)
    s2, d2, t2 = learners.boot_strength(engine, cols, nlev, 8, "bde", 1.0, seed=20261109)
    assert np.array_equal(s1, s2) and np.array_equal(d1, d2) and t1 == t2
    assert s1[0, 1] == 1.0        # A - B is in every replicate of learning.test


@pytest.mark.parametrize("ds", ["learning_test", "alarm", "insurance", "gaussian_test"])
def test_aracne_and_chow_liu_against_reference_golden(engine, ds):
    """Committed outputs of the reference's aracne() / chow_liu() (tests/golden/
    ref_mi_matrix.json); the coefficient matrix comes from bnl_mi_matrix."""
    import json
    import os
    from bnlearn_b200 import learners
    from util import GOLDEN
    with open(os.path.join(GOLDEN, "ref_mi_matrix.json")) as f:
        g = json.load(f)[ds]
    if ds == "gaussian_test":
        names, cols = load_gaussian(ds)
        d, est = engine.register(cols), "mi-g"
    else:
        names, cols, nlev = load_discrete(ds)
        d, est = engine.register(cols, nlev), "mi"
    p = len(names)
    assert learners.aracne(d, p, est) == [tuple(a) for a in g["aracne"]]
    tree = learners.chow_liu(d, p, est)
    assert tree == [tuple(a) for a in g["chow_liu"]]
    directed = learners.tree_directions(tree, p, 0)
    assert len(directed) == p - 1 and len({b for _, b in directed}) == p - 1 and all(b != 0 for _, b in directed)
    if "chow_liu_given_first" in g:
        got = learners.chow_liu(d, p, est, cols=list(range(1, p)), cond=0)
        assert got == [tuple(a) for a in g["chow_liu_given_first"]]
    d.release()


def test_mi_matrix_against_oracle(engine):
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("alarm")
    d = engine.register(cols, nlev)
    sel = [3, 7, 12, 20, 33]
    m = d.mi_matrix("mi", sel)
    mc = d.mi_matrix("mi", sel, cond=5)
    n = len(cols[0])
    for a in range(len(sel)):
        for b in range(a + 1, len(sel)):
            i, j = sel[a], sel[b]
            st, _, _ = orc.ci_discrete("mi", cols[i], nlev[i], cols[j], nlev[j])
            assert abs(m[a, b] - st / (2 * n)) <= 1e-12 * max(1.0, abs(m[a, b])) and m[a, b] == m[b, a]
            st, _, _ = orc.ci_discrete("mi", cols[i], nlev[i], cols[j], nlev[j], [cols[5]], [nlev[5]])
            assert abs(mc[a, b] - st / (2 * n)) <= 1e-12 * max(1.0, abs(mc[a, b]))
    assert d.mi_matrix("mi", [4]).shape == (1, 1)
    d.release()


@pytest.mark.parametrize("ds,score,arg,maxp", [("learning_test", "bic", None, np.inf), ("alarm", "bde", 1.0, np.inf),
                                               ("alarm", "bic", None, 2), ("insurance", "k2", 0.0, 3),
                                               ("gaussian_test", "bic-g", None, np.inf),
                                               ("gaussian_test", "bge", 0.0, np.inf)])
def test_native_hc_matches_the_stepwise_loop(engine, ds, score, arg, maxp):
    """bnl_hc (the whole search inside the library) against the same search driven
    one .Call-equivalent at a time (learners.hc, itself held to the reference's
    backend above): same graph, node scores, iteration and delta counts."""
    from bnlearn_b200 import callsites as cs, learners
    if ds == "gaussian_test":
        names, cols = load_gaussian(ds)
        d = engine.register(cols)
    else:
        names, cols, nlev = load_discrete(ds)
        d = engine.register(cols, nlev)
    n = len(cols[0])
    arg = np.log(n) / 2 if arg is None else arg
    be = learners.GpuBackend(d)
    a1, s1 = learners.hc(be, len(cols), score, arg, maxp=maxp)
    counter = cs.TestCounter()
    a2, s2, iters = d.hc(score, arg, maxp=maxp, counter=counter)
    assert np.array_equal(a1, a2)
    assert counter.value == be.counter.value
    assert np.array_equal(s1, s2)
    assert iters == a1.sum() + 1 or iters >= 1
    # start graph, blacklist, whitelist and max.iter
    p = len(cols)
    bl = np.zeros((p, p), dtype=np.int32)
    bl[0, 1] = bl[1, 0] = 1
    start = np.zeros((p, p), dtype=np.int32)
    start[2, 3] = 1
    wl = start.copy()
    a3, _, it3 = d.hc(score, arg, start=start, blmat=bl, wlmat=wl, maxp=maxp, max_iter=4)
    assert it3 <= 4 and a3[0, 1] == 0 and a3[1, 0] == 0 and (a3[2, 3] == 1 or a3[3, 2] == 1)
    d.release()


def test_hc_on_mixed_data_native_and_stepwise(engine):
    """hc with bic-cg on clgaussian.test: bnl_hc and the stepwise loop agree, no
    numeric node ends up as a parent of a factor (such arcs score -Inf), and the
    node scores are reproduced by the oracle."""
    from bnlearn_b200 import callsites as cs, learners
    from oracle import oracle as orc
    from util import load_mixed
    names, cols, nlev = load_mixed("clgaussian_test")
    n, p = len(cols[0]), len(cols)
    d = engine.register(cols, nlev)
    k = np.log(n) / 2
    be = learners.GpuBackend(d)
    a1, s1 = learners.hc(be, p, "bic-cg", k)
    counter = cs.TestCounter()
    a2, s2, _ = d.hc("bic-cg", k, counter=counter)
    assert np.array_equal(a1, a2) and counter.value == be.counter.value
    # the per-configuration moments are accumulated with double atomics: the last
    # bits of a score depend on the order of arrival
    assert np.allclose(s1, s2, rtol=1e-9, atol=0)
    assert a2.sum() >= 5
    for i in range(p):
        for j in range(p):
            if a2[i, j]:
                assert not (nlev[i] == 0 and nlev[j] > 0)
    for j in range(p):
        pa = [int(i) for i in np.flatnonzero(a2[:, j])]
        v = orc.score_cg("bic-cg", cols[j], nlev[j], [cols[i] for i in pa], [nlev[i] for i in pa], k)[0]
        assert abs(v - s2[j]) <= 1e-9 * abs(v)
    d.release()


def test_device_resample_equals_host_resample(engine):
    """bnl_data_resample (row gather on the device) gives the tables and the search
    result of registering the host-resampled data frame, with and without NA."""
    from bnlearn_b200 import learners
    for name in ("learning_test", "learning_test_na"):
        names, cols, nlev = load_discrete(name)
        rows = np.random.default_rng(5).integers(0, len(cols[0]), 4000)
        src = engine.register(cols, nlev)
        a = src.resample(rows)
        b = engine.register([c[rows] for c in cols], nlev)
        x, y, z = [0, 1, 2], [3, 4, 5], [[1], [], [0, 1]]
        for ta, tb in zip(a.counts(x, y, z), b.counts(x, y, z)):
            assert np.array_equal(ta, tb)
        score = "bic" if name == "learning_test" else "pnal"
        ra, sa, ia = a.hc(score, 4.0)
        rb_, sb, ib = b.hc(score, 4.0)
        assert np.array_equal(ra, rb_) and np.array_equal(sa, sb) and ia == ib
        with pytest.raises(Exception):
            src.resample([0, len(cols[0])])
        for d in (a, b, src):
            d.release()
    names, cols = load_gaussian("gaussian_test")
    rows = np.random.default_rng(6).integers(0, len(cols[0]), 3000)
    src = engine.register(cols)
    a, b = src.resample(rows), engine.register([c[rows] for c in cols])
    ma, mb = a.moments(), b.moments()
    assert np.array_equal(ma[0], mb[0]) and np.array_equal(ma[1], mb[1])
    for d in (a, b, src):
        d.release()


def test_boot_strength_native_equals_stepwise(engine):
    from bnlearn_b200 import learners
    names, cols, nlev = load_discrete("learning_test")
    s1 = learners.boot_strength(engine, cols, nlev, 6, "bde", 1.0, seed=11, native=True, workers=3)
    s3 = learners.boot_strength(engine, cols, nlev, 6, "bde", 1.0, seed=11, native=True, workers=1)
    assert np.array_equal(s1[0], s3[0]) and s1[2] == s3[2]
    import torch
    devs = list(range(torch.cuda.device_count()))
    s4 = learners.boot_strength(engine, cols, nlev, 6, "bde", 1.0, seed=11, workers=4, devices=devs)
    assert np.array_equal(s1[0], s4[0]) and s1[2] == s4[2]
    s2 = learners.boot_strength(engine, cols, nlev, 6, "bde", 1.0, seed=11, native=False)
    assert np.array_equal(s1[0], s2[0]) and np.array_equal(s1[1], s2[1]) and s1[2] == s2[2]


@pytest.mark.parametrize("ds,score,arg,tabu_len,max_iter", [("learning_test", "bde", 1.0, 5, np.inf),
                                                            ("learning_test", "bic", None, 3, np.inf),
                                                            ("alarm", "bic", None, 10, 40),
                                                            ("gaussian_test", "bic-g", None, 4, np.inf)])
def test_native_tabu_matches_the_stepwise_loop(engine, ds, score, arg, tabu_len, max_iter):
    """bnl_tabu against learners.tabu on the same engine (learners.tabu itself is
    held to the reference's backend above)."""
    from bnlearn_b200 import callsites as cs, learners
    if ds == "gaussian_test":
        names, cols = load_gaussian(ds)
        d = engine.register(cols)
    else:
        names, cols, nlev = load_discrete(ds)
        d = engine.register(cols, nlev)
    n = len(cols[0])
    arg = np.log(n) / 2 if arg is None else arg
    be = learners.GpuBackend(d)
    a1, s1 = learners.tabu(be, len(cols), score, arg, max_iter=max_iter, tabu=tabu_len)
    counter = cs.TestCounter()
    a2, s2, iters = d.tabu(score, arg, max_iter=0 if max_iter == np.inf else max_iter, tabu=tabu_len,
                           counter=counter)
    assert np.array_equal(a1, a2)
    assert counter.value == be.counter.value
    assert np.allclose(s1, s2, rtol=1e-12, atol=0)
    d.release()


@pytest.mark.parametrize("ds,test", [("learning_test", "mi"), ("alarm", "x2"), ("gaussian_test", "cor"),
                                     ("learning_test_na", "mi")])
def test_native_pc_stable_matches_the_level_loop(engine, ds, test):
    """bnl_pc_stable against learners.pc_stable on the same engine (itself held to the
    reference's backend above): skeleton, separating sets, p-values, test counter."""
    from bnlearn_b200 import callsites as cs, learners
    if ds == "gaussian_test":
        names, cols = load_gaussian(ds)
        d = engine.register(cols)
    else:
        names, cols, nlev = load_discrete(ds)
        if ds == "alarm":
            cols = [c[:5000] for c in cols]
        d = engine.register(cols, nlev)
    be = learners.GpuBackend(d)
    s1, d1, p1 = learners.pc_stable(be, len(cols), test, 0.05)
    counter = cs.TestCounter()
    s2, d2, p2 = d.pc_stable(test, 0.05, counter=counter)
    assert s1 == s2
    assert d1 == d2
    assert counter.value == be.counter.value
    for k in p1:
        assert (p1[k] is None and p2[k] == 0.0) or p1[k] == p2[k], (k, p1[k], p2[k])
    d.release()
