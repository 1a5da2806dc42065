"""GPU parity tests: the CUDA path, called through the C ABI, against (1) the
golden outputs of the reference's own C code and (2) the CPU oracle on seeded
synthetic inputs. Counts bit-exact; statistics, p-values and scores within 1e-9
relative (north_star)."""
import json
import os

import numpy as np
import pytest

DISCRETE_SETS = ["learning_test", "alarm", "learning_test_na", "learning_test_jt_sh", "alarm_jt_sh",
                 "learning_test_nml", "alarm_nml", "alarm_600_nml"]
from util import GOLDEN, load_discrete, load_gaussian, rel_close

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def engine():
    from bnlearn_b200 import Engine
    with Engine() as e:
        yield e


def _check(a, b, what, rtol=RTOL, atol=0.0):
    assert rel_close(a, b, rtol, atol), "%s: got %r, reference %r" % (what, a, b)


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_discrete_ci_against_reference_golden(engine, golden, ds):
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    d = engine.register(cols, nlev)
    recs = golden[ds]["tests"]
    for test in sorted({r["test"] for r in recs}):
        rs = [r for r in recs if r["test"] == test]
        st, df, pv = d.ci_discrete(test, [r["x"] for r in rs], [r["y"] for r in rs], [r["z"] for r in rs])
        for i, r in enumerate(rs):
            assert df[i] == r["df"], r
            _check(st[i], r["stat"], ("stat", r), atol=1e-11)
            _check(pv[i], r["p"], ("p", r), atol=1e-300)
    d.release()


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_discrete_scores_against_reference_golden(engine, golden, ds):
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    d = engine.register(cols, nlev)
    recs = golden[ds]["scores"]
    keys = sorted({(r["score"], r["arg"]) for r in recs})
    for score, arg in keys:
        rs = [r for r in recs if r["score"] == score and r["arg"] == arg]
        sc, _ = d.score_discrete(score, [r["node"] for r in rs], [r["parents"] for r in rs], arg)
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("score", r))
    d.release()


def test_nal_of_a_column_without_observed_values(engine):
    """-Inf, as nal_dnode_root / nal_dnode_parents return for nobs == 0
    (src/scores/discrete.nal.c:21-22, 52-53)."""
    empty = [np.full(64, np.iinfo(np.int32).min, dtype=np.int32), np.ones(64, dtype=np.int32)]
    d = engine.register(empty, [2, 2])
    sc, _ = d.score_discrete("nal", [0, 1], [[], [0]], 0.0)
    assert sc[0] == -np.inf and sc[1] == -np.inf
    d.release()


def test_counts_bit_exact_against_oracle(engine):
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("alarm")
    d = engine.register(cols, nlev)
    rng = np.random.default_rng(7)
    xs, ys, zs = [], [], []
    for _ in range(40):
        k = int(rng.integers(0, 5))
        idx = rng.choice(len(cols), size=2 + k, replace=False)
        xs.append(int(idx[0])); ys.append(int(idx[1]) if rng.random() < 0.8 else -1)
        zs.append([int(i) for i in idx[2:]])
    tabs = d.counts(xs, ys, zs)
    for x, y, z, t in zip(xs, ys, zs, tabs):
        ref = orc.counts(cols[x], nlev[x], cols[y] if y >= 0 else None, nlev[y] if y >= 0 else 1,
                         [cols[i] for i in z], [nlev[i] for i in z])
        assert np.array_equal(t, ref), (x, y, z)
    d.release()


def test_na_and_ragged_rows(engine):
    """NA rows are skipped; nobs is not a multiple of the 16-row vector width."""
    from oracle import oracle as orc
    rng = np.random.default_rng(11)
    n = 100003
    lev = [3, 2, 4, 5, 2]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    for c in cols:
        c[rng.random(n) < 0.03] = orc.NA_INT
    d = engine.register(cols, lev)
    xs, ys, zs = [0, 1, 2, 0], [1, 2, 3, 4], [[], [0], [0, 4], [1, 2, 3]]
    tabs = d.counts(xs, ys, zs)
    for x, y, z, t in zip(xs, ys, zs, tabs):
        ref = orc.counts(cols[x], lev[x], cols[y], lev[y], [cols[i] for i in z], [lev[i] for i in z])
        assert np.array_equal(t, ref)
    for test in ("mi", "x2", "mi-adf", "x2-adf"):
        st, df, pv = d.ci_discrete(test, xs, ys, zs)
        for i, (x, y, z) in enumerate(zip(xs, ys, zs)):
            rs, rdf, rp = orc.ci_discrete(test, cols[x], lev[x], cols[y], lev[y],
                                          [cols[j] for j in z], [lev[j] for j in z])
            assert df[i] == rdf
            _check(st[i], rs, (test, "stat", i), atol=1e-11)
            _check(pv[i], rp, (test, "p", i), atol=1e-300)
    d.release()


def test_empty_and_tiny_inputs(engine):
    d = engine.register([np.zeros(0, np.int32), np.zeros(0, np.int32)], [3, 2])
    st, df, pv = d.ci_discrete("mi", [0], [1], [[]])
    assert st[0] == 0.0 and df[0] == 2.0 and pv[0] == 1.0
    sc, npar = d.score_discrete("bic", [0], [[1]], 1.0)
    assert sc[0] == -4.0 and npar[0] == 4.0
    st, df, pv = d.ci_discrete("mi", [], [], [])
    assert len(st) == 0
    d.release()
    d = engine.register([np.array([1, 2, 1], np.int32), np.array([2, 2, 1], np.int32)], [2, 2])
    t = d.counts([0], [1], [[]])[0]
    assert t.tolist() == [[[1, 1], [0, 1]]]
    d.release()


def test_large_table_paths(engine):
    """Tables beyond the shared-memory capacity take the global-atomics path."""
    from oracle import oracle as orc
    rng = np.random.default_rng(5)
    n = 50000
    lev = [4, 3, 16, 16, 16, 16]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    d = engine.register(cols, lev)
    z = [2, 3, 4, 5]                      # 65536 configurations x 12 cells
    t = d.counts([0], [1], [z])[0]
    ref = orc.counts(cols[0], 4, cols[1], 3, [cols[i] for i in z], [lev[i] for i in z])
    assert np.array_equal(t, ref)
    for test in ("mi", "mi-adf"):
        st, df, pv = d.ci_discrete(test, [0], [1], [z])
        rs, rdf, rp = orc.ci_discrete(test, cols[0], 4, cols[1], 3, [cols[i] for i in z], [lev[i] for i in z])
        assert df[0] == rdf
        _check(st[0], rs, "stat", atol=1e-11)
        _check(pv[0], rp, "p", atol=1e-300)
    for score, arg in (("bic", 2.0), ("bde", 1.0), ("bds", 1.0)):
        sc, _ = d.score_discrete(score, [0], [z], arg)
        ref_s, _ = orc.score_discrete(score, cols[0], 4, [cols[i] for i in z], [lev[i] for i in z], arg)
        _check(sc[0], ref_s, score)
    d.release()


def test_error_paths(engine):
    from bnlearn_b200 import BnlError
    d = engine.register([np.ones(8, np.int32)] * 3, [2, 2, 2])
    with pytest.raises(BnlError):
        d.ci_discrete("mi", [0], [7], [[]])
    with pytest.raises(BnlError):
        d.ci_gaussian("cor", [0], [1], [[]])
    d.release()
    with pytest.raises(BnlError):
        engine.register([np.array([1, 5], np.int32)], [3])       # code outside 1..nlevels
    with pytest.raises(BnlError):
        engine.register([np.ones(4, np.int32)], [40000])         # beyond the uint16 layout
    big = engine.register([np.ones(4, np.int32)] * 8, [100] * 8)
    with pytest.raises(BnlError) as ei:
        big.ci_discrete("mi", [0], [1], [[2, 3, 4, 5, 6]])
    assert ei.value.code == 4                                    # BNL_EOVERFLOW
    big.release()


def test_wide_factors_uint16_layout(engine):
    """Factors with 128..32767 levels live in the uint16 block: counts bit-exact,
    tests and scores against the oracle, in every role (x, y, conditioning variable,
    node, parent), with NA, through a device-side resample, and next to uint8 columns."""
    from oracle import oracle as orc
    rng = np.random.default_rng(77)
    n = 60_011
    lev = [300, 3, 1500, 4, 130, 2]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    cols[1] = (1 + (cols[0] + rng.integers(0, 2, n)) % 3).astype(np.int32)     # depends on the wide column
    cols[4][rng.random(n) < 0.01] = np.iinfo(np.int32).min
    cols[3][rng.random(n) < 0.01] = np.iinfo(np.int32).min
    d = engine.register(cols, lev)
    x, y = [0, 1, 3, 4, 5, 0], [1, 2, 5, 3, 1, 4]
    z = [[], [], [0], [5], [4, 3], [3]]
    for t, (a, b, c) in zip(d.counts(x, y, z), zip(x, y, z)):
        ref = orc.counts(cols[a], lev[a], cols[b], lev[b], [cols[i] for i in c], [lev[i] for i in c])
        assert np.array_equal(t, ref), (a, b, c)
    for test in ("mi", "x2", "mi-adf", "jt", "mi-sh"):
        st, df, pv = d.ci_discrete(test, x, y, z)
        for i in range(len(x)):
            if test == "jt" and lev[x[i]] * lev[y[i]] > 5000:
                continue                                  # the oracle's JT is O(llx^2 lly^2)
            o = orc.ci_discrete(test, cols[x[i]], lev[x[i]], cols[y[i]], lev[y[i]], [cols[j] for j in z[i]],
                                [lev[j] for j in z[i]])
            assert df[i] == o[1], (test, i)
            # shrinkage MI over a 156 000-cell table is a sum of cancelling terms: two
            # double-precision evaluations agree to ~1e-8 there, not 1e-9
            loose = test == "mi-sh" and lev[x[i]] * lev[y[i]] > 5000
            _check(st[i], o[0], (test, i, "stat"), rtol=1e-6 if loose else RTOL, atol=1e-9)
            _check(pv[i], o[2], (test, i, "p"), rtol=1e-6 if loose else RTOL, atol=1e-300)
    tg, pa = [0, 1, 2, 5, 3], [[], [0], [5], [2, 1], [0, 5]]
    complete = [c.copy() for c in cols]
    for c, l in zip(complete, lev):
        c[c == np.iinfo(np.int32).min] = 1
    dc = engine.register(complete, lev)
    for score, arg in (("loglik", 0.0), ("bic", 5.5), ("bde", 1.0), ("bds", 1.0), ("k2", 0.0), ("qnml", 0.0)):
        sc, npar = dc.score_discrete(score, tg, pa, arg)
        for i in range(len(tg)):
            o = orc.score_discrete(score, complete[tg[i]], lev[tg[i]], [complete[j] for j in pa[i]],
                                   [lev[j] for j in pa[i]], arg)
            _check(sc[i], o[0], (score, i))
    rows = rng.integers(0, n, 20_000)
    rs = d.resample(rows)
    rr = engine.register([c[rows] for c in cols], lev)
    for ta, tb in zip(rs.counts(x, y, z), rr.counts(x, y, z)):
        assert np.array_equal(ta, tb)
    for q in (d, dc, rs, rr):
        q.release()


@pytest.mark.parametrize("ds", ["gaussian_test", "gaussian_test_na"])
def test_gaussian_ci_against_reference_golden(engine, golden, ds):
    names, cols = load_gaussian(ds)
    d = engine.register(cols)
    recs = golden[ds]["tests"]
    for test in sorted({r["test"] for r in recs}):
        rs = [r for r in recs if r["test"] == test]
        st, df, pv = d.ci_gaussian(test, [r["x"] for r in rs], [r["y"] for r in rs], [r["z"] for r in rs])
        for i, r in enumerate(rs):
            assert df[i] == r["df"]
            _check(st[i], r["stat"], ("stat", r), atol=1e-12)
            _check(pv[i], r["p"], ("p", r), atol=1e-300)
    d.release()


@pytest.mark.parametrize("ds", ["gaussian_test", "gaussian_test_na"])
def test_gaussian_scores_against_reference_golden(engine, golden, ds):
    names, cols = load_gaussian(ds)
    d = engine.register(cols)
    recs = golden[ds]["scores"]
    keys = sorted({(r["score"], r["arg"]) for r in recs})
    for score, arg in keys:
        rs = [r for r in recs if r["score"] == score and r["arg"] == arg]
        sc, _ = d.score_gaussian(score, [r["node"] for r in rs], [r["parents"] for r in rs], arg)
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("score", r))
    d.release()


def test_conditional_gaussian_scores_against_reference_golden(engine):
    """loglik-cg / aic-cg / bic-cg on the bundled mixed data set: discrete, Gaussian,
    conditional-Gaussian and impossible (-Inf) nodes in one batch."""
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_cg.json")) as f:
        recs = json.load(f)["clgaussian_test"]["scores"]
    names, cols, nlev = load_mixed("clgaussian_test")
    d = engine.register(cols, nlev)
    for score, arg in sorted({(r["score"], r["arg"]) for r in recs}):
        rs = [r for r in recs if r["score"] == score and r["arg"] == arg]
        sc, _ = d.score_cg(score, [r["node"] for r in rs], [r["parents"] for r in rs], arg)
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("score", r))
    # a configuration with too few rows makes the node singular; missing values are refused
    from oracle import oracle as orc
    n = 40
    rng = np.random.default_rng(3)
    f = np.array([1] * 37 + [2, 2, 3], dtype=np.int32)
    g1, g2 = rng.normal(size=n), rng.normal(size=n)
    small = engine.register([f, g1, g2], [3, 0, 0])
    sc, npar = small.score_cg("loglik-cg", [1, 1], [[0], [0, 2]])
    assert sc[0] == -np.inf and sc[1] == -np.inf and npar[0] == 6 and npar[1] == 9
    assert orc.score_cg("loglik-cg", g1, 0, [f], [3])[0] == -np.inf
    small.release()
    g1[5] = np.nan
    bad = engine.register([f, g1, g2], [3, 0, 0])
    with pytest.raises(Exception):
        bad.score_cg("loglik-cg", [1], [[0]])
    bad.release()
    d.release()


def test_conditional_gaussian_with_a_wide_discrete_parent(engine):
    """A factor with more than 127 levels (uint16 block) as discrete parent of a
    numeric node, alone and next to a uint8 factor and a numeric parent: the moment
    kernel reads its codes from the wide block. Against the oracle."""
    from oracle import oracle as orc
    rng = np.random.default_rng(41)
    n = 60000
    wide = rng.integers(1, 201, n).astype(np.int32)                       # 200 levels
    f = rng.integers(1, 4, n).astype(np.int32)
    g1 = rng.normal(size=n)
    g0 = 0.01 * wide + 0.5 * f + 0.7 * g1 + rng.normal(size=n)
    cols, lev = [g0, wide, f, g1], [0, 200, 3, 0]
    d = engine.register(cols, lev)
    units = [[1], [1, 3], [1, 2], [2, 1, 3]]
    for score, k in (("loglik-cg", 0.0), ("bic-cg", np.log(n) / 2)):
        sc, npar = d.score_cg(score, [0] * len(units), units, k)
        for i, pa in enumerate(units):
            ref, rnp = orc.score_cg(score, g0, 0, [cols[j] for j in pa], [lev[j] for j in pa], k)
            _check(sc[i], ref, (score, pa))
            assert npar[i] == rnp
    d.release()


def test_conditional_gaussian_nal_scores_on_incomplete_data(engine):
    """nal-cg / pnal-cg on clgaussian.test with 5 % of the cells missing (re-created
    from the fixture's seed): discrete nodes (nal), Gaussian nodes with numeric parents
    only (nal-g), mixed parents (c_fast_ccgnal over the conditional least squares on
    locally complete rows) and the -Inf cases, against nal_cgnode of the reference."""
    from util import load_mixed
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
    d = engine.register(na, nlev)
    recs = fx["scores"]
    for score, arg in sorted({(r["score"], r["arg"]) for r in recs}):
        rs = [r for r in recs if r["score"] == score and r["arg"] == arg]
        sc, _ = d.score_cg(score, [r["node"] for r in rs], [r["parents"] for r in rs], arg)
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("score", r))
    # the whole search runs on these scores (the data has missing values everywhere)
    amat, scores, iters = d.hc("pnal-cg", np.log(len(cols[0])) / 2)
    pa = [[int(i) for i in np.nonzero(amat[:, j])[0]] for j in range(len(cols))]
    chk, _ = d.score_cg("pnal-cg", list(range(len(cols))), pa, np.log(len(cols[0])) / 2)
    for j in range(len(cols)):
        _check(scores[j], chk[j], ("hc node score", j))
    d.release()


def test_conditional_gaussian_tests_against_reference_golden(engine):
    """mi-cg on the bundled mixed data: all 16 combinations of factor / numeric x, y
    and conditioning sets, in one batched call."""
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_cg.json")) as f:
        recs = json.load(f)["clgaussian_test"]["tests"]
    names, cols, nlev = load_mixed("clgaussian_test")
    d = engine.register(cols, nlev)
    st, df, pv = d.ci_cg([r["x"] for r in recs], [r["y"] for r in recs], [r["z"] for r in recs])
    for i, r in enumerate(recs):
        assert df[i] == r["df"], r
        _check(st[i], r["stat"], ("stat", r), atol=1e-9)
        _check(pv[i], r["p"], ("p", r), rtol=1e-8, atol=1e-300)
    d.release()


def test_conditional_gaussian_tests_on_incomplete_data(engine):
    """mi-cg with 5 % of the cells missing (ct_micg_with_missing): every test on the
    rows complete for all its columns -- a temporary data set gathered on the device --
    in one call next to tests that would be complete anyway; no data set is leaked."""
    from util import load_mixed
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
    na[0] = cols[0]                                     # one complete column pair: a complete test in the batch
    na[5] = cols[5]
    d = engine.register(na, nlev)
    recs = [r for r in fx["tests"] if 0 not in [r["x"], r["y"]] + r["z"] and 5 not in [r["x"], r["y"]] + r["z"]]
    assert len(recs) > 10
    xs, ys, zs = [r["x"] for r in recs] + [0], [r["y"] for r in recs] + [5], [r["z"] for r in recs] + [[]]
    st, df, pv = d.ci_cg(xs, ys, zs)
    for i, r in enumerate(recs):
        assert df[i] == r["df"], r
        _check(st[i], r["stat"], ("stat", r), atol=1e-9)
        _check(pv[i], r["p"], ("p", r), rtol=1e-8, atol=1e-300)
    whole = engine.register(cols, nlev)                 # the complete pair equals the complete data's test
    s2, d2, p2 = whole.ci_cg([0], [5], [[]])
    assert st[-1] == s2[0] and df[-1] == d2[0] and pv[-1] == p2[0]
    whole.release()
    d.release()


def test_ebic_against_reference_golden(engine):
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_ebic.json")) as f:
        g = json.load(f)
    for ds, recs in g.items():
        if ds == "learning_test":
            names, cols, nlev = load_discrete(ds)
        elif ds == "gaussian_test":
            names, cols = load_gaussian(ds)
            nlev = None
        else:
            names, cols, nlev = load_mixed(ds)
        d = engine.register(cols, nlev)
        for gamma in (0.5, 0.3):
            rs = [r for r in recs if r["gamma"] == gamma]
            if gamma != 0.5:
                d.set_ebic_gamma(gamma)                  # 0.5 is the default
            sc, _ = d.score(rs[0]["score"], [r["node"] for r in rs], [r["parents"] for r in rs], rs[0]["arg"])
            for i, r in enumerate(rs):
                _check(sc[i], r["value"], ("ebic", ds, r))
        d.release()


def test_bge_against_reference_golden(engine):
    with open(os.path.join(GOLDEN, "ref_bge.json")) as f:
        recs = json.load(f)["gaussian_test"]
    names, cols = load_gaussian("gaussian_test")
    d = engine.register(cols)
    default = [r for r in recs if r["iss_mu"] == 1.0]
    custom = [r for r in recs if r["iss_mu"] != 1.0]
    for rs, configure in ((default, lambda: None),
                          (custom, lambda: d.set_bge_prior(custom[0]["iss_mu"], custom[0]["iss_w"], custom[0]["nu"]))):
        configure()
        sc, _ = d.score_gaussian("bge", [r["node"] for r in rs], [r["parents"] for r in rs])
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("bge", r["node"], r["parents"]))
    d.set_bge_prior()                                  # back to the defaults
    sc, _ = d.score_gaussian("bge", [default[0]["node"]], [default[0]["parents"]])
    _check(sc[0], default[0]["value"], "bge defaults restored")
    with pytest.raises(Exception):
        d.set_bge_prior(1.0, len(names))               # iss.w must exceed ncol + 1
    d.release()


def test_gaussian_incomplete_wide_units_and_edge_cases(engine):
    """Per-unit complete-row moments: units wider than 8 columns (the spilled
    variant), a column with no observed value, NAL-G on complete data."""
    from oracle import oracle as orc
    rng = np.random.default_rng(21)
    n, p = 3000, 12
    base = rng.normal(size=(n, 3))
    cols = [base @ rng.normal(size=3) + rng.normal(size=n) for _ in range(p)]
    for c in cols[:6]:
        c[rng.random(n) < 0.03] = np.nan
    cols.append(np.full(n, np.nan))
    d = engine.register(cols)
    x, y = [0, 1, 2, 12], [3, 4, 5, 0]
    z = [[6, 7, 8, 9, 10, 11, 1, 2], [0], [], [1]]
    for test in ("cor", "zf", "mi-g", "mi-g-sh"):
        st, df, pv = d.ci_gaussian(test, x, y, z)
        for i in range(len(x)):
            o = orc.ci_gaussian(test, [cols[x[i]], cols[y[i]]] + [cols[j] for j in z[i]])
            assert df[i] == o[1]
            _check(st[i], o[0], (test, i, "stat"), atol=1e-12)
            _check(pv[i], o[2], (test, i, "p"), atol=1e-300)
    tg, pa = [0, 7, 12, 3], [[1, 2, 3, 4, 5, 6, 7, 8, 9], [], [], [12]]
    sc, _ = d.score_gaussian("pnal-g", tg, pa, 2.0)
    for i in range(len(tg)):
        o = orc.score_gaussian("pnal-g", cols[tg[i]], [cols[j] for j in pa[i]], 2.0)[0]
        _check(sc[i], o, ("pnal-g", i))
    assert sc[2] == -np.inf and sc[3] == -np.inf
    d.release()
    names, full = load_gaussian("gaussian_test")
    d = engine.register(full)
    sc, _ = d.score_gaussian("nal-g", [0, 5], [[], [0, 1, 2]], 0.0)
    ll, _ = d.score_gaussian("loglik-g", [0, 5], [[], [0, 1, 2]], 0.0)
    for a, b in zip(sc, ll):
        _check(a, b / len(full[0]), "nal-g on complete data")
    d.release()


def test_gaussian_moments_against_numpy(engine):
    rng = np.random.default_rng(3)
    n, p = 20011, 70
    x = rng.normal(size=(p, n)) * rng.uniform(0.5, 3, size=(p, 1)) + rng.normal(size=(p, 1)) * 10
    x[5] = 0.3 * x[1] - 2 * x[2] + rng.normal(size=n) * 0.1
    d = engine.register(list(x))
    mean, cov = d.moments()
    assert np.allclose(mean, x.mean(axis=1), rtol=1e-13, atol=1e-13)
    ref = np.cov(x)
    assert np.allclose(cov, ref, rtol=1e-11, atol=1e-12)
    assert np.array_equal(cov, cov.T)
    d.release()


@pytest.mark.parametrize("scratch", ["1", "0"])
def test_gram_routes_against_numpy(engine, monkeypatch, scratch):
    """The Gram kernel's two routes -- the data set centred into scratch memory first, or
    each stage centred in shared memory when there is no room for that -- on a shape with
    ragged rows (not a multiple of the 32-row stage), a ragged last column tile and means far
    from zero (c_covmat, src/core/covariance.matrix.c:65-97)."""
    monkeypatch.setenv("BNL_GRAM_SCRATCH", scratch)
    rng = np.random.default_rng(11)
    n, p = 70003, 161
    x = rng.normal(size=(p, n)) * rng.uniform(0.5, 3, size=(p, 1)) + rng.normal(size=(p, 1)) * 100
    x[7] = 0.5 * x[3] - x[150] + rng.normal(size=n) * 0.05
    d = engine.register(list(x))
    mean, cov = d.moments()
    d.release()
    assert np.allclose(mean, x.mean(axis=1), rtol=1e-13, atol=1e-13)
    xc = x - x.mean(axis=1, keepdims=True)
    ref = (xc @ xc.T) / (n - 1)
    assert np.allclose(cov, ref, rtol=1e-11, atol=1e-11)
    assert np.array_equal(cov, cov.T)


def test_gaussian_against_oracle_synthetic(engine):
    from oracle import oracle as orc
    rng = np.random.default_rng(9)
    n, p = 3001, 12
    x = rng.normal(size=(p, n))
    for j in range(2, p):
        x[j] += 0.6 * x[j - 1] - 0.4 * x[j - 2]
    cols = list(x)
    d = engine.register(cols)
    xs, ys, zs = [], [], []
    for _ in range(60):
        k = int(rng.integers(0, 7))
        idx = rng.choice(p, size=2 + k, replace=False)
        xs.append(int(idx[0])); ys.append(int(idx[1])); zs.append([int(i) for i in idx[2:]])
    for test in ("cor", "zf", "mi-g"):
        st, df, pv = d.ci_gaussian(test, xs, ys, zs)
        for i in range(len(xs)):
            rs, rdf, rp, _ = orc.ci_gaussian(test, [cols[xs[i]], cols[ys[i]]] + [cols[j] for j in zs[i]])
            assert df[i] == rdf
            _check(st[i], rs, (test, "stat", i), atol=1e-12)
            _check(pv[i], rp, (test, "p", i), atol=1e-300)
    for score in ("loglik-g", "bic-g"):
        sc, _ = d.score_gaussian(score, xs, zs, np.log(n) / 2)
        for i in range(len(xs)):
            ref, _ = orc.score_gaussian(score, cols[xs[i]], [cols[j] for j in zs[i]], np.log(n) / 2)
            _check(sc[i], ref, (score, i))
    d.release()


def test_special_functions_on_device_match_host_build(engine):
    """The device build of bnl_special.cuh agrees with mpmath through real tests:
    a t test at a million degrees of freedom (config D's regime)."""
    from oracle import oracle as orc
    rng = np.random.default_rng(21)
    n = 1000003
    a = rng.normal(size=n)
    b = 0.002 * a + rng.normal(size=n)
    c = rng.normal(size=n)
    d = engine.register([a, b, c])
    for test in ("cor", "zf"):
        st, df, pv = d.ci_gaussian(test, [0, 0], [1, 2], [[], [1]])
        for i, cols in enumerate(([a, b], [a, c, b])):
            rs, rdf, rp, _ = orc.ci_gaussian(test, cols)
            assert df[i] == rdf
            _check(st[i], rs, (test, "stat", i), rtol=1e-9, atol=1e-13)
            _check(pv[i], rp, (test, "p", i))
    d.release()


@pytest.mark.parametrize("with_na", [False, True])
def test_every_table_size_class(engine, with_na):
    """One unit per counting path: <= 256 cells, <= 880, <= ~1800, shared-memory
    table, global-memory table; with and without NAs; ragged row count."""
    from oracle import oracle as orc
    rng = np.random.default_rng(17)
    n = 70001
    lev = [5, 4, 5, 5, 6, 3, 7, 2, 9]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    if with_na:
        for c in cols:
            c[rng.random(n) < 0.01] = orc.NA_INT
    d = engine.register(cols, lev)
    units = [(0, 1, [5]),                 # 60 cells
             (0, 1, [2, 3]),              # 500 cells
             (0, 1, [2, 3, 5]),           # 1500 cells
             (0, 1, [2, 3, 4, 7]),        # 6000 cells
             (0, 1, [2, 3, 4, 6, 8]),     # 189000 cells: beyond shared memory
             (6, 8, [0, 1, 2, 3, 4, 5])]  # 6 conditioning columns, generic column loop
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    tabs = d.counts(xs, ys, zs)
    for (x, y, z), t in zip(units, tabs):
        ref = orc.counts(cols[x], lev[x], cols[y], lev[y], [cols[i] for i in z], [lev[i] for i in z])
        assert np.array_equal(t, ref), (x, y, z)
    st, df, pv = d.ci_discrete("mi", xs, ys, zs)
    for i, (x, y, z) in enumerate(units):
        rs, rdf, rp = orc.ci_discrete("mi", cols[x], lev[x], cols[y], lev[y], [cols[j] for j in z],
                                      [lev[j] for j in z])
        assert df[i] == rdf
        _check(st[i], rs, ("stat", i), atol=1e-11)
        _check(pv[i], rp, ("p", i), atol=1e-300)
    d.release()


@pytest.mark.parametrize("n", [31, 70001, 300003])
def test_nibble_copy_route_counts_bit_exact(engine, monkeypatch, n):
    """count_small_kernel and count_pair32_kernel (tables of <= 32 cells as row pairs) reading
    the nibble copy of the columns (two codes per byte; units of <= 256 cells whose columns have
    at most 15 levels and no NA): tables bit-exact against
    fill_1d/2d/3d_table (src/core/contingency.tables.c:20-223) for 1-6 columns per table,
    ragged row counts (the rows past the last 32-row group come from the byte columns), next
    to units that must stay on the byte route (a 16-level column, a column with NA)."""
    from oracle import oracle as orc
    monkeypatch.setenv("BNL_NIBBLES", "1")
    monkeypatch.setenv("BNL_NIBBLES_AFTER", "0")
    monkeypatch.setenv("BNL_PAIRS", "0")           # keep the one- and two-way tables in this kernel too
    rng = np.random.default_rng(23)
    lev = [2, 3, 4, 2, 2, 15, 16, 3, 2, 5]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    cols[7][rng.random(n) < 0.02] = orc.NA_INT      # column 7 holds NA
    d = engine.register(cols, lev)
    units = [(0, None, []), (1, 2, []), (5, 1, []), (0, 1, [2]), (1, 2, [0, 3]), (0, 1, [3, 4, 8]),
             (0, 3, [4, 8, 1, 2]), (9, 5, [0]),
             (6, 1, []), (1, 7, [0]), (0, 6, [3]),        # byte route: 16 levels, NA
             # enough tables of <= 32 cells for the row-pair launch (count_pair32_kernel): 1 - 5 columns
             (3, None, []), (0, 3, []), (3, 4, []), (0, 4, [8]), (1, 8, []), (2, 3, [4]), (8, 0, [3]), (4, 1, [0]),
             (2, 8, []), (9, 0, [3]), (0, 3, [4, 8]), (3, 4, [0, 8, 2]), (2, 9, []), (1, 9, [3])]
    xs, ys, zs = [u[0] for u in units], [u[1] if u[1] is not None else -1 for u in units], [u[2] for u in units]
    for rep in range(2):                             # second call: the copy exists already
        tabs = d.counts(xs, ys, zs)
        for (x, y, z), t in zip(units, tabs):
            if y is None:
                ref = np.bincount(cols[x] - 1, minlength=lev[x])
                assert np.array_equal(np.asarray(t).ravel(), ref), (x, n)
                continue
            ref = orc.counts(cols[x], lev[x], cols[y], lev[y], [cols[i] for i in z], [lev[i] for i in z])
            assert np.array_equal(t, ref), (x, y, z, n)
    d.release()


def test_nibble_copy_is_rebuilt_after_commit(engine, monkeypatch):
    """bnl_data_commit voids a nibble copy made from the previous contents of the columns."""
    import torch
    monkeypatch.setenv("BNL_NIBBLES", "1")
    monkeypatch.setenv("BNL_NIBBLES_AFTER", "0")
    monkeypatch.setenv("BNL_PAIRS", "0")
    rng = np.random.default_rng(5)
    n, lev = 50000, [3, 2, 4]
    d = engine.alloc(lev, n)
    u8, nu, _, _ = d.device_view(0)
    from bnlearn_b200.replicate import _as_tensor
    view = _as_tensor(u8, nu).view(len(lev), -1)
    for trial in range(2):
        cols = [rng.integers(0, l, n).astype(np.uint8) for l in lev]
        for i, c in enumerate(cols):
            view[i, :n] = torch.from_numpy(c).cuda()
        torch.cuda.synchronize()
        d.commit()
        t = d.counts([0], [1], [[2]])[0]
        ref = np.zeros((lev[2], lev[0], lev[1]), dtype=np.int64)
        np.add.at(ref, (cols[2], cols[0], cols[1]), 1)
        assert np.array_equal(np.asarray(t).reshape(ref.shape), ref), trial
    d.release()


def test_mbde_against_reference_golden(engine):
    """mBDe = BDe over the rows on which the node was not intervened upon: the
    interventions arrive as one-level mask columns (NA on those rows) registered
    next to the data. Against per_node_score("mbde") of the reference; nodes
    without a mask score as BDe; a resampled data set keeps its masks."""
    from oracle import oracle as orc
    with open(os.path.join(GOLDEN, "ref_mbde.json")) as f:
        fx = json.load(f)["learning_test"]
    names, cols, nlev = load_discrete("learning_test")
    n, m = len(cols[0]), len(cols)
    masks = {}
    for node, rows in fx["exp"].items():
        mk = np.ones(n, dtype=np.int32)
        mk[np.asarray(rows) - 1] = orc.NA_INT
        masks[int(node)] = mk
    order = sorted(masks)
    d = engine.register(cols + [masks[k] for k in order], nlev + [1] * len(order))
    for j, k in enumerate(order):
        d.set_interventions(k, m + j)
    for iss in (1.0, 10.0):
        rs = [r for r in fx["scores"] if r["iss"] == iss]
        sc, npar = d.score_discrete("mbde", [r["node"] for r in rs], [r["parents"] for r in rs], iss)
        for i, r in enumerate(rs):
            _check(sc[i], r["value"], ("mbde", r["node"], r["parents"], iss))
            assert npar[i] == (nlev[r["node"]] - 1) * int(np.prod([nlev[p] for p in r["parents"]]))
    a, _ = d.score_discrete("mbde", [1, 2], [[0], [1, 3]], 1.0)           # no mask on B, C
    b, _ = d.score_discrete("bde", [1, 2], [[0], [1, 3]], 1.0)
    assert np.array_equal(a, b)
    rows = np.arange(n, dtype=np.int32)[::-1].copy()                       # a permutation of the rows
    r = d.resample(rows)
    c, _ = r.score_discrete("mbde", [0, 3], [[1, 2], [0]], 1.0)
    e, _ = d.score_discrete("mbde", [0, 3], [[1, 2], [0]], 1.0)
    for i in range(2):
        _check(c[i], e[i], ("resampled", i))
    # the whole search on the six nodes (the masks are columns 6..8, never nodes):
    # the node scores it returns are those of its final graph
    amat, scores, iters = d.hc("mbde", 1.0, nnodes=m)
    assert iters > 0 and amat.sum() > 0
    pa = [[int(i) for i in np.nonzero(amat[:, j])[0]] for j in range(m)]
    chk, _ = d.score_discrete("mbde", list(range(m)), pa, 1.0)
    for j in range(m):
        _check(scores[j], chk[j], ("hc node score", j))
    from bnlearn_b200 import BnlError
    with pytest.raises(BnlError):
        d.set_interventions(0, 1)                                          # not a one-level column
    r.release()
    d.release()


def test_candidate_runs_with_mid_size_tables(engine):
    """score_cache_fill's batch for a node that already has 2-3 parents: the
    candidates share every column but the last, and their tables (240-3600 cells)
    are too large for several to share an SM. Such runs go through the base-index
    column (count_based_kernel, every counters-per-cell class); tables that stay
    small go to the group kernel and a short run to the per-unit kernels. Ragged
    row count, scores against the oracle."""
    from oracle import oracle as orc
    rng = np.random.default_rng(23)
    n = 70003
    lev = [3, 4, 5, 5, 2, 3, 4, 5, 6, 7, 9, 12, 4, 5, 6, 7, 3, 2]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    # some dependence, so that the tables are not flat
    cols[6] = ((cols[0] + cols[1] + rng.integers(0, 2, n)) % lev[6] + 1).astype(np.int32)
    d = engine.register(cols, lev)
    targets, parents = [], []
    for base in ([1, 2], [1, 2, 3]):          # 60 and 300 base cells
        for c in range(4, 18):
            targets.append(0)
            parents.append(base + [c])
    for c in (4, 5, 6):                       # a run of three: too short, per-unit kernels
        targets.append(1)
        parents.append([2, 3, c])
    for score, arg in (("loglik", 0.0), ("bde", 1.0)):
        got, npar = d.score_discrete(score, targets, parents, arg)
        for i, (t, pa) in enumerate(zip(targets, parents)):
            ref, _ = orc.score_discrete(score, cols[t], lev[t], [cols[j] for j in pa], [lev[j] for j in pa], arg)
            assert npar[i] == (lev[t] - 1) * int(np.prod([lev[j] for j in pa]))
            _check(got[i], ref, (score, t, pa))
    d.release()


@pytest.mark.parametrize("seed,n", [(1, 65536), (2, 65537), (3, 131071), (4, 200015)])
def test_candidate_runs_randomised(engine, seed, n):
    """Random score_cache_fill-like batches: several nodes, each with its own parents
    and a run of candidates, interleaved with unrelated units, at row counts around
    the 16-row group and chunk boundaries. Every unit against the oracle, so that the
    group kernel, the base-index kernel (all classes) and the per-unit kernels are
    each hit with odd shapes."""
    from oracle import oracle as orc
    rng = np.random.default_rng(seed)
    p = 22
    lev = [int(v) for v in rng.integers(2, 7, p)]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    if seed == 3:                                        # a column with NAs: its units leave the grouped paths
        cols[5][rng.random(n) < 0.02] = orc.NA_INT
    d = engine.register(cols, lev)
    targets, parents = [], []
    for _ in range(5):
        t = int(rng.integers(0, p))
        k = int(rng.integers(1, 5))
        base = [int(v) for v in rng.choice([c for c in range(p) if c != t], size=k, replace=False)]
        cands = [c for c in range(p) if c != t and c not in base]
        rng.shuffle(cands)
        for c in cands[:int(rng.integers(3, len(cands) + 1))]:
            targets.append(t); parents.append(base + [int(c)])
        targets.append(int(rng.integers(0, p))); parents.append([])          # breaks the run
    got, _ = d.score_discrete("loglik", targets, parents, 0.0)
    for i, (t, pa) in enumerate(zip(targets, parents)):
        if int(np.prod([lev[j] for j in pa] + [lev[t]])) > 200000:
            continue
        ref, _ = orc.score_discrete("loglik", cols[t], lev[t], [cols[j] for j in pa], [lev[j] for j in pa], 0.0)
        _check(got[i], ref, (seed, t, pa))
    d.release()


def test_multi_device_context_gives_identical_results():
    """A context on two devices splits the units in contiguous blocks; outputs must
    not change by a bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from bnlearn_b200 import Engine
    names, cols, nlev = load_discrete("alarm")
    gn, gcols = load_gaussian("gaussian_test")
    rng = np.random.default_rng(23)
    xs, ys, zs = [], [], []
    for _ in range(101):
        k = int(rng.integers(0, 4))
        idx = rng.choice(len(cols), size=2 + k, replace=False)
        xs.append(int(idx[0])); ys.append(int(idx[1])); zs.append([int(i) for i in idx[2:]])
    out = []
    for devs in ([0], [0, 1]):
        with Engine(devs) as e:
            d = e.register(cols, nlev)
            g = e.register(gcols)
            out.append((d.ci_discrete("mi", xs, ys, zs), d.score_discrete("bde", xs, zs, 1.0),
                        g.ci_gaussian("cor", [0, 1, 2], [3, 4, 5], [[6], [], [0, 1]])))
    for a, b in zip(out[0], out[1]):
        for u, v in zip(a, b):
            assert np.array_equal(u, v)
    # candidate runs (group and base-index kernels need >= 65536 rows): the split may cut a
    # run in two, every device works with its own scratch columns
    n = 70003
    lev = [3, 4, 5, 5, 2, 3, 4, 5, 6, 7, 9, 12, 4, 5, 6, 7, 3, 2]
    ccols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    targets, parents = [], []
    for base in ([1, 2], [1, 2, 3], [2]):
        for c in range(4, 18):
            targets.append(0); parents.append(base + [c])
    res = []
    for devs in ([0], [0, 1]):
        with Engine(devs) as e:
            d = e.register(ccols, lev)
            res.append(d.score_discrete("bde", targets, parents, 1.0)[0])
    assert np.array_equal(res[0], res[1])


@pytest.mark.parametrize("threads", [None, "0", "3"])
def test_large_registration_hybrid_upload(engine, threads, monkeypatch):
    """Data frames above 8M values are uploaded by two cooperating routes (host
    cores packing into a pinned ring + raw copies packed on the device); whatever
    the split, the packed data -- hence every count -- must be identical, NA
    flags included, and a bad code must still be refused."""
    from bnlearn_b200 import BnlError
    if threads is not None:
        monkeypatch.setenv("BNL_PACK_THREADS", threads)
    rng = np.random.default_rng(31)
    n = 2_100_003
    lev = [3, 4, 2, 5, 3]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    cols[1][rng.random(n) < 0.001] = np.iinfo(np.int32).min
    cols[4][-1] = np.iinfo(np.int32).min
    d = engine.register(cols, lev)
    xs, ys, zs = [0, 1, 3, 4], [2, 0, 1, 3], [[], [2], [0, 2], [0, 1, 2]]
    tabs = d.counts(xs, ys, zs)
    for x, y, z, t in zip(xs, ys, zs, tabs):
        keep = np.ones(n, dtype=bool)
        for c in [x, y] + z:
            keep &= cols[c] != np.iinfo(np.int32).min
        key = np.zeros(n, dtype=np.int64)
        mul = 1
        for c in [y, x] + z:
            key += (cols[c].astype(np.int64) - 1) * mul
            mul *= lev[c]
        ref = np.bincount(key[keep], minlength=t.size)
        assert np.array_equal(t.ravel(), ref), (x, y, z)
    d.release()
    cols[2][1_500_000] = 7
    with pytest.raises(BnlError):
        engine.register(cols, lev)


def test_full_size_properties(engine):
    """Config B's full size (37 factors x 10 M rows), where the oracle would take
    minutes: size-independent properties of the counts and statistics instead.
    Tables sum to the row count; summing the conditional table over the strata
    gives the marginal table; a test is symmetric in (x, y); counts are additive
    over a row split and double when every row appears twice (device-side row
    gathers)."""
    from bnlearn_b200 import synth
    rows = 10_000_000
    names, cols, lev = synth.alarm_shaped(rows)
    d = engine.register(cols, lev)
    x, y, z = [3, 12, 30, 7], [17, 4, 2, 21], [[], [8], [5, 22], [1, 9, 14]]
    tabs = d.counts(x, y, z)
    marg = d.counts(x, y, [[]] * 4)
    for t, m in zip(tabs, marg):
        assert int(t.sum(dtype=np.int64)) == rows
        assert np.array_equal(t.sum(axis=0, dtype=np.int64), m[0].astype(np.int64))
    for test in ("mi", "x2"):
        s1, d1, p1 = d.ci_discrete(test, x, y, z)
        s2, d2, p2 = d.ci_discrete(test, y, x, z)
        assert np.array_equal(d1, d2)
        assert np.allclose(s1, s2, rtol=1e-12, atol=0) and np.allclose(p1, p2, rtol=1e-9, atol=1e-300)
    # additivity over a row split, and doubling
    cut = 3_333_337
    lo = d.resample(np.arange(0, cut, dtype=np.int32))
    hi = d.resample(np.arange(cut, rows, dtype=np.int32))
    for t, a, b in zip(tabs, lo.counts(x, y, z), hi.counts(x, y, z)):
        assert np.array_equal(t, a + b)
    lo.release()
    hi.release()
    twice = d.resample(np.concatenate([np.arange(rows, dtype=np.int32)] * 2))
    for t, t2 in zip(tabs, twice.counts(x, y, z)):
        assert np.array_equal(2 * t, t2)
    # G^2 scales with the sample size when every row is duplicated
    s1, _, _ = d.ci_discrete("mi", x, y, z)
    s2, _, _ = twice.ci_discrete("mi", x, y, z)
    assert np.allclose(s2, 2 * s1, rtol=1e-11, atol=0)
    twice.release()
    d.release()


def test_row_sharded_tables_give_the_unsharded_statistics(engine):
    """Rows split into three blocks (as ranks of a row-sharded run would hold
    them): per-block bnl_counts tables summed, then bnl_stat_from_tables -- equal to
    the statistics and scores of the whole data set."""
    for name in ("alarm", "learning_test_na"):
        names, cols, lev = load_discrete(name)
        n = len(cols[0])
        whole = engine.register(cols, lev)
        cuts = [0, n // 3, n // 2 + 17, n]
        shards = [engine.register([c[a:b] for c in cols], lev) for a, b in zip(cuts[:-1], cuts[1:])]
        p = len(cols)
        x, y, z = [0, 2, 4, 1], [1, 3, 5, 0], [[], [0], [1, 3], [2, 3, 4]]
        summed = [sum(t.astype(np.int64) for t in ts).astype(np.int32)
                  for ts in zip(*[s.counts(x, y, z) for s in shards])]
        for t, w in zip(summed, whole.counts(x, y, z)):
            assert np.array_equal(t, w)
        for test in ("mi", "x2", "mi-adf", "jt", "mi-sh"):
            got = engine.stat_from_tables(test, summed, [len(q) for q in z])
            ref = whole.ci_discrete(test, x, y, z)
            for g, r in zip(got, ref):
                assert np.array_equal(g, r), test
        # scores: node x[u] with parents z[u]
        tabs = [sum(t.astype(np.int64) for t in ts).astype(np.int32)
                for ts in zip(*[s.counts(x, [-1] * 4, z) for s in shards])]
        for score, arg in (("bic", 4.2), ("bde", 1.0), ("bds", 2.0), ("k2", 0.0), ("fnml", 0.0), ("qnml", 0.0),
                           ("pnal", 1.5)):
            got = engine.stat_from_tables(score, tabs, [len(q) for q in z], arg, nrows=n)
            ref = whole.score_discrete(score, x, z, arg)
            assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]), score
        for d in shards + [whole]:
            d.release()


def test_row_sharded_moments_give_the_unsharded_gaussian_statistics(engine):
    """Gaussian half of the row-sharded run: three row blocks registered apart, their
    (mean, cov) pooled as replicate.pool_moments pools them over the ranks, handed to
    one of the shards with bnl_data_set_moments -- tests, scores and the subset
    patterns then equal those of the whole data set (1e-9; the pooling changes the
    order of the sums)."""
    from bnlearn_b200 import callsites as cs
    names, cols = load_gaussian("gaussian_test")
    n = len(cols[0])
    whole = engine.register(cols)
    cuts = [0, n // 3, n // 2 + 17, n]
    shards = [engine.register([c[a:b] for c in cols]) for a, b in zip(cuts[:-1], cuts[1:])]
    ns = [b - a for a, b in zip(cuts[:-1], cuts[1:])]
    moms = [s.moments() for s in shards]
    mean = sum(k * m[0] for k, m in zip(ns, moms)) / n
    ss = sum((k - 1) * m[1] + k * np.outer(m[0] - mean, m[0] - mean) for k, m in zip(ns, moms))
    d = shards[0]
    d.set_moments(n, mean, ss / (n - 1))
    x, y, z = [0, 2, 4, 1, 6], [1, 3, 5, 0, 2], [[], [0], [1, 3], [2, 3, 4], [0, 1, 3, 4, 5]]
    for test in ("cor", "zf", "mi-g"):
        got, ref = d.ci_gaussian(test, x, y, z), whole.ci_gaussian(test, x, y, z)
        assert np.array_equal(got[1], ref[1]), test                       # df: the pooled row count
        for i in range(len(x)):
            _check(got[0][i], ref[0][i], (test, "stat", i), atol=1e-12)
            _check(got[2][i], ref[2][i], (test, "p", i), atol=1e-300)
    for score, arg in (("loglik-g", 0.0), ("bic-g", np.log(n) / 2), ("bge", 0.0)):
        got, ref = d.score_gaussian(score, x, z, arg), whole.score_gaussian(score, x, z, arg)
        for i in range(len(x)):
            _check(got[0][i], ref[0][i], (score, i))
    c1, c2 = cs.TestCounter(), cs.TestCounter()
    a = cs.allsubs_test(d, 0, 5, [1, 2, 3], [], "cor", 0.05, 0, 3, c1)
    b = cs.allsubs_test(whole, 0, 5, [1, 2, 3], [], "cor", 0.05, 0, 3, c2)
    assert a[3] == b[3] and c1.value == c2.value and rel_close(a[0], b[0], 1e-9, 1e-300)
    from bnlearn_b200 import BnlError
    with pytest.raises(BnlError):
        d.ci_gaussian_mc("mc-cor", [0], [1], [[]], 10)                     # needs the rows
    for q in shards + [whole]:
        q.release()


def test_alloc_fill_commit_equals_register(engine):
    """bnl_data_alloc / device_view / commit: a data set whose packed blocks are
    copied in on the device (what an NCCL broadcast does between ranks) answers like
    the registered original, NA flags included; a bad code is refused at commit."""
    import torch
    from bnlearn_b200.replicate import _as_tensor
    from util import load_mixed
    names, cols, lev = load_discrete("learning_test_na")
    src = engine.register(cols, lev)
    dst = engine.alloc(lev, len(cols[0]))
    su, nu, _, _ = src.device_view()
    du, nu2, _, _ = dst.device_view()
    assert nu == nu2 and nu > 0
    _as_tensor(du, nu).copy_(_as_tensor(su, nu))
    torch.cuda.synchronize()
    dst.commit()
    x, y, z = [0, 1, 2], [3, 4, 5], [[], [0], [1, 3]]
    for a, b in zip(src.ci_discrete("mi", x, y, z), dst.ci_discrete("mi", x, y, z)):
        assert np.array_equal(a, b)
    for ta, tb in zip(src.counts(x, y, z), dst.counts(x, y, z)):
        assert np.array_equal(ta, tb)
    bad = engine.alloc(lev, len(cols[0]))
    bu, nb, _, _ = bad.device_view()
    t = _as_tensor(bu, nb)
    t.copy_(_as_tensor(su, nu))
    t[5] = 100
    torch.cuda.synchronize()
    with pytest.raises(Exception):
        bad.commit()
    for d in (src, dst, bad):
        d.release()
    # mixed data: both blocks, Gaussian moments formed at commit
    names, cols, lev = load_mixed("clgaussian_test")
    src = engine.register(cols, lev)
    dst = engine.alloc(lev, len(cols[0]))
    su, nu, sf, nf = src.device_view()
    du, _, dfp, _ = dst.device_view()
    _as_tensor(du, nu).copy_(_as_tensor(su, nu))
    _as_tensor(dfp, nf).copy_(_as_tensor(sf, nf))
    torch.cuda.synchronize()
    dst.commit()
    a = src.score_cg("bic-cg", [3, 6, 0], [[0, 7], [0, 3, 4], [1]], 4.0)[0]
    b = dst.score_cg("bic-cg", [3, 6, 0], [[0, 7], [0, 3, 4], [1]], 4.0)[0]
    assert np.allclose(a, b, rtol=1e-12, atol=0)
    src.release()
    dst.release()


def test_large_numeric_registration_staged_upload(engine):
    """Numeric data frames above 8 M values in pageable memory are uploaded in 4 MiB
    chunks through pinned buffers by host threads; pinned sources go straight
    through. Either way the resident columns -- hence the moments -- are the same."""
    import torch
    rng = np.random.default_rng(41)
    n, p = 1_200_007, 8
    cols = [rng.normal(size=n) * (1 + j) + j for j in range(p)]
    cols[3] = cols[0] * 0.5 + cols[3]
    d1 = engine.register(cols)
    pinned = torch.empty((p, n), dtype=torch.float64, pin_memory=True)
    host = pinned.numpy()
    for j in range(p):
        host[j] = cols[j]
    d2 = engine.register([host[j] for j in range(p)])
    m1, c1 = d1.moments()
    m2, c2 = d2.moments()
    assert np.array_equal(m1, m2) and np.array_equal(c1, c2)
    X = np.stack(cols, axis=1)
    assert np.allclose(m1, X.mean(axis=0), rtol=1e-12, atol=1e-12)
    assert np.allclose(c1, np.cov(X, rowvar=False), rtol=1e-10, atol=1e-12)
    a = d1.ci_gaussian("cor", [0, 1], [3, 2], [[], [4, 5]])
    b = d2.ci_gaussian("cor", [0, 1], [3, 2], [[], [4, 5]])
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
    d1.release()
    d2.release()


def test_no_device_memory_growth(engine):
    """Register / test / score / resample / release in a loop: device memory in use
    settles (pooled blocks are reused), it does not grow with the iteration count."""
    import torch
    names, cols, nlev = load_discrete("alarm")
    gn, gcols = load_gaussian("gaussian_test")

    def cycle():
        d = engine.register(cols, nlev)
        d.ci_discrete("mi", [0, 1, 2], [3, 4, 5], [[], [6], [7, 8]])
        d.score_discrete("bde", [0, 1], [[2], [3, 4]], 1.0)
        r = d.resample(np.arange(0, len(cols[0]), 2, dtype=np.int32))
        r.hc("bic", 4.0, max_iter=3)
        r.release()
        d.release()
        g = engine.register(gcols)
        g.ci_gaussian("zf", [0, 1], [2, 3], [[4], []])
        g.score_gaussian("bge", [0], [[1, 2]])
        g.release()

    for _ in range(5):
        cycle()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(40):
        cycle()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 < 32 << 20, (free0, free1)


@pytest.mark.parametrize("ncols,nobs,maxlev", [(20, 100003, 4), (60, 70001, 8), (130, 33000, 6)])
def test_all_pairs_tensor_core_tables(engine, monkeypatch, ncols, nobs, maxlev):
    """The all-pairs contraction (tcgen05 indicator cross-product, csrc/bnl_pairs.cuh)
    against numpy: every two-way table of the column set in both orientations, the
    one-way tables from the diagonal, NA rows, a row count that is no multiple of
    the 128-row stage, one / two / many output tiles, next to three-column units
    that stay on the atomic kernels in the same batch. Then statistics and scores
    bit-identical to the per-unit path (same tables, same epilogue)."""
    rng = np.random.default_rng(ncols)
    na = np.iinfo(np.int32).min
    lev = [int(v) for v in rng.integers(2, maxlev + 1, ncols)]
    lev[3] = 1                                   # a one-level factor is not eligible: stays on the atomic path
    cols = []
    for i, l in enumerate(lev):
        c = rng.integers(1, l + 1, nobs).astype(np.int32)
        if i % 3 == 0:
            c[rng.random(nobs) < 0.02] = na
        cols.append(c)
    d = engine.register(cols, lev)
    xs, ys, zs = [], [], []
    for i in range(ncols):
        xs.append(i); ys.append(-1); zs.append([])                      # one-way table
        for j in range(ncols):
            if i != j and (i < j or (i + j) % 5 == 0):
                xs.append(i); ys.append(j); zs.append([])
    for k in range(12):                                                 # conditional tables in the same batch
        i, j, z = (int(v) for v in rng.choice(ncols, 3, replace=False))
        xs.append(i); ys.append(j); zs.append([z])
    monkeypatch.setenv("BNL_PAIRS", "2")
    tabs = d.counts(xs, ys, zs)
    code = [np.where(c == na, -1, c - 1).astype(np.int64) for c in cols]
    for x, y, z, t in zip(xs, ys, zs, tabs):
        keep = code[x] >= 0
        key = code[x].copy()
        shape = [lev[x]]
        if y >= 0:
            keep &= code[y] >= 0
            key = key * lev[y] + code[y]
            shape.append(lev[y])
        if z:
            keep &= code[z[0]] >= 0
            key = code[z[0]] * int(np.prod(shape)) + key
            shape.insert(0, lev[z[0]])
        want = np.bincount(key[keep], minlength=int(np.prod(shape)))
        assert np.array_equal(np.asarray(t).ravel(), want), (x, y, z)
    # the same tables from a batch without the conditional units: with >= 4096 units this
    # takes the route that builds the unit descriptors on the device
    keep_u = [k for k, z in enumerate(zs) if not z]
    again = d.counts([xs[k] for k in keep_u], [ys[k] for k in keep_u], [[]] * len(keep_u))
    for k, t in zip(keep_u, again):
        assert np.array_equal(np.asarray(t), np.asarray(tabs[k])), (xs[k], ys[k])
    if ncols >= 130:
        # a rank's share of a sharded first iteration: a block of targets against every earlier
        # column -- far from all pairs, so only the output tiles that hold a table are computed
        where = {(x, y): k for k, (x, y, z) in enumerate(zip(xs, ys, zs)) if not z}
        sub = [(i, j) for j in range(88, ncols) for i in range(j) if i != 3 and j != 3]
        assert len(sub) >= 4096
        part = d.counts([p[0] for p in sub], [p[1] for p in sub], [[]] * len(sub))
        for (i, j), t in zip(sub, part):
            assert np.array_equal(np.asarray(t), np.asarray(tabs[where[(i, j)]])), (i, j)
    pairs = [(x, y) for x, y, z in zip(xs, ys, zs) if y >= 0 and not z]
    px, py = [p[0] for p in pairs], [p[1] for p in pairs]
    got = d.ci_discrete("mi", px, py, [[]] * len(pairs)) + d.score_discrete("bde", px, [[v] for v in py], 1.0)
    monkeypatch.setenv("BNL_PAIRS", "0")
    want = d.ci_discrete("mi", px, py, [[]] * len(pairs)) + d.score_discrete("bde", px, [[v] for v in py], 1.0)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    d.release()


def test_gaussian_allpairs_matches_per_pair_tests(engine):
    """bnl_ci_gaussian_allpairs (level 0 of pc.stable on numeric data in one call, pairs
    enumerated and filtered on the device) against bnl_ci_gaussian on the same pairs:
    identical p-values, survivors in row-major order, test counter n (n - 1) / 2."""
    from bnlearn_b200.callsites import TestCounter
    rng = np.random.default_rng(12)
    n, p = 3001, 67
    base = rng.normal(size=(n, 6))
    cols = [np.ascontiguousarray(base[:, i % 6] * (0.05 * (i % 5)) + rng.normal(size=n)) for i in range(p)]
    cols[5] = np.full(n, 1.25)                    # a constant column: r = 0 by c_fast_cor's guard
    d = engine.register(cols)
    iu = np.triu_indices(p, 1)
    for test in ("cor", "zf", "mi-g"):
        st, df, pv = d.ci_gaussian(test, iu[0].astype(np.int32), iu[1].astype(np.int32), [[]] * len(iu[0]))
        counter = TestCounter()
        keep, pm = d.ci_gaussian_allpairs(test, 0.05, counter=counter)
        assert counter.value == p * (p - 1) / 2
        assert np.array_equal(pm[iu], pv) and np.array_equal(pm, pm.T)
        want = [(int(i), int(j)) for i, j, q in zip(iu[0], iu[1], pv) if q <= 0.05]
        assert [tuple(r) for r in keep.tolist()] == want
        sub = [3, 10, 11, 40, 66]
        keep2, pm2 = d.ci_gaussian_allpairs(test, 1.0, cols=sub)
        assert len(keep2) == 10 and np.array_equal(pm2, pm[np.ix_(sub, sub)])
    d.release()


def test_all_pairs_route_every_statistic_and_refusals(engine, monkeypatch):
    """Batches of >= 4096 one- and two-way tables take the route without host-side
    descriptors (device-built units, warp-per-unit epilogue, no indicator row for the
    last level of NA-free columns). Every test and score it serves must equal the
    per-unit kernels bit for bit; eight-level factors (the second half of the level
    stores), columns with and without NA, BDs (block-wide epilogue) included. What it
    cannot serve -- a bad column index -- still reaches the general path's error."""
    from bnlearn_b200 import BnlError
    rng = np.random.default_rng(21)
    na = np.iinfo(np.int32).min
    ncols, nobs = 96, 40007
    lev = [int(v) for v in rng.integers(2, 9, ncols)]
    lev[:4] = [8, 8, 2, 5]
    cols = []
    for i, l in enumerate(lev):
        c = rng.integers(1, l + 1, nobs).astype(np.int32)
        if i % 4 == 1:
            c[rng.random(nobs) < 0.05] = na
        cols.append(c)
    d = engine.register(cols, lev)
    pairs = [(i, j) for i in range(ncols) for j in range(ncols) if i != j and (i < j or (i * j) % 3 == 0)]
    assert len(pairs) >= 4096
    px = np.array([p[0] for p in pairs], dtype=np.int32)
    py = np.array([p[1] for p in pairs], dtype=np.int32)
    no_z = [[]] * len(pairs)
    one_parent = [[int(v)] for v in py]

    def everything():
        out = []
        for test in ("mi", "mi-adf", "x2", "x2-adf"):
            out += list(d.ci_discrete(test, px, py, no_z))
        for score, arg in (("loglik", 0.0), ("bic", 3.5), ("bde", 2.0), ("bds", 1.0), ("k2", 0.0), ("qnml", 0.0)):
            out += list(d.score_discrete(score, px, one_parent, arg))
        # root nodes (one-way tables) in a batch of their own
        roots = np.tile(np.arange(ncols, dtype=np.int32), 50)
        out += list(d.score_discrete("bde", roots, [[]] * len(roots), 1.0))
        return out

    monkeypatch.setenv("BNL_PAIRS", "2")
    got = everything()
    monkeypatch.setenv("BNL_PAIRS", "0")
    want = everything()
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert np.array_equal(a, b, equal_nan=True)
    monkeypatch.setenv("BNL_PAIRS", "2")
    bad = py.copy()
    bad[17] = ncols + 5
    with pytest.raises(BnlError):
        d.ci_discrete("mi", px, bad, no_z)
    d.release()


def test_all_pairs_many_binary_columns(engine, monkeypatch):
    """NA-free binary factors take ONE indicator row each, so a 256-row tile of the indicator
    matrix could hold more columns than a producer thread has slots for; the layout closes a
    128-row block after 72 columns. 320 binary columns (some with NA, some three-level ones in
    between), every pair, against numpy."""
    rng = np.random.default_rng(33)
    na = np.iinfo(np.int32).min
    ncols, nobs = 320, 20011
    lev = [3 if i % 17 == 0 else 2 for i in range(ncols)]
    cols = []
    for i, l in enumerate(lev):
        c = rng.integers(1, l + 1, nobs).astype(np.int32)
        if i % 23 == 5:
            c[rng.random(nobs) < 0.03] = na
        cols.append(c)
    d = engine.register(cols, lev)
    pairs = [(i, j) for i in range(ncols) for j in range(i + 1, ncols)]
    monkeypatch.setenv("BNL_PAIRS", "2")
    tabs = d.counts([p[0] for p in pairs], [p[1] for p in pairs], [[]] * len(pairs))
    code = np.stack([np.where(c == na, -1, c - 1) for c in cols]).astype(np.int64)
    for (x, y), t in zip(pairs, tabs):
        keep = (code[x] >= 0) & (code[y] >= 0)
        want = np.bincount((code[x] * lev[y] + code[y])[keep], minlength=lev[x] * lev[y])
        assert np.array_equal(np.asarray(t).ravel(), want), (x, y)
    d.release()
