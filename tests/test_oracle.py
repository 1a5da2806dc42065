"""CPU tests: the C restatement (oracle/bnl_oracle.c) against the golden outputs
of the reference's own C code (tests/golden/ref_outputs.json) and, where the
compiled reference is present, against oracle/_ref directly. Tolerance on
floating point: 1e-9 relative (north_star); counts are exact."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as orc
DISCRETE_SETS = ["learning_test", "alarm", "learning_test_na", "learning_test_jt_sh", "alarm_jt_sh",
                 "learning_test_nml", "alarm_nml", "alarm_600_nml"]
from util import GOLDEN, load_discrete, load_gaussian, rel_close

RTOL = 1e-9


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_discrete_ci_tests_match_reference(golden, ds):
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    for rec in golden[ds]["tests"]:
        st, df, p = orc.ci_discrete(rec["test"], cols[rec["x"]], nlev[rec["x"]], cols[rec["y"]],
                                    nlev[rec["y"]], [cols[i] for i in rec["z"]],
                                    [nlev[i] for i in rec["z"]])
        assert df == rec["df"], rec
        assert rel_close(st, rec["stat"], RTOL, 1e-12), (rec, st)
        assert rel_close(p, rec["p"], RTOL, 1e-300), (rec, p)


@pytest.mark.parametrize("ds", DISCRETE_SETS)
def test_discrete_scores_match_reference(golden, ds):
    names, cols, nlev = load_discrete(golden[ds].get("data", ds))
    for rec in golden[ds]["scores"]:
        v, _ = orc.score_discrete(rec["score"], cols[rec["node"]], nlev[rec["node"]],
                                  [cols[i] for i in rec["parents"]],
                                  [nlev[i] for i in rec["parents"]], rec["arg"])
        assert rel_close(v, rec["value"], RTOL), (rec, v)


GAUSSIAN_SETS = ["gaussian_test", "gaussian_test_na"]


@pytest.mark.parametrize("ds", GAUSSIAN_SETS)
def test_gaussian_ci_tests_match_reference(golden, ds):
    names, cols = load_gaussian(ds)
    for rec in golden[ds]["tests"]:
        st, df, p, _ = orc.ci_gaussian(rec["test"], [cols[rec["x"]], cols[rec["y"]]] +
                                       [cols[i] for i in rec["z"]])
        assert df == rec["df"]
        assert rel_close(st, rec["stat"], RTOL, 1e-13), (rec, st)
        assert rel_close(p, rec["p"], RTOL, 1e-300), (rec, p)


@pytest.mark.parametrize("ds", GAUSSIAN_SETS)
def test_gaussian_scores_match_reference(golden, ds):
    names, cols = load_gaussian(ds)
    for rec in golden[ds]["scores"]:
        v, _ = orc.score_gaussian(rec["score"], cols[rec["node"]],
                                  [cols[i] for i in rec["parents"]], rec["arg"])
        assert rel_close(v, rec["value"], RTOL), (rec, v)


def test_conditional_gaussian_scores_match_reference():
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_cg.json")) as f:
        recs = json.load(f)["clgaussian_test"]["scores"]
    names, cols, nlev = load_mixed("clgaussian_test")
    for r in recs:
        v, _ = orc.score_cg(r["score"], cols[r["node"]], nlev[r["node"]], [cols[i] for i in r["parents"]],
                            [nlev[i] for i in r["parents"]], r["arg"])
        assert rel_close(v, r["value"], RTOL), (r, v)


def test_conditional_gaussian_tests_match_reference():
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_cg.json")) as f:
        recs = json.load(f)["clgaussian_test"]["tests"]
    names, cols, nlev = load_mixed("clgaussian_test")
    for r in recs:
        st, df, p = orc.ci_cg(cols[r["x"]], nlev[r["x"]], cols[r["y"]], nlev[r["y"]],
                              [cols[i] for i in r["z"]], [nlev[i] for i in r["z"]])
        assert df == r["df"], r
        assert rel_close(st, r["stat"], RTOL, 1e-9), (r, st)
        assert rel_close(p, r["p"], 1e-8, 1e-300), (r, p)


def test_ebic_matches_reference():
    """ebic = bic - 4 gamma #parents log(ncol) on top of the three bic flavours."""
    from util import load_mixed
    with open(os.path.join(GOLDEN, "ref_ebic.json")) as f:
        g = json.load(f)
    for ds, recs in g.items():
        if ds == "learning_test":
            names, cols, nlev = load_discrete(ds)
        elif ds == "gaussian_test":
            names, cols = load_gaussian(ds)
            nlev = [0] * len(cols)
        else:
            names, cols, nlev = load_mixed(ds)
        for r in recs:
            node, pa = r["node"], r["parents"]
            v = orc.score_cg("bic-cg", cols[node], nlev[node], [cols[i] for i in pa], [nlev[i] for i in pa],
                             r["arg"])[0]
            v -= 4.0 * r["gamma"] * len(pa) * np.log(len(cols))
            assert rel_close(v, r["value"], RTOL), (ds, r, v)


def test_bge_matches_reference():
    with open(os.path.join(GOLDEN, "ref_bge.json")) as f:
        recs = json.load(f)["gaussian_test"]
    names, cols = load_gaussian("gaussian_test")
    for r in recs:
        v = orc.score_bge(cols[r["node"]], [cols[i] for i in r["parents"]], len(names), r["iss_mu"],
                          r["iss_w"], [r["nu"][r["node"]]] + [r["nu"][i] for i in r["parents"]])
        assert rel_close(v, r["value"], RTOL), (r["node"], r["parents"], v, r["value"])


def test_counts_are_bincounts():
    names, cols, nlev = load_discrete("alarm")
    x, y, z1, z2 = 3, 12, 33, 7
    tab = orc.counts(cols[x], nlev[x], cols[y], nlev[y], [cols[z1], cols[z2]], [nlev[z1], nlev[z2]])
    cfg = (cols[z1] - 1) + nlev[z1] * (cols[z2] - 1)
    key = (cfg.astype(np.int64) * nlev[x] + (cols[x] - 1)) * nlev[y] + (cols[y] - 1)
    expect = np.bincount(key, minlength=tab.size).reshape(tab.shape)
    assert np.array_equal(tab, expect)
    assert tab.sum() == len(cols[x])


def test_na_rows_are_skipped():
    rng = np.random.default_rng(1)
    n = 5000
    x = rng.integers(1, 4, n).astype(np.int32)
    y = rng.integers(1, 3, n).astype(np.int32)
    z = rng.integers(1, 5, n).astype(np.int32)
    x[::7] = orc.NA_INT
    z[::11] = orc.NA_INT
    keep = (x != orc.NA_INT) & (z != orc.NA_INT)
    a = orc.ci_discrete("mi", x, 3, y, 2, [z], [4])
    b = orc.ci_discrete("mi", x[keep], 3, y[keep], 2, [z[keep]], [4])
    assert a == b
    cfg, nl = orc.config([x, z], [3, 4], 1)
    assert nl == 12 and np.all(cfg[~keep] == orc.NA_INT)


def test_adf_returns_independence_below_five_per_cell():
    rng = np.random.default_rng(2)
    n = 40
    x = rng.integers(1, 4, n).astype(np.int32)
    y = rng.integers(1, 4, n).astype(np.int32)
    st, df, p = orc.ci_discrete("mi-adf", x, 3, y, 3)
    assert st == 0.0 and p == 1.0 and df <= 4


def test_empty_input():
    e = np.zeros(0, dtype=np.int32)
    st, df, p = orc.ci_discrete("mi", e, 3, e, 2)
    assert st == 0.0 and df == 2.0 and p == 1.0


def test_config_overflow_is_reported():
    e = np.ones(4, dtype=np.int32)
    with pytest.raises(OverflowError):
        orc.ci_discrete("mi", e, 2, e, 2, [e] * 6, [64] * 6)


def test_special_functions_against_mpmath():
    with open(os.path.join(GOLDEN, "special_functions.json")) as f:
        sf = json.load(f)
    L = orc.lib()
    for x, v in sf["lgamma"]:
        assert rel_close(L.Rf_lgammafn(x), v, 1e-13, 1e-15), x
    for x, df, v in sf["pchisq_upper"]:
        assert rel_close(L.Rf_pchisq(x, df, 0, 0), v, 1e-12), (x, df)
    for t, df, v in sf["pt_upper2"]:
        assert rel_close(2 * L.Rf_pt(t, df, 0, 0), v, 1e-12), (t, df)
    for z, v in sf["pnorm_upper2"]:
        assert rel_close(2 * L.Rf_pnorm5(z, 0, 1, 0, 0), v, 1e-13), z


def test_bde_score_equivalence():
    """man/score.Rd:169-171: reversing a covered arc leaves BDe unchanged."""
    names, cols, nlev = load_discrete("learning_test")
    a, b = 0, 1
    s1 = orc.score_discrete("bde", cols[a], nlev[a], [], [], 1.0)[0] + \
        orc.score_discrete("bde", cols[b], nlev[b], [cols[a]], [nlev[a]], 1.0)[0]
    s2 = orc.score_discrete("bde", cols[b], nlev[b], [], [], 1.0)[0] + \
        orc.score_discrete("bde", cols[a], nlev[a], [cols[b]], [nlev[b]], 1.0)[0]
    assert abs(s1 - s2) < 1e-8


def test_bic_matches_hand_formula():
    """man/score.Rd:188-205: sum(counts * log(prop.table(counts, 2))) - nparams * log(n) / 2."""
    names, cols, nlev = load_discrete("learning_test")
    node, pa = 3, [0, 2]
    n = len(cols[0])
    cfg = (cols[pa[0]] - 1) + nlev[pa[0]] * (cols[pa[1]] - 1)
    tab = np.zeros((nlev[node], nlev[pa[0]] * nlev[pa[1]]))
    np.add.at(tab, (cols[node] - 1, cfg), 1)
    prop = tab / tab.sum(axis=0, keepdims=True)
    ll = np.sum(tab[tab > 0] * np.log(prop[tab > 0]))
    nparams = (nlev[node] - 1) * tab.shape[1]
    v, npar = orc.score_discrete("bic", cols[node], nlev[node], [cols[i] for i in pa],
                                 [nlev[i] for i in pa], np.log(n) / 2)
    assert npar == nparams
    assert rel_close(v, ll - nparams * np.log(n) / 2, 1e-12)
