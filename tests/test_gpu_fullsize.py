"""Parity at the configurations' REAL sizes (BASELINE.json configs B, C, D, E):
the launches the benchmark times -- row-pair tables, long row chunks, the hybrid
upload, candidate groups, every base-index counter class, the Gram of 2000 x 1 M
and the t tail at df ~ 10^6 -- against the reference's own C (oracle/_ref) on a
stratified sample of their units. Counts bit-exact; statistics, p-values and
scores within 1e-9 relative; graphs and test counters identical.

The reference does one 10 M-row test in ~20-60 ms per core, so a sample of a few
dozen units per configuration keeps the whole file to a couple of minutes.
Configs C and D are generated on the device (bnlearn_b200/synth.py) and copied
to the host for the reference."""
import numpy as np
import pytest

from util import load_gaussian, rel_close

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def engine():
    from bnlearn_b200 import Engine
    with Engine() as e:
        yield e


@pytest.fixture(scope="module")
def ref():
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libbnref.so is not built (needs the reference sources at build time)")
    from oracle import refapi
    return refapi


def _close(a, b, what, atol=0.0):
    assert rel_close(a, b, RTOL, atol), "%s: got %r, reference %r" % (what, a, b)


# --------------------------------------------------------------------------
# config B: pc.stable + mi, 37 factors x 10 M rows

@pytest.fixture(scope="module")
def config_b(engine):
    from bnlearn_b200 import synth
    names, codes, lev = synth.alarm_shaped_cached(10_000_000)
    cols = [np.ascontiguousarray(c.astype(np.int32)) for c in codes]
    d = engine.register(cols, lev)
    yield names, cols, lev, d
    d.release()


def _ref_ci(ref, cols, lev, x, y, z, test_e=1):
    from oracle import refbind as rb
    if z:
        cfg, llz = ref.fast_config([cols[j] for j in z], [lev[j] for j in z], 1)
        st, df = ref.chisqtest(cols[x], lev[x], cols[y], lev[y], test_e, cfg, llz)
    else:
        cfg, llz = None, 1
        st, df = ref.chisqtest(cols[x], lev[x], cols[y], lev[y], test_e)
    return st, df, rb.lib().Rf_pchisq(st, df, 0, 0), cfg, llz


def test_config_b_bench_batch_against_reference(engine, ref, config_b):
    """The 1081-test batch of bench.py in ONE call (so every unit runs in the mode
    the benchmark uses: row pairs for the marginal tests, chunk-major CTAs), 36 of
    its tests against c_fast_config + c_cchisqtest + pchisq
    (src/tests/discrete/discrete.tests.c:11-94), 12 per conditioning-set size."""
    from bnlearn_b200 import workloads
    names, cols, lev, d = config_b
    units = workloads.pc_stable_units(names, lev, max_level=2)
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    st0, df0, pv0 = d.ci_discrete("mi", xs, ys, zs)
    # from a data set's second counting call on the conditional tests read the nibble copy of the
    # columns (the route bench.py times): the same bits either way, and the reference's numbers
    st, df, pv = d.ci_discrete("mi", xs, ys, zs)
    assert np.array_equal(st, st0) and np.array_equal(df, df0) and np.array_equal(pv, pv0)
    rng = np.random.default_rng(20261106)
    picked = []
    for k in (0, 1, 2):
        idx = [i for i, u in enumerate(units) if len(u[2]) == k]
        picked += [int(i) for i in rng.choice(idx, size=12, replace=False)]
    for i in picked:
        x, y, z = units[i]
        rs, rdf, rp, _, _ = _ref_ci(ref, cols, lev, x, y, z)
        assert df[i] == rdf, units[i]
        _close(st[i], rs, ("stat", units[i]), atol=1e-9)
        _close(pv[i], rp, ("p", units[i]), atol=1e-300)


def test_config_b_tables_and_larger_tests_against_reference(engine, ref, config_b):
    """Tables bit-exact against fill_3d_table (src/core/contingency.tables.c:198-223)
    for every size class at 10 M rows: row pairs (<= 16 cells), byte-packed
    (<= 256), 16-bit packed with 64 and with 32 lane counters (768 and 1024 cells),
    and mi / x2 / mi-adf on the same units."""
    names, cols, lev, d = config_b
    four = [i for i, l in enumerate(lev) if l == 4]
    three = [i for i, l in enumerate(lev) if l == 3]
    two = [i for i, l in enumerate(lev) if l == 2]
    units = [(two[0], two[1], []), (three[0], four[0], []), (four[1], four[2], [three[1]]),
             (three[2], three[3], [four[3], two[2]]), (four[0], four[1], [four[2], four[3], three[0]]),
             (four[4], four[5], [four[6], four[0], four[1]]), (two[3], four[2], [three[4], three[5]])]
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    tabs = d.counts(xs, ys, zs)
    refs = []
    for (x, y, z), t in zip(units, tabs):
        _, _, _, cfg, llz = _ref_ci(ref, cols, lev, x, y, z)
        rt, nobs = ref.tables3d(cols[x], lev[x], cols[y], lev[y], cfg, llz)
        assert nobs == 10_000_000
        assert np.array_equal(np.asarray(t).reshape(rt.shape), rt), (x, y, z)
        refs.append(rt)
    for test, code in (("mi", 1), ("mi-adf", 2), ("x2", 3)):
        st, df, pv = d.ci_discrete(test, xs, ys, zs)
        for i, (x, y, z) in enumerate(units):
            rs, rdf, rp, _, _ = _ref_ci(ref, cols, lev, x, y, z, code)
            assert df[i] == rdf, (test, units[i])
            _close(st[i], rs, (test, "stat", units[i]), atol=1e-9)
            _close(pv[i], rp, (test, "p", units[i]), atol=1e-300)


# --------------------------------------------------------------------------
# config C: hc + BDe, 1000 factors x 1 M rows

@pytest.fixture(scope="module")
def config_c(engine):
    from bnlearn_b200 import synth
    host, lev, parents = synth.random_discrete_bn_device(1000, 1_000_000)
    cols = [host[i] for i in range(host.shape[0])]
    d = engine.register(cols, lev)
    yield cols, lev, parents, d
    d.release()


def _ref_bde(ref, cols, lev, target, pa):
    idx = [target] + list(pa)
    names = ["V%d" % i for i in idx]
    data = ref.RefData([cols[i] for i in idx], [lev[i] for i in idx], names)
    par = {nm: [] for nm in names}
    par[names[0]] = names[1:]
    return float(ref.per_node_score(data, par, [names[0]], "bde", iss=1.0)[0])


def _pick_levels(lev, want, avoid):
    out = []
    for l in want:
        i = next(i for i in range(len(lev)) if lev[i] == l and i not in avoid and i not in out)
        out.append(i)
    return out


@pytest.mark.parametrize("base_levels", [
    [],                 # two-column units: per-unit kernel, row-pair tables
    [3],                # candidate groups (count_group_kernel), three columns
    [4, 3],             # candidate groups, four columns
    [4, 3, 3],          # base-index kernel, 32 counters per cell
    [4, 4, 3, 3],       # 32 / 16 counters per cell in one launch
    [4, 4, 4, 3, 3],    # 16 / 8 counters per cell
    [4, 4, 4, 4, 4],    # largest tables past the base-index classes: generic kernel
])
def test_config_c_candidate_batches_against_reference(engine, ref, config_c, base_levels):
    """One score_cache_fill-shaped batch per kernel class -- every other node as the
    candidate parent of one target, as hc_opt_step's iteration does -- and 9
    candidates of it (three of each level count, so every table size of the launch)
    against per_node_score("bde") (src/scores/per.node.score.c:31-48 ->
    dirichlet_node, src/scores/dirichlet.posterior.c:152-202)."""
    cols, lev, parents, d = config_c
    target = 500
    base = _pick_levels(lev, base_levels, {target})
    cands = [i for i in range(len(cols)) if i != target and i not in base]
    sc, _ = d.score_discrete("bde", [target] * len(cands), [base + [c] for c in cands], 1.0)
    rng = np.random.default_rng(len(base_levels))
    for l in (2, 3, 4):
        pos = [k for k, c in enumerate(cands) if lev[c] == l]
        for k in rng.choice(pos, size=3, replace=False):
            want = _ref_bde(ref, cols, lev, target, base + [cands[int(k)]])
            _close(sc[int(k)], want, ("bde", base_levels, cands[int(k)]))


def test_config_c_first_iteration_against_reference(engine, ref, config_c):
    """The first hc iteration at full size -- every arc i -> j, i < j, of the empty graph,
    499 500 BDe scores from ONE tensor-core contraction (pair_onehot_kernel, unit descriptors
    built on the device, warp-per-unit epilogue) -- and 30 of them, spread over the level
    counts, against per_node_score("bde"); then the same through bnl_score_cache_fill with its
    mirrored copies and counter (src/learning/score/hc.cache.lookup.c:52-117)."""
    from bnlearn_b200 import callsites as cs
    cols, lev, parents, d = config_c
    n = len(cols)
    tg = np.array([j for j in range(n) for i in range(j)], dtype=np.int32)
    pa = np.array([i for j in range(n) for i in range(j)], dtype=np.int32)
    sc, npar = d.score_discrete("bde", tg, (np.arange(len(tg) + 1, dtype=np.int32), pa), 1.0)
    rng = np.random.default_rng(41)
    picked = [int(k) for k in rng.choice(len(tg), size=30, replace=False)]
    for k in picked:
        want = _ref_bde(ref, cols, lev, int(tg[k]), [int(pa[k])])
        _close(sc[k], want, ("bde", int(tg[k]), int(pa[k])))
        assert npar[k] == (lev[tg[k]] - 1) * lev[pa[k]]
    # the call site: deltas against the empty graph's node scores, both orientations filled
    reference = d.score_discrete("bde", list(range(n)), [[]] * n, 1.0)[0]
    amat = np.zeros((n, n), dtype=np.int32, order="F")
    cache = np.zeros((n, n), order="F")
    counter = cs.TestCounter()
    cs.score_cache_fill(d, "bde", 1.0, amat, amat.copy(), list(range(n)), True, reference, cache, counter)
    assert counter.value == n * (n - 1) // 2
    for k in picked:
        i, j = int(pa[k]), int(tg[k])
        want = sc[k] - reference[j]
        assert rel_close(cache[i, j], want, RTOL, 1e-6) and cache[j, i] == cache[i, j]


def test_config_c_score_cache_fill_block_against_reference(engine, ref, config_c):
    """score_cache_fill (src/learning/score/hc.cache.lookup.c:10-128) on the first
    64 nodes of the 1 M-row data set, four updated nodes with 0-3 parents: cache
    columns and the test counter against the reference's own entry point."""
    from bnlearn_b200 import callsites as cs
    from oracle import refbind as rb
    cols, lev, parents, d = config_c
    n = 64
    amat = np.zeros((n, n), dtype=np.int32, order="F")
    for j in (17, 40, 63):
        for i in [p for p in parents[j] if p < n][:3]:
            amat[i, j] = 1
    amat[3, 40] = amat[9, 40] = 1
    amat[20, 63] = 1
    par = [[int(i) for i in np.flatnonzero(amat[:, j])] for j in range(n)]
    reference = d.score_discrete("bde", list(range(n)), par, 1.0)[0]
    updated = [5, 17, 40, 63]
    blmat = np.zeros((n, n), dtype=np.int32, order="F")
    counter = cs.TestCounter()
    cache = np.zeros((n, n), order="F")
    cs.score_cache_fill(d, "bde", 1.0, amat, blmat, updated, False, reference, cache, counter)

    names = ["V%d" % i for i in range(n)]
    data = ref.RefData(cols[:n], lev[:n], names)
    rpar = {names[j]: [names[i] for i in par[j]] for j in range(n)}
    want_ref = ref.per_node_score(data, rpar, names, "bde", iss=1.0)
    for j in range(n):
        _close(reference[j], want_ref[j], ("node score", j))
    rb.reset_test_counter()
    want = np.zeros((n, n), order="F")
    ref.score_cache_fill(data, rpar, "bde", want_ref, False, updated, amat, want, blmat, iss=1.0)
    assert counter.value == rb.test_counter()
    for j in updated:
        for i in range(n):
            assert rel_close(cache[i, j], want[i, j], RTOL, 1e-6), (i, j, cache[i, j], want[i, j])


# --------------------------------------------------------------------------
# config D: pc.stable + cor / zf, 2000 numeric columns x 1 M rows

@pytest.fixture(scope="module")
def config_d(engine):
    from bnlearn_b200 import synth
    host, parents = synth.random_gaussian_bn_device(2000, 1_000_000)
    cols = [host[i] for i in range(host.shape[0])]
    d = engine.register(cols)
    yield cols, parents, d
    d.release()


def test_config_d_gram_against_long_double_covariance(engine, config_d):
    """64 entries of the Gram matrix of 2000 x 1 M against the two-pass covariance
    with long double accumulators that c_covmat computes
    (src/core/covariance.matrix.c:65-97; c_meanvec, src/core/moments.c:7-30)."""
    cols, parents, d = config_d
    mean, cov = d.moments()
    rng = np.random.default_rng(8)
    n = len(cols[0])
    pairs = [(int(a), int(b)) for a, b in rng.integers(0, 2000, size=(48, 2))]
    pairs += [(j, p) for j in (1500, 1900, 1999, 777) for p in parents[j]][:12]
    pairs += [(i, i) for i in (0, 999, 1999, 1234)]
    for a, b in pairs:
        ma = cols[a].sum(dtype=np.longdouble) / n
        mb = cols[b].sum(dtype=np.longdouble) / n
        want = float(((cols[a].astype(np.longdouble) - ma) * (cols[b].astype(np.longdouble) - mb)).sum() / (n - 1))
        _close(mean[a], float(ma), ("mean", a), atol=1e-15)
        scale = float(np.sqrt(cov[a, a] * cov[b, b]))
        assert abs(cov[a, b] - want) <= RTOL * max(abs(want), 1e-3 * scale), (a, b, cov[a, b], want)


def test_config_d_tests_against_reference(engine, ref, config_d):
    """64 cor / zf tests with 0-3 conditioning variables at 1 M rows against the
    reference's indep_test -> ct_gaustests_complete / ut_gaustests_complete
    (src/tests/patterns/ctest.c:58-142, utest.c:65-129): partial correlation by
    SVD pseudo-inverse, the t tail at df ~ 10^6 and the normal tail of Fisher's Z."""
    cols, parents, d = config_d
    rng = np.random.default_rng(9)
    units = []
    kids = [j for j in range(200, 2000) if len(parents[j]) >= 2]
    for k in range(32):
        j = kids[int(rng.integers(0, len(kids)))]
        s = k % 4
        if k % 2 == 0:      # child vs parent given other parents / random columns: strong dependence
            pa = list(parents[j])
            z = (pa[1:] + [int(v) for v in rng.choice(200, size=3, replace=False)])[:s]
            units.append((j, pa[0], z))
        else:               # unrelated columns: p-values spread over (0, 1)
            idx = [int(v) for v in rng.choice(2000, size=2 + s, replace=False)]
            units.append((idx[0], idx[1], idx[2:]))
    for test in ("cor", "zf"):
        st, df, pv = d.ci_gaussian(test, [u[0] for u in units], [u[1] for u in units], [u[2] for u in units])
        for i, (x, y, z) in enumerate(units):
            idx = [x, y] + z
            names = ["G%d" % c for c in idx]
            data = ref.RefData([cols[c] for c in idx], None, names)
            rs, rdf, rp = ref.ci_test(data, names[0], names[1], names[2:], test)
            if rdf is not None:                      # zf's htest carries no df
                assert df[i] == rdf, (test, units[i])
            _close(st[i], rs, (test, "stat", units[i]), atol=1e-12)
            _close(pv[i], rp, (test, "p", units[i]), atol=1e-300)


# --------------------------------------------------------------------------
# config E: boot.strength, insurance-shaped 27 x 20 000

def test_config_e_replicates_identical_to_reference_hc(engine, ref):
    """Replicates 0-3 of boot_strength (device-side resampling + bnl_hc) against hc
    driven by the reference's score_cache_fill / per_node_score on the same
    resampled rows (R/arc.strength.R:144-165): identical graphs and counters."""
    from bnlearn_b200 import learners, synth
    from bnlearn_b200.callsites import TestCounter
    from refbackend import RefBackend
    names, cols, lev = synth.insurance_shaped(20000)
    n, nobs, seed = len(cols), 20000, 20261109
    src = engine.register(cols, lev)
    acc = np.zeros((n, n))
    for r in range(4):
        rows = np.random.Generator(np.random.PCG64(seed + r)).integers(0, nobs, nobs)
        ds = src.resample(rows)
        counter = TestCounter()
        amat, _, _ = ds.hc("bde", 1.0, counter=counter)
        ds.release()
        rbk = RefBackend([c[rows] for c in cols], lev)
        want, _ = learners.hc(rbk, n, "bde", 1.0)
        assert np.array_equal(np.asarray(amat), np.asarray(want)), r
        assert counter.value == rbk.counter.value, r
        acc += np.asarray(want)
    src.release()
    strength, direction, _ = learners.boot_strength(engine, cols, lev, 4, "bde", 1.0, seed=seed)
    prob = acc / 4
    assert np.array_equal(strength, prob + prob.T)


# --------------------------------------------------------------------------

def test_iamb_on_gaussian_data_identical_to_reference(engine, ref):
    """iamb on gaussian.test with cor: every conditional test and the backward
    roundrobin_test go through rrd_gaustests_complete / ct_gaustests_complete
    (src/tests/patterns/roundrobin.test.c:106-219) on complete data."""
    from bnlearn_b200 import learners
    from bnlearn_b200.learners import GpuBackend
    from refbackend import RefBackend
    names, cols = load_gaussian("gaussian_test")
    d = engine.register(cols)
    gpu, rbk = GpuBackend(d), RefBackend(cols)
    for algo in ("iamb", "mmpc"):
        gpu.counter.reset(); rbk.counter.reset()
        n1, m1 = learners.local_discovery(gpu, len(cols), algo, "cor", 0.05)
        n2, m2 = learners.local_discovery(rbk, len(cols), algo, "cor", 0.05)
        assert n1 == n2 and m1 == m2
        assert gpu.counter.value == rbk.counter.value
    d.release()


def test_gaussian_permutation_test_past_four_million_rows(engine, ref):
    """One permutation no longer fits the 8 M-index batch twice: the first batch
    still holds the identity next to the first drawn permutation. Against
    c_gauss_mcarlo (src/tests/permutation/gaussian.monte.carlo.c:24-129) on the
    same uniform stream."""
    rng = np.random.default_rng(4)
    n = 4_300_000
    x = rng.normal(size=n)
    y = 0.0005 * x + rng.normal(size=n)
    d = engine.register([x, y])
    st, pv = d.ci_gaussian_mc("mc-cor", [0], [1], [[]], 6, 1.0, seed=3)
    rs, rp = ref.mc_gaussian("mc-cor", [x, y], B=6, alpha=1.0, seed=3)
    _close(st[0], rs, "statistic", atol=1e-12)
    assert pv[0] == rp
    d.release()
