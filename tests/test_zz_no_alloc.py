"""bnl_ci_discrete_noalloc: the allocation-free CI test for callers whose device
memory belongs to a framework pool (here PyTorch's caching allocator plays that
caller: the columns and the result live in torch tensors, the library gets raw
addresses and must not allocate anything itself).

The result must equal the regular entry point's (bnl_ci_discrete, which keeps its
count table in a cudaMalloc'ed buffer) within the suite's 1e-9 relative
tolerance -- on two calls in a row, because the built-in scratch table keeps the
previous call's counts and has to be cleared by each call. This file sorts last
so that a failure here cannot hide another test's result under `pytest -x`.
"""
import numpy as np
import pytest

from util import load_discrete, rel_close

pytestmark = pytest.mark.gpu


def _run_pair(engine_ds, cols_dev, nlev, nobs, out, test, x, y, z):
    import torch
    from bnlearn_b200.noalloc import ci_discrete_noalloc
    idx = [x, y] + list(z)
    stream = torch.cuda.current_stream().cuda_stream
    ci_discrete_noalloc(test, [cols_dev[i].data_ptr() for i in idx], [nlev[i] for i in idx], nobs,
                        out.data_ptr(), stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy().copy()
    st, df, pv = engine_ds.ci_discrete(test, [x], [y], [list(z)])
    assert got[1] == df[0], (test, x, y, z)
    assert rel_close(got[0], st[0], 1e-9, 1e-11), (test, x, y, z, got[0], st[0])
    assert rel_close(got[2], pv[0], 1e-9, 1e-300), (test, x, y, z, got[2], pv[0])


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    import torch
    from bnlearn_b200 import Engine
    names, cols, nlev = load_discrete("alarm")
    nobs = len(cols[0]) - 5                       # not a multiple of 16: exercises the tail rows
    cols = [c[:nobs] for c in cols]
    packed = [torch.from_numpy((c - 1).astype(np.uint8)).cuda() for c in cols]
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    with Engine() as eng:
        ds = eng.register(cols, nlev)
        # two calls in a row with DIFFERENT tables: the second must not see the first's counts.
        _run_pair(ds, packed, nlev, nobs, out, "mi", 3, 12, [33, 7])
        _run_pair(ds, packed, nlev, nobs, out, "mi", 0, 1, [])
        _run_pair(ds, packed, nlev, nobs, out, "x2-adf", 12, 13, [14, 33, 34, 35])      # 4096 cells: shared-memory path
        _run_pair(ds, packed, nlev, nobs, out, "mi-adf", 12, 13, [14, 33, 34, 35, 36])  # 16384 cells: global-atomics path
        _run_pair(ds, packed, nlev, nobs, out, "x2", 5, 4, [3])
        ds.release()


def test_noalloc_handles_missing_values():
    import torch
    from bnlearn_b200 import Engine
    rng = np.random.default_rng(4)
    n = 40001
    lev = [3, 4, 2]
    cols = [rng.integers(1, l + 1, n).astype(np.int32) for l in lev]
    for c in cols:
        c[rng.random(n) < 0.05] = np.iinfo(np.int32).min
    packed = [torch.from_numpy(np.where(c == np.iinfo(np.int32).min, 255, c - 1).astype(np.uint8)).cuda()
              for c in cols]
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    with Engine() as eng:
        ds = eng.register(cols, lev)
        _run_pair(ds, packed, lev, n, out, "mi", 0, 1, [2])
        _run_pair(ds, packed, lev, n, out, "x2", 0, 1, [2])
        ds.release()


def test_noalloc_rejects_tables_beyond_the_builtin_scratch():
    import torch
    from bnlearn_b200 import BnlError
    from bnlearn_b200.noalloc import MAX_CELLS, ci_discrete_noalloc
    col = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    levels = [16] * 5                                  # 16^5 cells > 65536
    assert 16 ** 5 > MAX_CELLS
    with pytest.raises(BnlError) as ei:
        ci_discrete_noalloc("mi", [col.data_ptr()] * 5, levels, 64, out.data_ptr())
    assert ei.value.code == 5 and "scratch" in str(ei.value)
