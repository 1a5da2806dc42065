"""The C ABI from C: tests/c_abi_client.c is compiled with gcc against
include/bnl_b200.h and libbnl_b200.so, run, and its output compared with the
oracle on the same (LCG-generated) data."""
import os
import subprocess

import numpy as np
import pytest

from util import rel_close

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lcg_data(nobs, seed):
    s = seed
    def nxt():
        nonlocal s
        s = (s * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
        return (s >> 33) & 0xFFFFFFFF
    cols = [np.zeros(nobs, np.int32) for _ in range(4)]
    for i in range(nobs):
        cols[0][i] = 1 + nxt() % 3
        cols[1][i] = 1 + (cols[0][i] + nxt() % 2) % 2
        cols[2][i] = 1 + nxt() % 4
        r = nxt()
        cols[3][i] = np.iinfo(np.int32).min if r % 50 == 0 else 1 + (cols[2][i] + nxt() % 3) % 3
        nxt()
    return cols


def test_c_client(tmp_path):
    from oracle import oracle as orc
    exe = str(tmp_path / "c_abi_client")
    libdir = os.path.join(ROOT, "bnlearn_b200")
    subprocess.check_call(["gcc", "-O1", "-std=c11", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c_abi_client.c"), "-o", exe,
                           "-L", libdir, "-l:libbnl_b200.so", "-Wl,-rpath," + libdir])
    nobs, seed = 5003, 42
    out = subprocess.check_output([exe, str(nobs), str(seed)], text=True)
    lines = [l.split() for l in out.strip().splitlines()]
    cols = _lcg_data(nobs, seed)
    lev = [3, 2, 4, 3]
    units = [(0, 1, []), (2, 3, [0]), (0, 3, [1, 2])]
    for i, (x, y, z) in enumerate(units):
        st, df, pv = orc.ci_discrete("mi", cols[x], lev[x], cols[y], lev[y], [cols[j] for j in z], [lev[j] for j in z])
        got = [float(v) for v in lines[i][2:5]]
        assert got[1] == df and rel_close(got[0], st, 1e-9, 1e-11) and rel_close(got[2], pv, 1e-9, 1e-300)
    # BDe with NA rows skipped (the reference refuses NA for BDe; here they are dropped).
    s0 = orc.score_discrete("bde", cols[1], 2, [cols[0]], [3], 1.0)[0]
    assert rel_close(float(lines[3][2]), s0, 1e-9)
    assert lines[5][0] == "indep" and lines[5][4] == "2"
    for k, x in enumerate((0, 2)):
        pv = orc.ci_discrete("x2", cols[x], lev[x], cols[1], 2)[2]
        assert rel_close(float(lines[5][1 + k]), pv, 1e-9, 1e-300)
    assert lines[6][0] == "error" and "column" in out
