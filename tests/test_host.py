"""CPU tests of the product's host-side pieces: the C-ABI library loads and
exports every symbol the header declares; the fp64 special functions (host
build of the very code the kernels run) match mpmath fixtures."""
import json
import os
import re

import pytest

from util import GOLDEN, rel_close

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from bnlearn_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "bnl_b200.h")).read()
    declared = set(re.findall(r"\b(bnl_[a-z_0-9]+)\s*\(", header))
    assert declared, "no declarations found in the header"
    for name in declared:
        assert hasattr(lib, name), "libbnl_b200.so does not export %s" % name
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from bnlearn_b200 import BnlError, Engine
    with pytest.raises(BnlError):
        Engine()


def test_device_special_functions_host_build_against_mpmath():
    from bnlearn_b200 import _lib
    L = _lib.load_special_host()
    with open(os.path.join(GOLDEN, "special_functions.json")) as f:
        sf = json.load(f)
    for x, v in sf["lgamma"]:
        assert rel_close(L.bnl_host_lgamma(x), v, 1e-12, 1e-14), x
    for x, df, v in sf["pchisq_upper"]:
        assert rel_close(L.bnl_host_pchisq_upper(x, df), v, 1e-12), (x, df)
    for t, df, v in sf["pt_upper2"]:
        assert rel_close(L.bnl_host_pt_two_sided(t, df), v, 1e-12), (t, df)
    for z, v in sf["pnorm_upper2"]:
        assert rel_close(L.bnl_host_pnorm_two_sided(z), v, 1e-12), z
    assert L.bnl_host_pchisq_upper(0.0, 0.0) == 1.0      # df = 0: point mass at zero
    assert L.bnl_host_pchisq_upper(3.0, 0.0) == 0.0


def test_cost_blocks_cover_the_list_in_order():
    """replicate.cost_blocks: contiguous, ordered, complete; equal costs give near-equal
    blocks; a leading run that must stay together goes to rank 0 once it costs a rank's share."""
    from bnlearn_b200.replicate import cost_blocks
    b = cost_blocks([1.0] * 103, 4)
    assert b[0] == 0 and b[-1] == 103 and all(x <= y for x, y in zip(b, b[1:]))
    assert max(y - x for x, y in zip(b, b[1:])) - min(y - x for x, y in zip(b, b[1:])) <= 1
    costs = [0.125] * 666 + [1.0] * 415
    assert cost_blocks(costs, 1) == [0, 1081]
    b8 = cost_blocks(costs, 8, together=666)
    assert b8[1] == 666 and b8[-1] == 1081 and len(b8) == 9
    b2 = cost_blocks(costs, 2, together=666)
    assert b2[0] == 0 and b2[1] > 666 and b2[2] == 1081        # cheaper than a rank's share: shares a rank
    assert cost_blocks([], 3) == [0, 0, 0, 0]


def test_search_core_opt_step_against_plain_scan(tmp_path):
    """Search::opt_step (csrc/bnl_search_core.h: row bounds over valid and acyclic additions,
    reachability closure, the exact cut below the best acceptable addition) gives the move of a
    plain scan in the reference's order (hc_opt_step, src/learning/score/hc.cache.lookup.c:131-363)
    on 32 000 steps of random DAG walks (hill climbing and tabu search), with exact ties and values a fraction
    of sqrt(eps) apart."""
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "search_core_test")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(root, "bnlearn_b200", "csrc"), "-o", exe,
                    os.path.join(root, "tests", "search_core_test.cpp")], check=True)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.startswith("ok:")
