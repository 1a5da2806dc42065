import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "ref_outputs.json")) as f:
        out = json.load(f)
    # learning.test with missing cells: the reference's *_with_missing arms and NAL / PNAL
    with open(os.path.join(ROOT, "tests", "golden", "ref_incomplete.json")) as f:
        out.update(json.load(f))
    # Jonckheere-Terpstra and shrinkage MI on complete data ("data" names the data set)
    with open(os.path.join(ROOT, "tests", "golden", "ref_jt_shrinkage.json")) as f:
        out.update(json.load(f))
    with open(os.path.join(ROOT, "tests", "golden", "ref_nml.json")) as f:          # fNML / qNML scores
        out.update(json.load(f))
    return out
