"""bnlearn_b200 -- B200 (sm_100a) engine for bnlearn's structure-learning
statistics: contingency-table counting, discrete / Gaussian node scores and
conditional-independence tests, behind the C ABI in include/bnl_b200.h.

    engine.py     Engine / DataSet: thin ctypes face of libbnl_b200.so
    callsites.py  host-side mirror of the reference's call sites
                  (score_cache_fill, indep_test, allsubs_test, roundrobin_test,
                  per_node_score, test counter)
    learners.py   restatements of the unchanged R search loops (hc, tabu,
                  pc.stable, iamb, mmpc, si.hiton.pc, boot.strength) used to
                  check graphs and test counts end to end without R
    synth.py      seeded synthetic data of the BASELINE.json shapes
    csrc/         CUDA kernels + the C ABI (hand-written for sm_100a)
"""
from .engine import (BnlError, DataSet, Engine, SCORES, TESTS, csr)  # noqa: F401

__all__ = ["Engine", "DataSet", "BnlError", "TESTS", "SCORES", "csr"]
