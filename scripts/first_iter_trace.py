"""Host-side time breakdown of the first score_cache_fill of hc on config C (BNL_TRACE=1)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("BNL_TRACE", "1")
from bnlearn_b200 import Engine, synth, callsites
n = 1000
host, lev, _ = synth.random_discrete_bn_device(n, 1_000_000)
with Engine() as eng:
    d = eng.register([host[i] for i in range(n)], lev)
    amat = np.zeros((n, n), dtype=np.int32, order="F")
    bl = np.zeros((n, n), dtype=np.int32, order="F")
    ref, _ = d.score_discrete("bde", list(range(n)), [[] for _ in range(n)], 1.0)
    cache = np.zeros((n, n), order="F")
    for rep in range(4):
        cnt = callsites.TestCounter()
        t0 = time.perf_counter()
        callsites.score_cache_fill(d, "bde", 1.0, amat, bl, np.arange(n, dtype=np.int32), True, ref, cache, cnt)
        print("score_cache_fill wall %.2f ms, deltas %d" % ((time.perf_counter() - t0) * 1e3, cnt.value), eng.last_call_profile())
    d.release()
