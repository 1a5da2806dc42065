"""Host-side time breakdown of bnl_hc on config C (BNL_TRACE=1 prints the totals at shutdown)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("BNL_TRACE", "1")
from bnlearn_b200 import Engine, synth
host, lev, _ = synth.random_discrete_bn_device(1000, 1_000_000)
with Engine() as eng:
    d = eng.register([host[i] for i in range(1000)], lev)
    t0 = time.perf_counter()
    amat, _, it = d.hc("bde", 1.0, maxp=5)
    print("hc wall", time.perf_counter() - t0, "iterations", it, "arcs", int(amat.sum()))
    d.release()
