"""Config D's registration-time moments (col sums + centring + Gram + finish) on 2000 x 1 M doubles."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bnlearn_b200 import Engine, synth
rows, nvars = 1_000_000, 2000
host, _ = synth.random_gaussian_bn_device(nvars, rows)
with Engine() as eng:
    for rep in range(3):
        t0 = time.perf_counter()
        ds = eng.register([host[i] for i in range(nvars)])
        reg = time.perf_counter() - t0
        ms = eng.last_call_profile()["count_ms"]
        fl = float(rows) * nvars * (nvars + 1)
        print("register %.3f s  moments kernels %.2f ms  %.2f TFLOP/s algorithmic  (BNL_GRAM_SCRATCH_MB=%s)"
              % (reg, ms, fl / ms / 1e9, os.environ.get("BNL_GRAM_SCRATCH_MB")))
        ds.release()
