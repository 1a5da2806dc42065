"""Times the all-pairs contraction (csrc/bnl_pairs.cuh) on the two shapes that use it:
config B level 0 (37 factors x 10 M rows, 666 pairs) and config C's first hc iteration
(1000 factors x 1 M rows, 499 500 pairs). Prints count-kernel and device time per call
from the engine's CUDA events; BNL_PAIRS=0 gives the per-unit atomic kernels."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bnlearn_b200 import Engine, csr, synth  # noqa: E402


def timed(eng, fn, reps=5):
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        p = eng.last_call_profile()
        rec = (p["count_ms"], p["total_ms"], dt * 1e3, p["launches"])
        best = rec if best is None or rec[0] < best[0] else best
    return dict(zip(("count_ms", "device_ms", "wall_ms", "launches"), best))


def main():
    out = {}
    which = sys.argv[1] if len(sys.argv) > 1 else "BC"
    with Engine() as eng:
        if "B" in which:
            names, codes, lev = synth.alarm_shaped_cached(10_000_000)
            cols = [np.ascontiguousarray(c.astype(np.int32)) for c in codes]
            d = eng.register(cols, lev)
            n = len(lev)
            xs = np.array([i for i in range(n) for j in range(i + 1, n)], dtype=np.int32)
            ys = np.array([j for i in range(n) for j in range(i + 1, n)], dtype=np.int32)
            z = csr([[]] * len(xs))
            for knob in ("1", "0"):
                os.environ["BNL_PAIRS"] = knob
                out["B_level0_pairs=%s" % knob] = timed(eng, lambda: d.ci_discrete("mi", xs, ys, z))
            d.release()
        if "C" in which:
            host, lev, _ = synth.random_discrete_bn_device(1000, 1_000_000)
            d = eng.register([host[i] for i in range(1000)], lev)
            tg = np.array([j for j in range(1000) for i in range(j)], dtype=np.int32)
            pa = csr([[i] for j in range(1000) for i in range(j)])
            for knob in ("1", "0"):
                os.environ["BNL_PAIRS"] = knob
                out["C_first_iteration_pairs=%s" % knob] = timed(eng, lambda: d.score_discrete("bde", tg, pa, 1.0), 3)
            d.release()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
