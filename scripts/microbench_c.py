#!/usr/bin/env python
"""Count-kernel timings on the config-C shape under the library's tuning knobs
(BNL_PAIR, BNL_CHUNK_ROWS, BNL_L2_MB). One process per setting."""
import json, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def child():
    import torch
    from bnlearn_b200 import Engine, callsites as cs
    rows, nnodes = 1_000_000, 1000
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    cpu = np.random.Generator(np.random.PCG64(1))
    levels = [int(v) for v in cpu.integers(2, 5, nnodes)]
    host = np.empty((nnodes, rows), dtype=np.int32)
    for i in range(nnodes):
        host[i] = (torch.randint(0, levels[i], (rows,), generator=g, device="cuda", dtype=torch.int32) + 1).cpu().numpy()
    eng = Engine()
    ds = eng.register([host[i] for i in range(nnodes)], levels)
    out = {}
    tgt = 500
    for npar in (0, 1, 2, 3, 4):
        units_t = [tgt] * (nnodes - 1 - npar)
        pa = [[tgt - 1 - k for k in range(npar)] + [i] for i in range(nnodes) if i != tgt and not (tgt - npar <= i < tgt)]
        best = 1e9
        for _ in range(5):
            ds.score_discrete("bde", units_t[:len(pa)], pa, 1.0)
            best = min(best, eng.last_call_profile()["count_ms"])
        out["cand_%dp" % npar] = best
    # all pairs of the first 400 nodes
    iu = np.triu_indices(400, 1)
    best = 1e9
    for _ in range(3):
        ds.score_discrete("bde", iu[0].astype(np.int32), [[int(j)] for j in iu[1]], 1.0)
        best = min(best, eng.last_call_profile()["count_ms"])
    out["allpairs400_ms"] = best
    print(json.dumps(out))

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child()
    else:
        settings = [{}, {"BNL_GROUP": "0"}]
        for st in settings:
            env = dict(os.environ); env.update(st)
            r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
            print(st, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-500:], flush=True)
