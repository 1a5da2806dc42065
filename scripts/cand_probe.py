#!/usr/bin/env python
"""One hc candidate batch (999 candidates of a node with 2 parents, 1M rows) -- a
short command for ncu on the counting kernels."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bnlearn_b200 import Engine
rows, nnodes = 1_000_000, 1000
g = torch.Generator(device="cuda"); g.manual_seed(1)
cpu = np.random.Generator(np.random.PCG64(1))
levels = [int(v) for v in cpu.integers(2, 5, nnodes)]
host = np.empty((nnodes, rows), dtype=np.int32)
for i in range(nnodes):
    host[i] = (torch.randint(0, levels[i], (rows,), generator=g, device="cuda", dtype=torch.int32) + 1).cpu().numpy()
eng = Engine()
ds = eng.register([host[i] for i in range(nnodes)], levels)
tgt, npar = 500, int(os.environ.get("NPAR", "2"))
pa = [[tgt - 1 - k for k in range(npar)] + [i] for i in range(nnodes) if i != tgt and not (tgt - npar <= i < tgt)]
for _ in range(3):
    ds.score_discrete("bde", [tgt] * len(pa), pa, 1.0)
    print(eng.last_call_profile())
