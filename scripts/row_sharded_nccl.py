#!/usr/bin/env python
"""Row-sharded CI tests over N GPUs (one process per GPU, torchrun): every rank
registers its block of rows, counts with bnl_counts, the int32 tables are summed
with one NCCL all-reduce, and rank 0 turns them into statistics with
bnl_stat_from_tables and checks them against an unsharded run on its own GPU.
SURVEY.md section 8e: rows are sharded only when N is very large; units are the
normal sharding axis (bench.py).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29544 scripts/row_sharded_nccl.py --rows 10000000
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bnlearn_b200 import Engine, synth, workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    names, cols, lev = synth.alarm_shaped(args.rows)
    units = workloads.pc_stable_units(names, lev)
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    a, b = args.rows * rank // world, args.rows * (rank + 1) // world
    eng = Engine([local])
    mine = eng.register([c[a:b] for c in cols], lev)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tabs = mine.counts(xs, ys, zs)
    flat = torch.from_numpy(np.concatenate([t.ravel() for t in tabs])).cuda()
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    torch.cuda.synchronize()
    summed = flat.cpu().numpy()
    out = None
    if rank == 0:
        off = np.cumsum([0] + [t.size for t in tabs])
        full = [summed[off[i]:off[i + 1]].reshape(tabs[i].shape) for i in range(len(tabs))]
        out = eng.stat_from_tables("mi", full, [len(z) for z in zs])
    dt = time.perf_counter() - t0
    if rank == 0:
        whole = eng.register(cols, lev)
        ref = whole.ci_discrete("mi", xs, ys, zs)
        same = all(np.array_equal(g, r) for g, r in zip(out, ref))
        print(json.dumps({"world": world, "rows": args.rows, "tests": len(xs), "identical_to_unsharded": bool(same),
                          "table_bytes_allreduced": int(flat.numel() * 4), "sharded_step_s": dt}))
        whole.release()
    mine.release()
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
