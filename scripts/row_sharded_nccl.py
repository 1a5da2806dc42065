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
    ap.add_argument("--gaussian", action="store_true",
                    help="the Gaussian half: per-rank Gram, pooled with two NCCL all-reduces (replicate.pool_moments)")
    ap.add_argument("--gcols", type=int, default=256)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    if args.gaussian:
        return gaussian(args, rank, world, local)
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


def gaussian(args, rank, world, local):
    """p numeric columns x rows, rows sharded: each rank forms the Gram of its block
    at registration, the moments are pooled over NCCL, every rank then serves the
    level-0 correlation tests from the pooled matrix."""
    from bnlearn_b200.replicate import pool_moments
    p, rows = args.gcols, args.rows
    rng = np.random.default_rng(20261115)                 # the same data on every rank; each keeps its block
    base = rng.standard_normal((8, rows))
    w = rng.standard_normal((p, 8)) * 0.4
    a, b = rows * rank // world, rows * (rank + 1) // world
    cols = [np.ascontiguousarray(w[j] @ base[:, a:b] + np.random.default_rng(100 + j).standard_normal(rows)[a:b])
            for j in range(p)]
    eng = Engine([local])
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mine = eng.register(cols)
    t_reg = time.perf_counter() - t0
    mean, cov = mine.moments()
    n, pm, pc = pool_moments(b - a, mean, cov, torch.device("cuda", local))
    mine.set_moments(n, pm, pc)
    torch.cuda.synchronize()
    t_pool = time.perf_counter() - t0 - t_reg
    xs = [i for i in range(p) for j in range(i + 1, p)]
    ys = [j for i in range(p) for j in range(i + 1, p)]
    st, df, pv = mine.ci_gaussian("cor", xs, ys, [[]] * len(xs))
    dt = time.perf_counter() - t0
    if rank == 0:
        full = [np.ascontiguousarray(w[j] @ base + np.random.default_rng(100 + j).standard_normal(rows))
                for j in range(p)]
        whole = eng.register(full)
        rst, rdf, rpv = whole.ci_gaussian("cor", xs, ys, [[]] * len(xs))
        err = float(np.max(np.abs(st - rst) / np.maximum(np.abs(rst), 1e-300)))
        print(json.dumps({"world": world, "rows": rows, "columns": p, "tests": len(xs), "pooled_rows": n,
                          "max_rel_diff_statistic_vs_unsharded": err, "df_identical": bool(np.array_equal(df, rdf)),
                          "moment_bytes_allreduced": int(8 * (1 + p + p * p)), "register_block_s": t_reg,
                          "pool_s": t_pool, "sharded_step_s": dt}))
        whole.release()
    mine.release()
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
