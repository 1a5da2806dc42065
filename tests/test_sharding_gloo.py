"""CPU test of the N > 1 path with world_size = 2 over gloo: units are split into
contiguous blocks per rank with no data-path collective, every rank evaluates
its block (here with the CPU oracle standing in for the device), and the
gathered, index-ordered results equal the unsharded evaluation. Also checks the
barrier / max-over-ranks timing reduction bench.py performs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import load_discrete


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _block(n, world, rank):
    return n * rank // world, n * (rank + 1) // world


def _worker(rank, world, port, units, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("learning_test")
    b, e = _block(len(units), world, rank)
    mine = np.zeros((e - b, 3))
    for i, (x, y, z) in enumerate(units[b:e]):
        mine[i] = orc.ci_discrete("mi", cols[x], nlev[x], cols[y], nlev[y], [cols[j] for j in z],
                                  [nlev[j] for j in z])
    gathered = [None] * world
    dist.all_gather_object(gathered, (b, mine))
    t = torch.tensor([0.010 * (rank + 1)], dtype=torch.float64)
    dist.barrier()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        full = np.concatenate([g[1] for g in sorted(gathered, key=lambda g: g[0])])
        np.save(out_path, np.concatenate([full.ravel(), t.numpy()]))
    dist.destroy_process_group()


def test_unit_sharding_world_size_2(tmp_path):
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("learning_test")
    rng = np.random.default_rng(3)
    units = []
    for _ in range(25):
        k = int(rng.integers(0, 3))
        idx = rng.choice(len(cols), size=2 + k, replace=False)
        units.append((int(idx[0]), int(idx[1]), [int(i) for i in idx[2:]]))
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, _free_port(), units, out), nprocs=2, join=True)
    got = np.load(out)
    ref = np.array([orc.ci_discrete("mi", cols[x], nlev[x], cols[y], nlev[y], [cols[j] for j in z],
                                    [nlev[j] for j in z]) for x, y, z in units]).ravel()
    assert np.array_equal(got[:-1], ref)
    assert got[-1] == 0.020          # max over ranks


def test_block_partition_matches_the_library():
    """bnl_block() (csrc/bnl_internal.h) and the Python partition agree and tile [0, n)."""
    for n in (0, 1, 7, 100, 1081):
        for world in (1, 2, 3, 8):
            edges = [_block(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))


def _row_worker(rank, world, port, units, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("learning_test")
    n = len(cols[0])
    a, b = _block(n, world, rank)                       # this rank's block of ROWS
    flat = []
    for x, y, z in units:
        t = orc.counts(cols[x][a:b], nlev[x], cols[y][a:b], nlev[y], [cols[j][a:b] for j in z],
                       [nlev[j] for j in z])
        flat.append(t.ravel())
    buf = torch.from_numpy(np.concatenate(flat).astype(np.int32))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)           # what ncclAllReduce(sum, int32) does on the GPUs
    if rank == 0:
        np.save(out_path, buf.numpy())
    dist.destroy_process_group()


def test_row_sharding_allreduce_world_size_2(tmp_path):
    """Rows sharded over two ranks: per-rank tables, one int32 sum all-reduce, and
    the result is the table of the whole data set (the statistics then come from
    bnl_stat_from_tables; tests/test_gpu_parity.py covers that half on the device)."""
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("learning_test")
    units = [(0, 1, []), (2, 3, [0]), (4, 5, [1, 2]), (1, 0, [3, 4, 5])]
    out = str(tmp_path / "tables.npy")
    mp.spawn(_row_worker, args=(2, _free_port(), units, out), nprocs=2, join=True)
    got = np.load(out)
    ref = np.concatenate([orc.counts(cols[x], nlev[x], cols[y], nlev[y], [cols[j] for j in z],
                                     [nlev[j] for j in z]).ravel() for x, y, z in units])
    assert np.array_equal(got, ref)


def _perm_worker(rank, world, port, batch, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import ctypes as C
    from oracle import oracle as orc
    names, cols, nlev = load_discrete("learning_test")
    n = len(cols[0])
    a, b = _block(n, world, rank)
    res = batch["results"]
    flat = [orc.counts(cols[r["x"]][a:b], nlev[r["x"]], cols[r["y"]][a:b], nlev[r["y"]],
                       [cols[j][a:b] for j in r["z"]], [nlev[j] for j in r["z"]]).ravel() for r in res]
    off = np.concatenate([[0], np.cumsum([len(t) for t in flat])]).astype(np.int64)
    buf = torch.from_numpy(np.concatenate(flat).astype(np.int32))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    if rank == 0:                                         # the random stream lives on one rank
        from bnlearn_b200 import _lib
        lib = _lib.load()
        tab = np.ascontiguousarray(buf.numpy(), dtype=np.int32)
        m = len(res)
        llx = np.array([nlev[r["x"]] for r in res], dtype=np.int32)
        lly = np.array([nlev[r["y"]] for r in res], dtype=np.int32)
        nz = np.array([len(r["z"]) for r in res], dtype=np.int32)
        st, df, pv = (np.zeros(m) for _ in range(3))
        state = C.c_uint64(batch["seed"])
        ip = lambda v: v.ctypes.data_as(C.POINTER(C.c_int))
        dp = lambda v: v.ctypes.data_as(C.POINTER(C.c_double))
        rc = lib.bnl_mc_from_tables(None, 50, batch["B"], batch["alpha"], m, ip(llx), ip(lly), ip(nz),
                                    tab.ctypes.data_as(C.POINTER(C.c_int32)),
                                    off.ctypes.data_as(C.POINTER(C.c_int64)),
                                    C.cast(lib.bnl_unif_splitmix64, C.c_void_p),
                                    C.cast(C.pointer(state), C.c_void_p), dp(st), dp(df), dp(pv))
        assert rc == 0
        np.save(out_path, np.concatenate([st, pv]))
    dist.destroy_process_group()


def test_row_sharded_permutation_test_world_size_2(tmp_path):
    """mc-mi with the rows split over two ranks: tables all-reduced, then
    bnl_mc_from_tables on rank 0 gives the statistics and Monte Carlo p-values of the
    reference's c_mcarlo / c_cmcarlo on the whole data (tests/golden/ref_permutation.json)."""
    import json
    from util import GOLDEN
    with open(os.path.join(GOLDEN, "ref_permutation.json")) as f:
        batch = [b for b in json.load(f)["learning_test"] if b["test"] == "mc-mi"][0]
    out = str(tmp_path / "perm.npy")
    mp.spawn(_perm_worker, args=(2, _free_port(), batch, out), nprocs=2, join=True)
    got = np.load(out)
    m = len(batch["results"])
    assert np.array_equal(got[:m], np.array([r["stat"] for r in batch["results"]]))
    assert np.array_equal(got[m:], np.array([r["p"] for r in batch["results"]]))


def _moments_worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bnlearn_b200.replicate import pool_moments
    from util import load_gaussian
    names, cols = load_gaussian("gaussian_test")
    x = np.array(cols)
    a, b = _block(x.shape[1], world, rank)
    blk = x[:, a:b]
    n, mean, cov = pool_moments(b - a, blk.mean(axis=1), np.cov(blk), None)
    if rank == 1:                                         # every rank holds the pooled moments
        np.save(out_path, np.concatenate([[n], mean, cov.ravel()]))
    dist.destroy_process_group()


def test_row_sharded_gaussian_moments_world_size_2(tmp_path):
    """Rows of numeric data sharded over two ranks: pool_moments (two sum
    all-reduces) gives the means and the covariance matrix of the whole data."""
    from util import load_gaussian
    names, cols = load_gaussian("gaussian_test")
    x = np.array(cols)
    out = str(tmp_path / "moments.npy")
    mp.spawn(_moments_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    p = x.shape[0]
    assert got[0] == x.shape[1]
    assert np.allclose(got[1:1 + p], x.mean(axis=1), rtol=1e-13, atol=0)
    assert np.allclose(got[1 + p:].reshape(p, p), np.cov(x), rtol=1e-12, atol=1e-15)


# ---------------------------------------------------------------------------
# The level batches of pc.stable split over two ranks (replicate.ShardedBackend,
# what `bench.py --gpus N` runs over NCCL): here over gloo, with the reference's
# own C evaluating each rank's block, against the unsharded run.

def _pc_worker(rank, world, port, out_path):
    import pickle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bnlearn_b200 import learners
    from bnlearn_b200.replicate import ShardedBackend, block, gather_rows
    from refbackend import RefBackend
    names, cols, nlev = load_discrete("learning_test")
    sb = ShardedBackend(RefBackend(cols, nlev))
    skel, dsep, pvals = learners.pc_stable(sb, len(cols), "mi", 0.05)
    tests = sb.total_tests()
    # gather_rows on a ragged split: 7 rows over 2 ranks
    b, e = block(7, world, rank)
    rows = gather_rows(np.arange(b, e, dtype=np.float64).reshape(-1, 1) * np.array([[1.0, 10.0]]), 7)
    if rank == 0:
        with open(out_path, "wb") as f:
            pickle.dump((skel, dsep, pvals, tests, sb.gathers, rows), f)
    dist.destroy_process_group()


def test_sharded_pc_stable_world_size_2(tmp_path):
    import pickle
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libbnref.so is not built")
    from bnlearn_b200 import learners
    from refbackend import RefBackend
    names, cols, nlev = load_discrete("learning_test")
    rbk = RefBackend(cols, nlev)
    want = learners.pc_stable(rbk, len(cols), "mi", 0.05)
    want_tests = rbk.counter.value
    out = str(tmp_path / "pc.pkl")
    mp.start_processes(_pc_worker, args=(2, _free_port(), out), nprocs=2, join=True, start_method="fork")
    with open(out, "rb") as f:
        skel, dsep, pvals, tests, gathers, rows = pickle.load(f)
    assert skel == want[0] and dsep == want[1] and pvals == want[2]
    assert tests == want_tests and gathers >= 2
    assert np.array_equal(rows, np.arange(7.0).reshape(-1, 1) * np.array([[1.0, 10.0]]))
