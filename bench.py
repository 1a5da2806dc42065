#!/usr/bin/env python
"""bench.py -- headline benchmark: conditional-independence tests per second on
BASELINE.json configs[1] (pc.stable() + mi on alarm-shaped discrete data,
37 factors x 10M rows), one process per GPU.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA engine
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own C on the host cores

A step = one pass of the hot path over one batch: every CI test pc.stable issues
on this data at levels 0-2 (bnlearn_b200/workloads.py, 1081 tests). With N ranks
the ONE fixed test list is cut into N contiguous blocks, every rank evaluates its
block on its own replica of the data set and the statistics are gathered to rank
0 in index order inside the timed region (strong scaling).
`value` has the data set resident in HBM; `e2e` registers it from host memory
inside every step. Rank 0 prints ONE JSON line. Timing rules: W >= 3 warm-up
steps, the data set (370 MB packed) is larger than the 126 MB L2 so no explicit
flush is needed, steps are bracketed by barrier + synchronize, max over ranks;
kernel times come from CUDA events on the engine's own stream.

The line also carries, measured in the same run: `roofline` (per launch class,
SURVEY.md section 8d: compulsory bytes / kernel time against the measured HBM
peak, the effective rate and the shared-memory-atomic bound as separate keys),
`cpu_baseline`, a parity check of the timed outputs against the reference's C,
the score-delta (config C) and Gaussian (config D) numbers, whole-learner wall
times (`wall`) and the sharded batches of configs B, C and E (`strong`).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from bnlearn_b200 import synth, workloads  # noqa: E402

ATOMIC_CEILING = 5.13e12      # shared-memory atomic lane updates/s on one B200 (profiles/r1_atom_probe.txt)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--cpu-sample", type=int, default=0, help="tests in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only: skip configs A, C, D, E")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc, self.lines, self.index = None, [], index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(smax) if smax else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_traffic(same_workload, compulsory):
    """`roofline.traffic`: DRAM bytes of the step's counting launches from the ncu
    capture under profiles/ (scripts/traffic_stamp.py) -- only while the capture
    belongs to the kernel sources this run was built from and to this workload."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        from traffic_stamp import source_stamp
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if not same_workload:
            return {"traffic": None, "traffic_note": "profiles/traffic.json holds the N = 1, 10M-row step only"}
        if t.get("source_stamp") != source_stamp():
            return {"traffic": None, "traffic_note": "profiles/traffic.json was captured on other kernel sources "
                                                     "(stamp %s); re-run scripts/traffic_stamp.py" % t.get("source_stamp")}
        k = {n: v["dram_bytes_per_launch"] for n, v in t["kernels"].items()
             if n.startswith(("count_", "pair_onehot"))}
        total = int(sum(k.values()))
        return {"traffic": total, "traffic_by_kernel": k, "traffic_over_compulsory": total / compulsory,
                "traffic_note": "ncu --set full capture of this command (profiles/traffic.json, "
                                "profiles/r2_bench_step_kernel_details.csv): each counting launch of the step reads "
                                "its columns from DRAM once -- level 0 the byte columns in the pair kernel, levels "
                                "1-2 the nibble copy (half the bytes), once in count_pair32 (tables of <= 32 cells) and "
                                "once in count_small -- and nothing more"}
    except Exception as exc:                         # a missing capture must not cost the bench line
        return {"traffic": None, "traffic_note": "no capture: %r" % (exc,)}


# --------------------------------------------------------------------------
# CPU legs: the reference's own C (oracle/_ref), forked over all host cores.

_CPU = {}


def _cpu_worker(args):
    lo, hi = args
    cols, lev, units = _CPU["cols"], _CPU["lev"], _CPU["units"]
    if _CPU["kind"] == "reference":
        from oracle import refapi as ra, refbind as rb
        L = rb.lib()
        for (x, y, z) in units[lo:hi]:
            if z:
                cfg, llz = ra.fast_config([cols[j] for j in z], [lev[j] for j in z], 1)
                st, df = ra.chisqtest(cols[x], lev[x], cols[y], lev[y], 1, cfg, llz)
            else:
                st, df = ra.chisqtest(cols[x], lev[x], cols[y], lev[y], 1)
            L.Rf_pchisq(st, df, 0, 0)
    else:
        from oracle import oracle as orc
        for (x, y, z) in units[lo:hi]:
            orc.ci_discrete("mi", cols[x], lev[x], cols[y], lev[y], [cols[j] for j in z],
                            [lev[j] for j in z])
    return hi - lo


def cpu_run(cols32, lev, units, cores):
    """Time `units` through the CPU implementation on `cores` forked workers."""
    import multiprocessing as mp
    from oracle import refbind
    _CPU.update(cols=cols32, lev=lev, units=units,
                kind="reference" if refbind.available() else "port")
    per = (len(units) + cores - 1) // cores
    spans = [(i, min(i + per, len(units))) for i in range(0, len(units), per)]
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_worker, [(0, 0)] * cores)          # workers forked and warm before the clock starts
        t0 = time.perf_counter()
        done = sum(pool.map(_cpu_worker, spans, chunksize=1))
        dt = time.perf_counter() - t0
    return done / dt, dt, _CPU["kind"]


def reference_ci(cols32, lev, unit):
    """(statistic, df, p-value) of one mi test by the reference's C (oracle/_ref), or
    by the oracle port when the reference was not compiled."""
    from oracle import refbind
    x, y, z = unit
    if refbind.available():
        from oracle import refapi as ra
        if z:
            cfg, llz = ra.fast_config([cols32[j] for j in z], [lev[j] for j in z], 1)
            st, df = ra.chisqtest(cols32[x], lev[x], cols32[y], lev[y], 1, cfg, llz)
        else:
            st, df = ra.chisqtest(cols32[x], lev[x], cols32[y], lev[y], 1)
        return st, df, refbind.lib().Rf_pchisq(st, df, 0, 0)
    from oracle import oracle as orc
    return orc.ci_discrete("mi", cols32[x], lev[x], cols32[y], lev[y], [cols32[j] for j in z], [lev[j] for j in z])


def _cpu_score_worker(span):
    """BDe scores of candidate parent sets (target | base parents + one candidate)
    through the reference's per_node_score -- c_fast_config, c_configurations /
    int2fac and cdpost (src/scores/dirichlet.posterior.c:152-202) -- or, when
    oracle/_ref is not built, the plain-C port."""
    lo, hi = span
    host, lev, tgt, base, cands = _CPU["score"]
    if hi <= lo:
        return 0
    if _CPU["kind"] == "reference":
        from oracle import refapi as ra, refbind as rb
        idx = [tgt] + base + cands[lo:hi]
        names = ["V%d" % i for i in idx]
        data = ra.RefData([host[i] for i in idx], [lev[i] for i in idx], names)
        for c in cands[lo:hi]:
            par = {nm: [] for nm in names}
            par["V%d" % tgt] = ["V%d" % i for i in base + [c]]
            ra.per_node_score(data, par, ["V%d" % tgt], "bde", iss=1.0)
        rb.reset()
    else:
        from oracle import oracle as orc
        for c in cands[lo:hi]:
            pa = base + [c]
            orc.score_discrete("bde", host[tgt], lev[tgt], [host[i] for i in pa], [lev[i] for i in pa], 1.0)
    return hi - lo


def cpu_score_run(host, lev, tgt, base, cands, cores):
    """Score deltas per second of the CPU implementation on `cores` forked workers."""
    import multiprocessing as mp
    from oracle import refbind
    _CPU.update(score=(host, lev, tgt, base, cands), kind="reference" if refbind.available() else "port")
    per = (len(cands) + cores - 1) // cores
    spans = [(i, min(i + per, len(cands))) for i in range(0, len(cands), per)]
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_score_worker, [(0, 0)] * cores)
        t0 = time.perf_counter()
        done = sum(pool.map(_cpu_score_worker, spans, chunksize=1))
        dt = time.perf_counter() - t0
    return done / dt, dt, _CPU["kind"]


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, ValueError, KeyError):
        return 6650.0, "fallback 6650 (B200_PROFILING.md)"


def hbm_class(name, kernel_ms, compulsory_bytes, unit_bytes, units, rows, peak, note=None, atomic=True):
    """One roofline entry of a counting launch. achieved =
    compulsory bytes (each distinct column once) / kernel time, frac <= 1; effective =
    the sum of the units' own bytes / time; the >= 50 % HBM target applies to launches
    with at most 2 units per distinct column. atomic_frac = row updates per second
    over the shared-memory-atomic ceiling, for launches counted with atomics."""
    t = kernel_ms * 1e-3
    per_col = float(units) * float(rows) / max(compulsory_bytes, 1.0)
    e = {"class": name, "bound": "hbm", "units": int(units), "kernel_ms": kernel_ms,
         "compulsory_bytes": float(compulsory_bytes), "achieved": compulsory_bytes / t / 1e9,
         "peak": peak, "unit": "GB/s", "frac": compulsory_bytes / t / 1e9 / peak,
         "effective": unit_bytes / t / 1e9, "units_per_distinct_column": per_col,
         "in_target_domain": bool(per_col <= 2.0)}
    if atomic:
        e["atomic_frac"] = float(units) * float(rows) / t / ATOMIC_CEILING
        e["atomic_ceiling_updates_per_s"] = ATOMIC_CEILING
    if note:
        e["note"] = note
    return e


# --------------------------------------------------------------------------
# config C: score deltas per second (hc/tabu + BDe, 1000 factors x 1M rows).

def config_c_extras(eng, world, rank, device, cpu_cores=0, reps=5, want_hc=True):
    import torch.distributed as dist
    from bnlearn_b200 import callsites as cs
    from bnlearn_b200.replicate import broadcast_dataset
    rows, nnodes = 1_000_000, 1000
    peak, _ = hbm_peak()
    if rank == 0:
        host, levels, _ = synth.random_discrete_bn_device(nnodes, rows)
        cols = [host[i] for i in range(nnodes)]
    else:
        host, cols = None, [None] * nnodes
        levels = [int(v) for v in np.random.Generator(np.random.PCG64(20261107)).integers(2, 5, nnodes)]
    ds = broadcast_dataset(eng, cols, levels, rows, src=0) if world > 1 else eng.register(cols, levels)
    out = {"shape": "%d factors x %d rows, BDe iss=1" % (nnodes, rows)}
    classes = []
    counter = cs.TestCounter()
    blmat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")

    def timed(amat, updated, equivalence, bl=blmat):
        par = [[int(i) for i in np.flatnonzero(amat[:, j])] for j in updated]
        ref = np.zeros(nnodes)
        ref[updated] = ds.score_discrete("bde", list(updated), par, 1.0)[0]
        cache = np.zeros((nnodes, nnodes), order="F")
        best, kms, n, dev_ms = None, None, 0, None
        for _ in range(reps):
            counter.reset()
            t0 = time.perf_counter()
            cs.score_cache_fill(ds, "bde", 1.0, amat, bl, updated, equivalence, ref, cache, counter)
            dt = time.perf_counter() - t0
            prof = eng.last_call_profile()
            if best is None or dt < best:
                best, kms, n, dev_ms = dt, prof["count_ms"], counter.value, prof["total_ms"]
        return n, best, kms, dev_ms, cache

    if rank == 0:
        # (a) mid-run iterations: every other node as the candidate parent of a node that
        #     already has k parents, k = 0..5 (tables of k + 2 columns).
        tgt = nnodes // 2
        amat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")
        for k in range(0, 6):
            if k:
                amat[tgt - k, tgt] = 1
            n, dt, kms, dms, _ = timed(amat, [tgt], False)
            comp = float(rows) * (n + k + 1)        # every candidate + target + shared parents, once
            ubytes = float(rows) * n * (k + 2)
            key = "iteration_%d_parents" % k
            out[key] = {"existing_parents": k, "columns_per_table": k + 2, "deltas": n, "ms": dt * 1e3,
                        "deltas_per_sec": n / dt, "count_kernel_ms": kms, "device_ms": dms,
                        "compulsory_gb": comp / 1e9, "hbm_frac_compulsory": comp / (kms * 1e-3) / 1e9 / peak}
            classes.append(hbm_class("config C: %d candidate parents of a node that has %d (tables of %d columns)"
                                     % (n, k, k + 2), kms, comp, ubytes, n, rows, peak))
            if k == 2 and cpu_cores > 0:
                # the same deltas on the host: a bounded sample of the candidates, 4 per core
                cands = [i for i in range(nnodes) if i not in (tgt, tgt - 1, tgt - 2)][:4 * cpu_cores]
                try:
                    v, cdt, kind = cpu_score_run(host, levels, tgt, [tgt - 2, tgt - 1], cands, cpu_cores)
                    out[key]["cpu_reference"] = {"deltas_per_sec": v, "cores": cpu_cores, "kind": kind,
                                                 "sample": "%d of the %d candidates (%.1f s)" % (len(cands), int(n), cdt)}
                except Exception as exc:          # the CPU leg must never take the GPU numbers down
                    out[key]["cpu_reference"] = {"unavailable": str(exc)[:200]}
        out["score_deltas_per_sec"] = out["iteration_2_parents"]["deltas_per_sec"]

    # (b) the first iteration, sharded by target node: every arc into every node of the
    #     empty graph, one orientation per pair (score equivalence, hc.cache.lookup.c:72-89).
    #     Rank r takes a contiguous block of targets holding ~1/N of the pairs i < j.
    amat = np.zeros((nnodes, nnodes), dtype=np.int32, order="F")
    lower = np.asfortranarray(np.tril(np.ones((nnodes, nnodes), dtype=np.int32), -1))   # i > j: the mirrored half
    bounds = [int(round(nnodes * np.sqrt(r / world))) for r in range(world + 1)]
    mine = list(range(bounds[rank], bounds[rank + 1]))
    if world > 1:
        dist.barrier()
    n, dt, kms, dms, cache = timed(amat, mine, False, lower)
    if world > 1:
        dist.barrier()                  # the ranks finish their repetitions at different times
    t0 = time.perf_counter()
    if world > 1:
        # blocks are uneven (triangular work): gather the whole matrix, each rank's columns filled
        import torch
        t = torch.from_numpy(np.ascontiguousarray(cache)).to(device)
        dist.all_reduce(t)
        full = t.cpu().numpy()
    else:
        full = cache
    gather_s = time.perf_counter() - t0
    tt = np.array([dt + gather_s, kms, float(n)])
    if world > 1:
        import torch
        t = torch.tensor(tt, dtype=torch.float64, device=device)
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm)
        wall, kmax, total = float(mx[0]), float(mx[1]), float(sm[2])
    else:
        wall, kmax, total = tt[0], tt[1], tt[2]
    if rank == 0:
        full = full + np.tril(full.T, -1)          # the mirrored half (score equivalence)
        out["first_iteration"] = {"deltas": total, "ms": wall * 1e3, "deltas_per_sec": total / wall,
                                  "count_kernel_ms_max_over_ranks": kmax, "n_gpus": world,
                                  "sharding": "contiguous blocks of target nodes, ~equal pair counts; "
                                              "cache columns summed to every rank with one NCCL all-reduce",
                                  "cache_sha1": hashlib.sha1(np.ascontiguousarray(full).tobytes()).hexdigest()[:16]}
        comp = float(rows) * nnodes
        classes.append(hbm_class("config C: first iteration, all %d pairs (tcgen05 indicator cross-product)" % int(total),
                                 kmax, comp, 2.0 * rows * total / world, total / world, rows, peak, atomic=False,
                                 note="max over %d rank(s); every two-way table from one contraction, so the "
                                      "columns are read once per output tile row/column, not once per pair" % world))
    # (c) the whole search on one GPU
    if rank == 0 and want_hc:
        # twice: the first run on a data set also pays for what is made once per data set or context
        # (scratch buffers, the nibble copy of the columns; a cudaMalloc of 0.5 GB was seen to stall
        # for 0.6 s right after another process had released the GPU) -- both are reported
        t0 = time.perf_counter()
        ds.hc("bde", 1.0, maxp=5)
        first = time.perf_counter() - t0
        t0 = time.perf_counter()
        famat, _, fiter = ds.hc("bde", 1.0, maxp=5)
        out["hc_to_convergence"] = {"wall_s": time.perf_counter() - t0, "wall_s_first_run": first,
                                    "iterations": int(fiter), "arcs": int(np.asarray(famat).sum())}
    ds.release()
    return out, classes


# --------------------------------------------------------------------------
# config D: Gaussian CI tests (pc.stable + cor, 2000 numeric columns x 1M rows): one
# Gram matrix on the fp64 tensor cores at registration, then batched solves from it.

def config_d_extras(eng):
    import torch
    rows, nvars = 1_000_000, 2000
    host, _ = synth.random_gaussian_bn_device(nvars, rows)
    # fp64 GEMM probe for the DMMA roofline denominator (cuBLAS DGEMM 8192^3, best of 5).
    a = torch.randn((8192, 8192), dtype=torch.float64, device="cuda")
    torch.matmul(a, a)
    torch.cuda.synchronize()
    probe = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, a); e1.record()
        torch.cuda.synchronize()
        probe.append(2 * 8192 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    dgemm_tf = max(probe)
    del a
    torch.cuda.empty_cache()

    t0 = time.perf_counter()
    ds = eng.register([host[i] for i in range(nvars)])
    reg_s = time.perf_counter() - t0
    gram_ms = eng.last_call_profile()["count_ms"]
    flops = float(rows) * nvars * (nvars + 1)
    out = {"shape": "%d numeric columns x %d rows" % (nvars, rows), "register_s": reg_s,
           "gram": {"kernel_ms": gram_ms, "algorithmic_flops": flops,
                    "achieved_tflops": flops / (gram_ms * 1e-3) / 1e12,
                    "dgemm_probe_tflops": dgemm_tf, "dgemm_probe_all": probe,
                    "frac_of_dgemm_probe": flops / (gram_ms * 1e-3) / 1e12 / dgemm_tf}}
    cls = {"class": "config D: Gram 2000 x 1 M (gram_dmma_kernel)", "bound": "tensor", "kernel_ms": gram_ms,
           "achieved": flops / (gram_ms * 1e-3) / 1e12, "peak": dgemm_tf, "unit": "TFLOP/s",
           "frac": flops / (gram_ms * 1e-3) / 1e12 / dgemm_tf,
           "peak_source": "cuBLAS DGEMM 8192^3 in this run, best of 5 (no fp64 entry in MEASURED_PEAKS.json)"}
    iu = np.triu_indices(nvars, 1)
    xs, ys = iu[0].astype(np.int32), iu[1].astype(np.int32)
    empty = (np.zeros(len(xs) + 1, dtype=np.int32), np.zeros(1, dtype=np.int32))
    for test in ("cor", "zf"):
        best = None
        for _ in range(3):
            t0 = time.perf_counter()
            st, df, pv = ds.ci_gaussian(test, xs, ys, empty)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        out["level0_" + test] = {"tests": len(xs), "ms": best * 1e3, "tests_per_sec": len(xs) / best,
                                 "kernel_ms": eng.last_call_profile()["count_ms"],
                                 "edges_kept_at_0.05": int((pv < 0.05).sum())}
    # the same level on the device end to end: pairs enumerated there, p-value matrix (32 MB) and the
    # surviving pairs back (bnl_ci_gaussian_allpairs); without the matrix only the survivors cross PCIe
    for want_p in (True, False):
        best = None
        for _ in range(3):
            t0 = time.perf_counter()
            keep, pm = ds.ci_gaussian_allpairs("cor", 0.05, want_p=want_p)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        out["level0_cor_allpairs" + ("" if want_p else "_edges_only")] = {
            "tests": len(xs), "ms": best * 1e3, "tests_per_sec": len(xs) / best,
            "kernel_ms": eng.last_call_profile()["count_ms"], "edges_kept_at_0.05": int(len(keep))}
    rng = np.random.default_rng(5)
    for k in (1, 3):
        n = 400000
        idx = np.stack([rng.permutation(nvars)[:2 + k] for _ in range(n)]).astype(np.int32)
        off = np.arange(0, (n + 1) * k, k, dtype=np.int32)
        zi = np.ascontiguousarray(idx[:, 2:]).ravel()
        t0 = time.perf_counter()
        ds.ci_gaussian("cor", idx[:, 0].copy(), idx[:, 1].copy(), (off, zi))
        dt = time.perf_counter() - t0
        out["cond_size_%d" % k] = {"tests": n, "ms": dt * 1e3, "tests_per_sec": n / dt,
                                   "kernel_ms": eng.last_call_profile()["count_ms"]}
    ds.release()
    # the reference recomputes the covariance from the raw columns for every test.
    from oracle import oracle as orc
    t0 = time.perf_counter()
    for i in range(8):
        orc.ci_gaussian("cor", [host[2 * i], host[2 * i + 1]])
    out["cpu_port_level0_tests_per_sec_per_core"] = 8 / (time.perf_counter() - t0)
    return out, cls


# --------------------------------------------------------------------------
# whole learners: wall time here, the reference's C beside it.

def wall_times(eng, ds_b, cols_b, lev_b, cores, with_cpu):
    from bnlearn_b200 import callsites as cs, learners
    out = {}
    # B: pc.stable + mi on the 10 M-row data set, in one library call
    counter = cs.TestCounter()
    t0 = time.perf_counter()
    skel, dsep, _ = ds_b.pc_stable("mi", 0.05, counter=counter)
    tb = time.perf_counter() - t0
    out["config_B_pc_stable"] = {"wall_s": tb, "tests": counter.value, "edges": len(skel)}
    # tests per conditioning-set size, for the reference's extrapolation: the level loop once more,
    # call by call, counting what each level issues
    from bnlearn_b200.replicate import ShardedBackend
    sb = ShardedBackend(learners.GpuBackend(ds_b))
    learners.pc_stable(sb, len(lev_b), "mi", 0.05)
    by_level = dict(sb.tests_by_size)
    # A: hc + BIC on the bundled alarm sample
    z = np.load(os.path.join(ROOT, "tests", "golden", "data_alarm.npz"))
    acols = [np.ascontiguousarray(c.astype(np.int32) + 1) for c in z["codes"]]
    alev = [int(v) for v in z["nlevels"]]
    da = eng.register(acols, alev)
    k = np.log(len(acols[0])) / 2
    da.hc("bic", k)
    counter.reset()
    t0 = time.perf_counter()
    amat_a, _, it_a = da.hc("bic", k, counter=counter)
    ta = time.perf_counter() - t0
    out["config_A_hc_bic_alarm"] = {"wall_s": ta, "deltas": counter.value, "iterations": int(it_a),
                                    "arcs": int(np.asarray(amat_a).sum())}
    da.release()
    # E: boot.strength, 256 x hc(bde) on insurance-shaped data
    _, ecols, elev = synth.insurance_shaped(20000)
    learners.boot_strength(eng, ecols, elev, 8, "bde", 1.0, seed=20261109, workers=8)
    t0 = time.perf_counter()
    strength, _, etests = learners.boot_strength(eng, ecols, elev, 256, "bde", 1.0, seed=20261109, workers=8)
    te = time.perf_counter() - t0
    out["config_E_boot_strength_256"] = {"wall_s": te, "replicates_per_sec": 256 / te, "deltas": etests,
                                         "host_threads": 8}
    if not with_cpu:
        return out
    try:
        from oracle import refbind
        if not refbind.available():
            raise RuntimeError("oracle/_ref not built")
        from oracle.refbackend import RefBackend
        # A in full, one core (hc has no cluster mode in the reference)
        rbk = RefBackend(acols, alev)
        t0 = time.perf_counter()
        ramat, _ = learners.hc(rbk, len(acols), "bic", k)
        out["config_A_hc_bic_alarm"]["reference"] = {
            "wall_s": time.perf_counter() - t0, "cores": 1, "deltas": rbk.counter.value,
            "identical_graph": bool(np.array_equal(np.asarray(ramat), np.asarray(amat_a)))}
        # E: one replicate in full, one core; 256 replicates = 256 x that, ceil(256 / P) x that on P cores
        rows = np.random.Generator(np.random.PCG64(20261109)).integers(0, 20000, 20000)
        rbk = RefBackend([c[rows] for c in ecols], elev)
        t0 = time.perf_counter()
        learners.hc(rbk, len(ecols), "bde", 1.0)
        per = time.perf_counter() - t0
        out["config_E_boot_strength_256"]["reference"] = {
            "wall_s_per_replicate": per, "cores": 1, "serial_wall_s_extrapolated": 256 * per,
            "cluster_wall_s_extrapolated": -(-256 // cores) * per, "cluster_cores": cores,
            "how": "replicate 0 timed in full through the reference's score_cache_fill / per_node_score; "
                   "R = 256 extrapolated (serial, and ceil(R/P) replicates per worker, R/arc.strength.R:210)"}
        # B: sampled and extrapolated -- 16 tests per conditioning-set size of the tests this
        # run issued, one core; the cluster figure divides each level's pairs over P workers.
        by_size = {}
        for u in workloads.pc_stable_units(list(synth.ALARM_NAMES), lev_b, max_level=3):
            by_size.setdefault(len(u[2]), []).append(u)
        rng = np.random.default_rng(2)
        per_size = {}
        for s, us in sorted(by_size.items()):
            pick = [us[i] for i in rng.choice(len(us), size=min(16, len(us)), replace=False)]
            t0 = time.perf_counter()
            for u in pick:
                reference_ci(cols_b, lev_b, u)
            per_size[s] = (time.perf_counter() - t0) / len(pick)
        last = max(per_size)
        serial = sum(n * per_size.get(k, per_size[last]) for k, n in by_level.items())
        out["config_B_pc_stable"]["reference"] = {
            "serial_wall_s_extrapolated": serial, "cluster_wall_s_extrapolated": serial / cores,
            "cluster_cores": cores, "seconds_per_test_by_conditioning_size": per_size,
            "tests_by_conditioning_size": by_level,
            "how": "16 tests per conditioning-set size timed on one core through c_fast_config + c_cchisqtest "
                   "+ pchisq; wall = sum over sizes of (tests this run issued) x (seconds per test); the cluster "
                   "figure divides every level over P workers (R/pcalgo.R:24-28), no imbalance counted"}
    except Exception as exc:                  # the CPU legs must never take the GPU numbers down
        out["reference_error"] = repr(exc)[:300]
    return out


# --------------------------------------------------------------------------
# sharded learners at N ranks: pc.stable's level batches and bootstrap replicates.

def strong_extras(eng, ds_b, world, rank, device):
    import torch
    import torch.distributed as dist
    from bnlearn_b200 import learners
    from bnlearn_b200.replicate import ShardedBackend
    out = {}

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # B: the whole skeleton phase, every level's pairs split over the ranks (R/pcalgo.R:24-28)
    sb = ShardedBackend(learners.GpuBackend(ds_b), device)
    learners.pc_stable(sb, 37, "mi", 0.05)
    sb.backend.counter.reset()
    sync()
    t0 = time.perf_counter()
    skel, dsep, pvals = learners.pc_stable(sb, 37, "mi", 0.05)
    sync()
    dt = time.perf_counter() - t0
    tests = sb.total_tests()
    h = hashlib.sha1(repr((skel, sorted(dsep.items()), sorted(pvals.items()))).encode()).hexdigest()[:16]
    out["config_B_pc_stable_sharded"] = {"wall_s": dt, "tests": tests, "tests_per_sec": tests / dt,
                                         "gathers": sb.gathers, "edges": len(skel), "result_sha1": h,
                                         "how": "level loop on every rank (Python), each level's pairs in contiguous "
                                                "blocks per rank, one all-gather per batched call"}
    # E: 256 bootstrap replicates, replicate r on rank r mod N, arc counts summed (bootstrap_reduce)
    _, ecols, elev = synth.insurance_shaped(20000)
    ids = list(range(rank, 256, world))
    learners.boot_arc_counts(eng, ecols, elev, ids[:8], "bde", 1.0, seed=20261109, workers=8)
    sync()
    t0 = time.perf_counter()
    arcs, etests = learners.boot_arc_counts(eng, ecols, elev, ids, "bde", 1.0, seed=20261109, workers=8)
    t = torch.from_numpy(np.concatenate([arcs.ravel(), [etests]])).to(device)
    if world > 1:
        dist.all_reduce(t)
    tot = t.cpu().numpy()
    sync()
    dt = time.perf_counter() - t0
    strength, _ = learners.strength_from_counts(tot[:-1].reshape(arcs.shape), 256)
    out["config_E_boot_strength_sharded"] = {
        "wall_s": dt, "replicates": 256, "replicates_per_sec": 256 / dt, "deltas": float(tot[-1]),
        "strength_sha1": hashlib.sha1(np.ascontiguousarray(strength).tobytes()).hexdigest()[:16],
        "how": "replicate r on rank r mod N, 8 host threads per rank; arc counts summed with one NCCL all-reduce "
               "(bootstrap_reduce, src/learning/averaging/bootstrap.c:126-186)"}
    if world > 1:
        mx = torch.tensor([out["config_B_pc_stable_sharded"]["wall_s"], dt], dtype=torch.float64, device=device)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        out["config_B_pc_stable_sharded"]["wall_s"] = float(mx[0])
        out["config_B_pc_stable_sharded"]["tests_per_sec"] = tests / float(mx[0])
        out["config_E_boot_strength_sharded"]["wall_s"] = float(mx[1])
        out["config_E_boot_strength_sharded"]["replicates_per_sec"] = 256 / float(mx[1])
    return out


def multi_device_context_check(units, lev):
    """One context on two devices against one on a single device (the split inside
    run_request / bnl_block), on a 50 000-row sample: identical outputs expected."""
    from bnlearn_b200 import Engine
    _, cols, _ = synth.alarm_shaped(50_000, 7)
    xs, ys, zs = [u[0] for u in units], [u[1] for u in units], [u[2] for u in units]
    res = []
    for devs in ([0], [0, 1]):
        with Engine(devs) as e:
            d = e.register(cols, lev)
            res.append(d.ci_discrete("mi", xs, ys, zs) + d.score_discrete("bde", xs, zs, 1.0))
            d.release()
    same = all(np.array_equal(a, b) for a, b in zip(res[0], res[1]))
    return "identical" if same else "DIFFERENT"


# --------------------------------------------------------------------------

def emit(line):
    """The JSON line is the ONLY thing written to the real stdout: libraries that
    print from C (NCCL's version banner) had fd 1 redirected to stderr in main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    # In the timed end-to-end steps only rank 0 uploads (the others receive the
    # packed columns by NCCL broadcast), so it gets every core for packing; the
    # other ranks upload once, untimed, and take a share.
    if "BNL_PACK_THREADS" not in os.environ:
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
        ncores = os.cpu_count() or 8
        os.environ["BNL_PACK_THREADS"] = str(ncores if rank == 0 else max(1, ncores // max(1, local_world)))

    if args.impl == "reference" and rank != 0:
        return 0

    names, codes, lev = synth.alarm_shaped_cached(args.rows)
    units = workloads.pc_stable_units(names, lev, max_level=2)
    n_units = len(units)
    workload = ("pc.stable(test='mi') CI-test batch, levels 0-2 (%d tests), alarm-shaped synthetic "
                "37 factors x %d rows, seed 20261106" % (n_units, args.rows))
    base = {"metric": "ci_tests_per_sec", "unit": "tests/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload, "rows": args.rows, "columns": len(names),
                       "tests_per_step": n_units,
                       "l2": "packed data set 370 MB > 126 MB L2, no explicit flush",
                       "sharding": "strong: the one fixed test list is cut into N contiguous blocks of about "
                                   "equal estimated cost, one per rank (each on its own replica of the data set); "
                                   "statistics gathered in index order (one NCCL all-gather) inside the timed region"}}

    cores = os.cpu_count() or 1

    # ---------------- reference arm: the reference's C on the host cores -----
    if args.impl == "reference":
        cols32 = [np.ascontiguousarray(c.astype(np.int32)) for c in codes]
        # a step = the whole 1081-test batch unless --cpu-sample bounds it; the workers
        # are forked and warm before the clock starts (cpu_run), tests are handed out
        # one at a time so that the cores stay balanced.
        sample = args.cpu_sample or n_units
        rng = np.random.default_rng(1)
        pick = [units[i] for i in rng.choice(n_units, size=min(sample, n_units), replace=False)]
        for _ in range(min(args.warmup, 1)):
            cpu_run(cols32, lev, pick[:cores * 4], cores)
        kind, total = "reference", 0.0
        for _ in range(args.steps):
            _, secs, kind = cpu_run(cols32, lev, pick, cores)
            total += secs
        dt = total / args.steps
        val = len(pick) / dt
        line = dict(base)
        line.update({"impl": "reference", "value": val, "ms_per_step": dt * 1e3, "n_gpus": args.gpus,
                     "cpu_baseline": {"value": val, "unit": "tests/s", "cores": cores, "kind": kind,
                                      "sample": "%d of the %d tests of one step, per step" % (len(pick), n_units)},
                     "e2e": {"value": val, "unit": "tests/s", "h2d_bytes_per_step": 0,
                             "d2h_bytes_per_step": 0},
                     "gpu_launches": 0})
        emit(line)
        return 0

    # ---------------- our arm ------------------------------------------------
    import torch
    import torch.distributed as dist
    from bnlearn_b200 import Engine, csr
    from bnlearn_b200.replicate import Gatherer, cost_blocks

    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # host copy of the data frame as the reference holds it (int32 factors), pinned;
    # a pageable copy too (R's own vectors are pageable: tabular_from_SEXP borrows them).
    ncol, rows = codes.shape
    pinned = torch.empty((ncol, rows), dtype=torch.int32, pin_memory=True)
    host = pinned.numpy()
    np.copyto(host, codes, casting="unsafe")
    cols32 = [host[i] for i in range(ncol)]

    # strong scaling: rank r evaluates units [b, e) of the one list. Blocks are cut by
    # estimated cost: the marginal tests come out of one tensor-core contraction and
    # cost ~1/8 of a conditional test each.
    n_marginal = sum(1 for u in units if not u[2])          # they lead the list
    bounds = cost_blocks([1.0 if u[2] else 0.125 for u in units], world, together=n_marginal)
    b, e = bounds[rank], bounds[rank + 1]
    my_units = units[b:e]
    gather = Gatherer(bounds, 3, device)
    xs = np.array([u[0] for u in my_units], dtype=np.int32)
    ys = np.array([u[1] for u in my_units], dtype=np.int32)
    z = csr([u[2] for u in my_units])
    unit_bytes = float(args.rows) * float(sum(2 + len(u[2]) for u in my_units))
    my_cols = sorted({c for u in my_units for c in (u[0], u[1], *u[2])})

    def step(ds):
        st, df, pv = ds.ci_discrete("mi", xs, ys, z)
        return gather(np.stack([st, df, pv], axis=1))

    # clocks are sampled from the warm-up to the end of the end-to-end loop (the
    # timed steps alone can be shorter than one nvidia-smi sampling period).
    sampler = ClockSampler(local_rank)
    sampler.start()
    eng = Engine([local_rank])
    ds = eng.register(cols32, lev)
    for _ in range(max(args.warmup, 3)):
        step(ds)
    barrier()
    count_ms, launches = [], 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        full = step(ds)
        prof = eng.last_call_profile()
        count_ms.append(prof["count_ms"])
        launches += prof["launches"]
    barrier()
    dt = (time.perf_counter() - t0) / args.steps
    outputs_sha1 = hashlib.sha1(np.ascontiguousarray(full).tobytes()).hexdigest()[:16]

    # the same list, every rank the whole of it on its replica (round 1's weak-scaling figure)
    replica = None
    if world > 1:
        xa = np.array([u[0] for u in units], dtype=np.int32)
        ya = np.array([u[1] for u in units], dtype=np.int32)
        za = csr([u[2] for u in units])
        ds.ci_discrete("mi", xa, ya, za)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ds.ci_discrete("mi", xa, ya, za)
        barrier()
        replica = (time.perf_counter() - t0) / args.steps

    # end to end: host buffers in, p-values out, H2D + packing inside the step. With
    # several ranks the job has ONE data frame: rank 0 uploads it and the packed
    # columns reach the other GPUs by an NCCL broadcast over NVLink (370 MB instead
    # of another 1.48 GB through the shared host memory per rank), every step.
    if world > 1:
        from bnlearn_b200.replicate import broadcast_dataset

        def bring_up(src_cols):
            return broadcast_dataset(eng, src_cols if rank == 0 else [None] * ncol, lev, rows, src=0)
    else:
        def bring_up(src_cols):
            return eng.register(src_cols, lev)

    def e2e_loop(src_cols):
        d2 = bring_up(src_cols); step(d2); d2.release()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            d2 = bring_up(src_cols)
            step(d2)
            d2.release()
        barrier()
        return (time.perf_counter() - t0) / args.steps

    dt_e2e = e2e_loop(cols32)
    pageable = [np.array(c) for c in cols32] if rank == 0 else cols32
    dt_e2e_pageable = e2e_loop(pageable)
    del pageable

    # N > 1: the same step with the ROWS sharded (reported beside the replicated one). Every rank
    # uploads and packs 1/N of the rows, counts all 1081 tables on them, the int32 tables (137 KB)
    # are summed on rank 0 by one NCCL reduce and rank 0 turns them into statistics with the same
    # epilogue (bnl_counts + bnl_stat_from_tables): bit-identical outputs.
    dt_e2e_rows, rows_sha1 = None, None
    if world > 1:
        xa = np.array([u[0] for u in units], dtype=np.int32)
        ya = np.array([u[1] for u in units], dtype=np.int32)
        za = csr([u[2] for u in units])
        llx, lly, llz, nz, toff = ds.table_shapes(xa, ya, za)
        h_tab = torch.zeros(int(toff[-1]), dtype=torch.int32).pin_memory()
        d_tab = torch.zeros(int(toff[-1]), dtype=torch.int32, device=device)
        ra, rb_ = rows * rank // world, rows * (rank + 1) // world
        my_rows = [c[ra:rb_] for c in cols32]

        def rows_step():
            d3 = eng.register(my_rows, lev)
            d3.counts_flat(xa, ya, za, toff, h_tab.numpy())
            d_tab.copy_(h_tab, non_blocking=True)
            dist.reduce(d_tab, dst=0)
            out = None
            if rank == 0:
                h_tab.copy_(d_tab, non_blocking=True)
                torch.cuda.current_stream().synchronize()
                out = eng.stat_from_flat("mi", llx, lly, llz, nz, h_tab.numpy(), toff, nrows=rows)
            d3.release()
            return out

        rows_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            out = rows_step()
        barrier()
        dt_e2e_rows = (time.perf_counter() - t0) / args.steps
        if rank == 0:
            rows_sha1 = hashlib.sha1(np.ascontiguousarray(np.stack(out, axis=1)).tobytes()).hexdigest()[:16]
    clocks = sampler.stop()

    # diagnostic: plain pinned host -> device copy rate of this box.
    dev_buf = torch.empty((8, rows), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dev_buf.copy_(pinned[:8], non_blocking=True)
    torch.cuda.synchronize()
    h2d_gbs = 8 * rows * 4 / (time.perf_counter() - t0) / 1e9
    del dev_buf

    kms_local = float(np.mean(count_ms))
    if world > 1:
        t = torch.tensor([dt, dt_e2e, dt_e2e_pageable, replica, kms_local, dt_e2e_rows], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt, dt_e2e, dt_e2e_pageable, replica, kms, dt_e2e_rows = (float(v) for v in t)
        lt = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(lt)
        launches = int(lt[0])
    else:
        kms = kms_local

    line = None
    if rank == 0:
        peak, peak_src = hbm_peak()
        # parity of the timed outputs: a stratified sample against the reference's C
        rng = np.random.default_rng(3)
        picked, worst = [], 0.0
        for s in (0, 1, 2):
            idx = [i for i, u in enumerate(units) if len(u[2]) == s]
            picked += [int(i) for i in rng.choice(idx, size=8, replace=False)]
        for i in picked:
            rs, rdf, rp = reference_ci(cols32, lev, units[i])
            for got, want in ((full[i, 0], rs), (full[i, 2], rp)):
                if got != want:
                    worst = max(worst, abs(got - want) / max(abs(got), abs(want), 1e-300))
            if full[i, 1] != rdf:
                worst = float("inf")
        comp_local = float(args.rows) * len(my_cols)
        # (no atomic_frac for the mixed step: its marginal tests are counted on the tensor cores; the
        #  conditional tests' figure is copied in from their own class below)
        roof = hbm_class("config B: %d of the %d tests on rank 0 (levels 0-2 of pc.stable, mixed)" % (e - b, n_units),
                         kms_local, comp_local, unit_bytes, e - b, args.rows, peak, atomic=False,
                         note="the step's counting launches together (CUDA events around them on the engine's "
                              "stream); %.0f units per distinct column, so the >= 50 %% HBM target does not apply "
                              "to this batch: L2 serves the column reuse and the shared-memory atomic rate binds"
                              % (unit_bytes / comp_local))
        roof.pop("class")
        roof.update({"kernel": "counting launches of one step (rank 0)", "peak_source": peak_src})
        roof.update(measured_traffic(world == 1 and args.rows == 10_000_000, comp_local))
        line = dict(base)
        line.update({
            "value": n_units / dt, "ms_per_step": dt * 1e3, "clocks": clocks,
            "e2e": {"value": n_units / dt_e2e, "unit": "tests/s",
                    "h2d_bytes_per_step": int(ncol * rows * 4 + xs.nbytes + ys.nbytes + z[0].nbytes + z[1].nbytes),
                    "d2h_bytes_per_step": int(3 * 8 * n_units), "ms_per_step": dt_e2e * 1e3,
                    "source": "pinned int32 columns",
                    "pageable": {"value": n_units / dt_e2e_pageable, "ms_per_step": dt_e2e_pageable * 1e3,
                                 "source": "pageable int32 columns (what R hands over)"},
                    "replication": ("single GPU" if world == 1 else
                                    "rank 0 uploads the data frame (h2d_bytes_per_step), the packed columns "
                                    "(%d bytes) are broadcast to the other ranks with NCCL every step"
                                    % (ncol * ((rows + 256) // 256 * 256))),
                    "pinned_h2d_gbs_probe": h2d_gbs},
            "gpu_launches": launches,
            "outputs_sha1": outputs_sha1,
            "parity_checked": len(picked), "parity_max_rel": worst,
            "parity_against": "oracle/_ref (c_fast_config + c_cchisqtest + pchisq), statistic / df / p-value of "
                              "the LAST timed step's gathered outputs, 8 tests per conditioning-set size",
            "roofline": roof,
        })
        if dt_e2e_rows is not None:
            # the row-sharded form of the same step, beside the replicated one: no faster on one node --
            # the upload is bound by the HOST cores that pack the int32 codes, and N ranks share them --
            # but it is the route for data that does not fit one GPU, and its outputs must be the same bits
            line["e2e"]["rows_sharded"] = {
                "value": n_units / dt_e2e_rows, "ms_per_step": dt_e2e_rows * 1e3, "outputs_sha1": rows_sha1,
                "identical_to_unit_sharded": bool(rows_sha1 == outputs_sha1),
                "how": "every rank registers 1/N of the rows from the pinned data frame, counts all %d tables on "
                       "them, one NCCL reduce of the int32 tables (%d bytes) to rank 0, statistics there "
                       "(bnl_counts + bnl_stat_from_tables)" % (n_units, 4 * int(toff[-1]))}
        if replica is not None:
            line["replicas"] = {"value": world * n_units / replica, "ms_per_step": replica * 1e3, "scaling": "weak",
                                "what": "every rank runs the full list on its own replica (round 1's figure)"}
    extras = not args.no_extras
    classes = []
    if extras:
        try:
            res = strong_extras(eng, ds, world, rank, device)
            if rank == 0:
                line["strong"] = res
        except Exception as exc:                      # the headline line must still be printed
            if rank == 0:
                line["strong"] = {"error": repr(exc)[:300]}
        if rank == 0 and world == 1:
            try:
                line["wall"] = wall_times(eng, ds, cols32, lev, cores, not args.no_cpu_baseline)
            except Exception as exc:
                line["wall"] = {"error": repr(exc)[:300]}
    # level 0 alone and levels 1-2 alone, for the per-class roofline
    if rank == 0:
        for nm, sel in (("config B: level 0, all 666 pairs (tcgen05 indicator cross-product)", [u for u in units if not u[2]]),
                        ("config B: levels 1-2, 415 conditional tests", [u for u in units if u[2]])):
            zx = csr([u[2] for u in sel])
            ax = np.array([u[0] for u in sel], dtype=np.int32)
            ay = np.array([u[1] for u in sel], dtype=np.int32)
            best = None
            for _ in range(4):
                ds.ci_discrete("mi", ax, ay, zx)
                c = eng.last_call_profile()["count_ms"]
                best = c if best is None else min(best, c)
            ncols_used = len({c for u in sel for c in (u[0], u[1], *u[2])})
            classes.append(hbm_class(nm, best, float(args.rows) * ncols_used,
                                     float(args.rows) * sum(2 + len(u[2]) for u in sel), len(sel), args.rows,
                                     hbm_peak()[0], atomic=bool(sel[0][2])))
    ds.release()
    if extras:
        try:
            res, ccls = config_c_extras(eng, world, rank, device,
                                        cpu_cores=0 if (args.no_cpu_baseline or world > 1) else cores,
                                        want_hc=world == 1)
            if rank == 0:
                line["score_deltas"] = res
                classes += ccls
        except Exception as exc:
            if rank == 0:
                line["score_deltas"] = {"error": repr(exc)[:300]}
        if rank == 0 and world == 1:
            try:
                line["gaussian"], gcls = config_d_extras(eng)
                classes.append(gcls)
            except Exception as exc:
                line["gaussian"] = {"error": repr(exc)[:300]}
    if rank == 0:
        line["roofline"]["classes"] = classes
        for c in classes:
            if c["class"].startswith("config B: levels 1-2") and "atomic_frac" in c:
                line["roofline"]["atomic_frac"] = c["atomic_frac"]
                line["roofline"]["atomic_ceiling_updates_per_s"] = ATOMIC_CEILING
                line["roofline"]["atomic_frac_of"] = ("the conditional tests of the step, timed alone (classes[1]): rows x units "
                                                      "per second over the shared-memory update ceiling; level 0 has no "
                                                      "atomics, and the units of <= 32 cells (a third) count two rows per update")
        if world > 1 and torch.cuda.device_count() >= 2:
            try:
                line["multi_device_context_check"] = multi_device_context_check(units[::7], lev)
            except Exception as exc:
                line["multi_device_context_check"] = repr(exc)[:200]
        if world == 1 and not args.no_cpu_baseline:
            sample = args.cpu_sample or max(cores * 64, 1024)       # one whole step here: ~4 s on 16 cores
            rng = np.random.default_rng(1)
            pick = [units[i] for i in rng.choice(n_units, size=min(sample, n_units), replace=False)]
            val, secs, kind = cpu_run(cols32, lev, pick, cores)
            # the reference as it ships is serial on this path: one core, a smaller sample
            sval, ssecs, _ = cpu_run(cols32, lev, pick[:max(16, len(pick) // 16)], 1)
            line["cpu_baseline"] = {"value": val, "unit": "tests/s", "cores": cores, "kind": kind,
                                    "sample": "%d of the %d tests of one step (%.1f s)" % (len(pick), n_units, secs),
                                    "serial": {"value": sval, "cores": 1,
                                               "sample": "%d tests (%.1f s)" % (max(16, len(pick) // 16), ssecs)}}
        try:
            r = subprocess.run(["R", "--version"], capture_output=True, text=True, timeout=10)
            line["r_front_end"] = r.stdout.splitlines()[0] if r.returncode == 0 else "unavailable on this box"
        except (OSError, subprocess.SubprocessError):
            line["r_front_end"] = "R front-end timing: unavailable on this box (no R interpreter)"
        emit(line)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
