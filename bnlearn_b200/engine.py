"""Python face of the C ABI (include/bnl_b200.h): Engine = bnl_ctx, DataSet =
bnl_data. Everything here goes through libbnl_b200.so; numpy only marshals the
caller's host buffers.

Column conventions are the reference's: factor columns are int32, 1-based, with
NA = INT_MIN; numeric columns are float64.
"""
import ctypes as C

import numpy as np

from . import _lib

TESTS = {"mi": 1, "mi-adf": 2, "x2": 3, "x2-adf": 4, "jt": 10, "mi-sh": 40, "cor": 20, "zf": 21, "mi-g": 22, "mi-cg": 30, "mi-g-sh": 41}
SCORES = {"loglik": 1, "aic": 3, "bic": 4, "ebic": 5, "bde": 6, "bds": 7, "bdj": 8, "k2": 9, "mbde": 10, "fnml": 11, "qnml": 12, "nal": 13, "pnal": 14,
          "loglik-g": 100, "aic-g": 102, "bic-g": 103, "ebic-g": 104, "bge": 105, "nal-g": 106, "pnal-g": 107,
          "loglik-cg": 200, "aic-cg": 202, "bic-cg": 203, "ebic-cg": 204, "nal-cg": 205, "pnal-cg": 206}
MC_TESTS = {"mc-mi": 50, "mc-x2": 51, "sp-mi": 52, "sp-x2": 53, "mc-jt": 54, "smc-mi": 60, "smc-x2": 61,
            "smc-jt": 62, "mc-cor": 70, "mc-mi-g": 71, "mc-zf": 72, "smc-cor": 80, "smc-mi-g": 81, "smc-zf": 82}
DISCRETE_TESTS = ("mi", "mi-adf", "x2", "x2-adf", "jt", "mi-sh")
GAUSSIAN_TESTS = ("cor", "zf", "mi-g", "mi-g-sh")
DISCRETE_SCORES = ("loglik", "aic", "bic", "ebic", "bde", "mbde", "bds", "bdj", "k2", "fnml", "qnml", "nal", "pnal")
GAUSSIAN_SCORES = ("loglik-g", "aic-g", "bic-g", "ebic-g", "bge", "nal-g", "pnal-g")
NA_INT = np.iinfo(np.int32).min


class BnlError(RuntimeError):
    def __init__(self, code, message):
        RuntimeError.__init__(self, "bnl error %d: %s" % (code, message))
        self.code = code


def _ip(a):
    return a.ctypes.data_as(_lib.c_int_p)


def _dp(a):
    return a.ctypes.data_as(_lib.c_dbl_p)


def csr(lists):
    """list of index lists -> (offsets int32[n+1], indices int32[total])."""
    off = np.zeros(len(lists) + 1, dtype=np.int32)
    if len(lists):
        off[1:] = np.cumsum([len(l) for l in lists])
    idx = np.fromiter((i for l in lists for i in l), dtype=np.int32, count=int(off[-1]))
    if idx.size == 0:
        idx = np.zeros(1, dtype=np.int32)
    return off, idx


def _as_csr(z):
    if isinstance(z, tuple) and len(z) == 2 and isinstance(z[0], np.ndarray):
        return (np.ascontiguousarray(z[0], dtype=np.int32),
                np.ascontiguousarray(z[1], dtype=np.int32))
    return csr(z)


class Engine:
    """One bnl_ctx. devices: list of CUDA ordinals (None = the current device)."""

    def __init__(self, devices=None):
        self._lib = _lib.load()
        self._ctx = C.c_void_p()
        # remembered so that helpers which open more contexts from other host threads
        # (learners.boot_arc_counts) land on the same GPU: a new thread's current device is 0
        self.devices = None if devices is None else [int(d) for d in devices]
        if devices is None:
            rc = self._lib.bnl_init(C.byref(self._ctx), None, 0)
        else:
            d = np.asarray(devices, dtype=np.int32)
            rc = self._lib.bnl_init(C.byref(self._ctx), _ip(d), len(d))
        if rc != 0:
            msg = self._lib.bnl_last_error(self._ctx).decode() if self._ctx else "bnl_init failed"
            if self._ctx:
                self._lib.bnl_shutdown(self._ctx)
                self._ctx = None
            raise BnlError(rc, msg)

    def _check(self, rc):
        if rc != 0:
            raise BnlError(rc, self._lib.bnl_last_error(self._ctx).decode())

    def stat_from_flat(self, name, llx, lly, llz, nz, flat, off, arg=0.0, nrows=0):
        """bnl_stat_from_tables on a flat int32 buffer (tables at off[u], shapes as
        DataSet.table_shapes gives them) -> three float64 arrays."""
        kind = 0 if name in TESTS else 1
        code = TESTS[name] if kind == 0 else SCORES[name]
        n = len(llx)
        o = [np.zeros(max(n, 1)) for _ in range(3)]
        self._check(self._lib.bnl_stat_from_tables(self._ctx, kind, code, float(arg), int(nrows), n, _ip(llx),
                                                   _ip(lly), _ip(llz), _ip(nz),
                                                   flat.ctypes.data_as(_lib.c_i32_p),
                                                   off.ctypes.data_as(_lib.c_i64_p), _dp(o[0]), _dp(o[1]),
                                                   _dp(o[2])))
        return o[0][:n], o[1][:n], o[2][:n]

    def stat_from_tables(self, name, tables, nz, arg=0.0, nrows=0):
        """Statistics from count tables shaped (llz, llx, lly) as DataSet.counts
        returns them (bnl_stat_from_tables): name is a discrete test -> (stat, df,
        p.value) or a discrete score -> (score, nparams). nz[u] = number of
        conditioning columns / parents of unit u."""
        kind = 0 if name in TESTS else 1
        code = TESTS[name] if kind == 0 else SCORES[name]
        n = len(tables)
        shp = np.array([t.shape for t in tables], dtype=np.int32).reshape(n, 3)
        llz, llx, lly = (np.ascontiguousarray(shp[:, i]) for i in range(3))
        off = np.zeros(n + 1, dtype=np.int64)
        off[1:] = np.cumsum([t.size for t in tables])
        flat = np.concatenate([np.ascontiguousarray(t, dtype=np.int32).ravel() for t in tables]) if n else \
            np.zeros(1, dtype=np.int32)
        nzs = np.ascontiguousarray(nz, dtype=np.int32)
        o = [np.zeros(max(n, 1)) for _ in range(3)]
        self._check(self._lib.bnl_stat_from_tables(self._ctx, kind, code, float(arg), int(nrows), n, _ip(llx),
                                                   _ip(lly), _ip(llz), _ip(nzs),
                                                   flat.ctypes.data_as(_lib.c_i32_p),
                                                   off.ctypes.data_as(_lib.c_i64_p), _dp(o[0]), _dp(o[1]),
                                                   _dp(o[2])))
        return (o[0][:n], o[1][:n], o[2][:n]) if kind == 0 else (o[0][:n], o[1][:n])

    def register(self, columns, nlevels=None):
        """Upload a data frame. nlevels[i] > 0 marks column i as a factor."""
        n = len(columns)
        nlevels = [0] * n if nlevels is None else [int(v) for v in nlevels]
        cols = []
        for i, c in enumerate(columns):
            cols.append(np.ascontiguousarray(c, dtype=np.int32 if nlevels[i] > 0 else np.float64))
        nobs = len(cols[0]) if n else 0
        for c in cols:
            if len(c) != nobs:
                raise ValueError("columns differ in length")
        ptrs = (C.c_void_p * n)(*[c.ctypes.data for c in cols])
        isnum = np.array([0 if v > 0 else 1 for v in nlevels], dtype=np.int32)
        lv = np.array(nlevels, dtype=np.int32)
        handle = C.c_uint64()
        self._check(self._lib.bnl_data_register(self._ctx, n, ptrs, _ip(isnum), _ip(lv), nobs,
                                                C.byref(handle)))
        return DataSet(self, handle.value, nobs, nlevels)

    def alloc(self, nlevels, nobs):
        """A data set of the given shape whose packed columns the caller fills on the
        device (bnl_data_alloc); see DataSet.device_view / commit."""
        lv = np.array([int(v) for v in nlevels], dtype=np.int32)
        isnum = np.array([0 if v > 0 else 1 for v in lv], dtype=np.int32)
        handle = C.c_uint64()
        self._check(self._lib.bnl_data_alloc(self._ctx, len(lv), _ip(isnum), _ip(lv), int(nobs), C.byref(handle)))
        return DataSet(self, handle.value, int(nobs), [int(v) for v in lv])

    def flush(self):
        self._check(self._lib.bnl_data_flush(self._ctx))

    def last_call_profile(self):
        """Device-side timing of the most recent compute call:
        dict(total_ms, count_ms, epilogue_ms, launches)."""
        t, c, e, n = C.c_double(), C.c_double(), C.c_double(), C.c_int()
        self._check(self._lib.bnl_last_call_profile(self._ctx, C.byref(t), C.byref(c), C.byref(e),
                                                    C.byref(n)))
        return {"total_ms": t.value, "count_ms": c.value, "epilogue_ms": e.value, "launches": n.value}

    def close(self):
        if self._ctx:
            self._lib.bnl_shutdown(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DataSet:
    def __init__(self, engine, handle, nobs, nlevels):
        self.engine, self.handle, self.nobs, self.nlevels = engine, handle, nobs, list(nlevels)
        self.ncols = len(nlevels)

    def resample(self, rows):
        """A new DataSet holding rows[...] (0-based, with repeats) of this one,
        gathered on the device (bnl_data_resample)."""
        e = self.engine
        r = np.ascontiguousarray(rows, dtype=np.int32)
        h = C.c_uint64(0)
        e._check(e._lib.bnl_data_resample(e._ctx, self.handle, len(r), _ip(r), C.byref(h)))
        return DataSet(e, h.value, len(r), self.nlevels)

    def device_view(self, dev_index=0):
        """(u8 pointer, bytes, f64 pointer, bytes) of the packed blocks on one device."""
        e = self.engine
        u8, f64 = C.c_void_p(), C.c_void_p()
        nu, nf = C.c_size_t(), C.c_size_t()
        e._check(e._lib.bnl_data_device_view(e._ctx, self.handle, int(dev_index), C.byref(u8), C.byref(nu),
                                             C.byref(f64), C.byref(nf)))
        return u8.value or 0, nu.value, f64.value or 0, nf.value

    def commit(self):
        """Finish a data set made by Engine.alloc once its blocks are filled (bnl_data_commit)."""
        e = self.engine
        e._check(e._lib.bnl_data_commit(e._ctx, self.handle))

    def release(self):
        if self.handle:
            e = self.engine
            e._check(e._lib.bnl_data_release(e._ctx, self.handle))
            self.handle = 0

    # ---- discrete ---------------------------------------------------------
    def counts(self, x, y, z):
        """Raw tables; returns a list of (llz, llx, lly) int32 arrays. y[u] = -1: no y."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        shapes, off = [], np.zeros(len(x) + 1, dtype=np.int64)
        for u in range(len(x)):
            llz = 1
            for j in zi[zo[u]:zo[u + 1]]:
                llz *= self.nlevels[j]
            shp = (llz, self.nlevels[x[u]], self.nlevels[y[u]] if y[u] >= 0 else 1)
            shapes.append(shp)
            off[u + 1] = off[u] + shp[0] * shp[1] * shp[2]
        tab = np.zeros(max(int(off[-1]), 1), dtype=np.int32)
        e._check(e._lib.bnl_counts(e._ctx, self.handle, len(x), _ip(x), _ip(y), _ip(zo), _ip(zi),
                                   tab.ctypes.data_as(_lib.c_i32_p),
                                   off.ctypes.data_as(_lib.c_i64_p)))
        return [tab[off[u]:off[u + 1]].reshape(shapes[u]) for u in range(len(x))]

    def table_shapes(self, x, y, z):
        """(llx, lly, llz, nz, off) int arrays of the tables counts() would return, for the flat calls."""
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        lev = np.asarray(self.nlevels, dtype=np.int64)
        n = len(x)
        llx = lev[x].astype(np.int32)
        lly = np.where(y >= 0, lev[np.maximum(y, 0)], 1).astype(np.int32)
        llz = np.ones(n, dtype=np.int64)
        for u in range(n):
            for j in zi[zo[u]:zo[u + 1]]:
                llz[u] *= lev[j]
        off = np.zeros(n + 1, dtype=np.int64)
        off[1:] = np.cumsum(llx.astype(np.int64) * lly * llz)
        return llx, lly, llz.astype(np.int32), np.diff(zo).astype(np.int32), off

    def counts_flat(self, x, y, z, off, out):
        """bnl_counts into the caller's flat int32 buffer `out` (tables at off[u]): no per-unit arrays."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        e._check(e._lib.bnl_counts(e._ctx, self.handle, len(x), _ip(x), _ip(y), _ip(zo), _ip(zi),
                                   out.ctypes.data_as(_lib.c_i32_p), off.ctypes.data_as(_lib.c_i64_p)))
        return out

    def ci_discrete(self, test, x, y, z):
        """Batch of discrete CI tests -> (statistic, df, p.value) arrays."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        n = len(x)
        st, df, pv = (np.zeros(max(n, 1)) for _ in range(3))
        e._check(e._lib.bnl_ci_discrete(e._ctx, self.handle, TESTS[test], n, _ip(x), _ip(y), _ip(zo),
                                        _ip(zi), _dp(st), _dp(df), _dp(pv)))
        return st[:n], df[:n], pv[:n]

    def set_interventions(self, node, mask_col):
        """mBDe: `mask_col` (a one-level factor, NA on the rows where `node` was
        intervened upon) masks every unit of `node`; -1 clears."""
        e = self.engine
        e._check(e._lib.bnl_data_set_interventions(e._ctx, self.handle, int(node), int(mask_col)))

    def _unif(self, seed, unif):
        """(function pointer, state pointer, keep-alive) of the uniform stream: a
        ctypes callback `unif(state) -> float`, or the library's splitmix64 on `seed`."""
        lib = self.engine._lib
        if unif is not None:
            cb = C.CFUNCTYPE(C.c_double, C.c_void_p)(unif)
            return C.cast(cb, C.c_void_p), None, cb
        state = C.c_uint64(int(seed))
        return C.cast(lib.bnl_unif_splitmix64, C.c_void_p), C.cast(C.pointer(state), C.c_void_p), state

    def ci_discrete_mc(self, test, x, y, z, B, alpha=1.0, seed=1, unif=None):
        """Batch of discrete permutation / semiparametric tests -> (statistic, df,
        p.value); the uniform stream is `unif` (a Python callable taking the state
        pointer) or splitmix64 seeded with `seed`, consumed test after test."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        n = len(x)
        st, df, pv = (np.zeros(max(n, 1)) for _ in range(3))
        fn, state, keep = self._unif(seed, unif)
        e._check(e._lib.bnl_ci_discrete_mc(e._ctx, self.handle, MC_TESTS[test], int(B), float(alpha), n, _ip(x),
                                           _ip(y), _ip(zo), _ip(zi), fn, state, _dp(st), _dp(df), _dp(pv)))
        self.last_stream_state = keep.value if unif is None else None     # splitmix64 state after the call
        return st[:n], df[:n], pv[:n]

    def ci_gaussian_mc(self, test, x, y, z, B, alpha=1.0, seed=1, unif=None):
        """Batch of Gaussian permutation tests -> (statistic, p.value)."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        n = len(x)
        st, pv = np.zeros(max(n, 1)), np.zeros(max(n, 1))
        fn, state, keep = self._unif(seed, unif)
        e._check(e._lib.bnl_ci_gaussian_mc(e._ctx, self.handle, MC_TESTS[test], int(B), float(alpha), n, _ip(x),
                                           _ip(y), _ip(zo), _ip(zi), fn, state, _dp(st), _dp(pv)))
        self.last_stream_state = keep.value if unif is None else None
        return st[:n], pv[:n]

    def score_discrete(self, score, target, parents, arg=0.0):
        """Batch of node scores -> (score, nparams); arg = k (aic/bic) or iss (bde/bds)."""
        e = self.engine
        t = np.ascontiguousarray(target, dtype=np.int32)
        po, pi = _as_csr(parents)
        n = len(t)
        sc, npar = np.zeros(max(n, 1)), np.zeros(max(n, 1))
        e._check(e._lib.bnl_score_discrete(e._ctx, self.handle, SCORES[score], float(arg), n, _ip(t),
                                           _ip(po), _ip(pi), _dp(sc), _dp(npar)))
        return sc[:n], npar[:n]

    # ---- Gaussian ---------------------------------------------------------
    def ci_gaussian(self, test, x, y, z):
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        n = len(x)
        st, df, pv = (np.zeros(max(n, 1)) for _ in range(3))
        e._check(e._lib.bnl_ci_gaussian(e._ctx, self.handle, TESTS[test], n, _ip(x), _ip(y), _ip(zo),
                                        _ip(zi), _dp(st), _dp(df), _dp(pv)))
        return st[:n], df[:n], pv[:n]

    def ci_gaussian_allpairs(self, test, alpha, cols=None, want_p=True, counter=None):
        """bnl_ci_gaussian_allpairs: every pair of `cols` (default: all numeric columns)
        tested marginally on the device. Returns (keep [m, 2] positions in cols of the
        pairs with p <= alpha, in row-major order; p-value matrix or None)."""
        e = self.engine
        if cols is None:
            cols = [i for i, v in enumerate(self.nlevels) if v <= 0]
        c = np.ascontiguousarray(cols, dtype=np.int32)
        n = len(c)
        pm = np.zeros((n, n)) if want_p else None
        cap = max(n * (n - 1) // 2, 1)
        ki, kj = np.zeros(cap, dtype=np.int32), np.zeros(cap, dtype=np.int32)
        nk = C.c_longlong(0)
        cnt = C.c_double(0.0)
        e._check(e._lib.bnl_ci_gaussian_allpairs(e._ctx, self.handle, TESTS[test], float(alpha), n, _ip(c),
                                                 None if pm is None else _dp(pm), _ip(ki), _ip(kj), cap,
                                                 C.byref(nk), C.byref(cnt)))
        if counter is not None:
            counter.add(cnt.value)
        return np.stack([ki[:nk.value], kj[:nk.value]], axis=1), pm

    def score_gaussian(self, score, target, parents, k=0.0):
        e = self.engine
        t = np.ascontiguousarray(target, dtype=np.int32)
        po, pi = _as_csr(parents)
        n = len(t)
        sc, npar = np.zeros(max(n, 1)), np.zeros(max(n, 1))
        e._check(e._lib.bnl_score_gaussian(e._ctx, self.handle, SCORES[score], float(k), n, _ip(t),
                                           _ip(po), _ip(pi), _dp(sc), _dp(npar)))
        return sc[:n], npar[:n]

    def hc(self, score, arg=0.0, nnodes=None, start=None, blmat=None, wlmat=None, maxp=np.inf,
           max_iter=0, counter=None):
        """bnl_hc: the whole greedy search in one library call. Returns
        (amat, per-node scores, iterations); `counter` (callsites.TestCounter) is
        advanced by the number of score deltas computed."""
        e = self.engine
        n = len(self.nlevels) if nnodes is None else int(nnodes)

        def mat(m):
            return None if m is None else np.asfortranarray(m, dtype=np.int32)
        s, b, w = mat(start), mat(blmat), mat(wlmat)
        amat = np.zeros((n, n), dtype=np.int32, order="F")
        scores = np.zeros(n)
        iters = C.c_longlong(0)
        cnt = C.c_double(0.0)
        e._check(e._lib.bnl_hc(e._ctx, self.handle, SCORES[score], float(arg), n,
                               None if s is None else _ip(s), None if b is None else _ip(b),
                               None if w is None else _ip(w), float(maxp), int(max_iter), _ip(amat),
                               _dp(scores), C.byref(iters), C.byref(cnt)))
        if counter is not None:
            counter.add(cnt.value)
        return amat, scores, iters.value

    def pc_stable(self, test, alpha=0.05, nnodes=None, max_sx=-1, counter=None):
        """bnl_pc_stable: skeleton, p-values and separating sets in one call. Returns
        (sorted edge list, {(i, j): separating set}, {(i, j): p-value}) like learners.pc_stable."""
        e = self.engine
        n = len(self.nlevels) if nnodes is None else int(nnodes)
        npairs = n * (n - 1) // 2
        adj = np.zeros((n, n), dtype=np.int32, order="F")
        pv = np.zeros((n, n), order="F")
        nds = np.zeros(max(npairs, 1), dtype=np.int32)
        off = np.zeros(npairs + 1, dtype=np.int64)
        cap = max(npairs * 4, 1024)
        cnt = C.c_double(0.0)
        while True:
            idx = np.zeros(cap, dtype=np.int32)
            cnt.value = 0.0
            rc = e._lib.bnl_pc_stable(e._ctx, self.handle, TESTS[test], float(alpha), n, int(max_sx), _ip(adj),
                                      _dp(pv), _ip(nds), off.ctypes.data_as(C.POINTER(C.c_longlong)), _ip(idx),
                                      cap, C.byref(cnt))
            if rc != 0 and b"dsep_idx too small" in e._lib.bnl_last_error(e._ctx) and cap < (1 << 28):
                cap *= 8
                continue
            e._check(rc)
            break
        if counter is not None:
            counter.add(cnt.value)
        pairs = [(i, j) for i in range(n) for j in range(i + 1, n)]
        skeleton = [p for k, p in enumerate(pairs) if adj[p[0], p[1]]]
        dsep = {p: [int(v) for v in idx[off[k]:off[k + 1]]] for k, p in enumerate(pairs) if nds[k] >= 0}
        pvals = {p: float(pv[p[0], p[1]]) for p in pairs}
        return skeleton, dsep, pvals

    def tabu(self, score, arg=0.0, nnodes=None, start=None, blmat=None, wlmat=None, maxp=np.inf,
             max_iter=0, tabu=10, max_tabu=None, counter=None):
        """bnl_tabu: tabu search in one library call; returns (amat, scores, iterations)."""
        e = self.engine
        n = len(self.nlevels) if nnodes is None else int(nnodes)

        def mat(m):
            return None if m is None else np.asfortranarray(m, dtype=np.int32)
        s, b, w = mat(start), mat(blmat), mat(wlmat)
        amat = np.zeros((n, n), dtype=np.int32, order="F")
        scores = np.zeros(n)
        iters = C.c_longlong(0)
        cnt = C.c_double(0.0)
        e._check(e._lib.bnl_tabu(e._ctx, self.handle, SCORES[score], float(arg), n,
                                 None if s is None else _ip(s), None if b is None else _ip(b),
                                 None if w is None else _ip(w), float(maxp), int(max_iter), int(tabu),
                                 int(tabu if max_tabu is None else max_tabu), _ip(amat), _dp(scores),
                                 C.byref(iters), C.byref(cnt)))
        if counter is not None:
            counter.add(cnt.value)
        return amat, scores, iters.value

    def mi_matrix(self, estimator, cols, cond=-1):
        """Pairwise mutual-information coefficients of `cols` (optionally given the
        factor `cond`) as a symmetric matrix with a zero diagonal."""
        e = self.engine
        c = np.ascontiguousarray(cols, dtype=np.int32)
        p = len(c)
        tri = np.zeros(max(p * (p - 1) // 2, 1))
        e._check(e._lib.bnl_mi_matrix(e._ctx, self.handle, TESTS[estimator], p, _ip(c), int(cond), _dp(tri)))
        m = np.zeros((p, p))
        iu = np.triu_indices(p, 1)
        m[iu] = tri[:len(iu[0])]
        return m + m.T

    def ci_cg(self, x, y, z):
        """mi-cg tests on mixed data -> (statistic, df, p.value) arrays (bnl_ci_cg)."""
        e = self.engine
        x = np.ascontiguousarray(x, dtype=np.int32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        zo, zi = _as_csr(z)
        n = len(x)
        st, df, pv = (np.zeros(max(n, 1)) for _ in range(3))
        e._check(e._lib.bnl_ci_cg(e._ctx, self.handle, n, _ip(x), _ip(y), _ip(zo), _ip(zi), _dp(st), _dp(df),
                                  _dp(pv)))
        return st[:n], df[:n], pv[:n]

    def score_cg(self, score, target, parents, k=0.0):
        """Conditional Gaussian node scores on mixed data (bnl_score_cg)."""
        e = self.engine
        t = np.ascontiguousarray(target, dtype=np.int32)
        po, pi = _as_csr(parents)
        n = len(t)
        sc, npar = np.zeros(max(n, 1)), np.zeros(max(n, 1))
        e._check(e._lib.bnl_score_cg(e._ctx, self.handle, SCORES[score], float(k), n, _ip(t), _ip(po), _ip(pi),
                                     _dp(sc), _dp(npar)))
        return sc[:n], npar[:n]

    def set_ebic_gamma(self, gamma):
        e = self.engine
        e._check(e._lib.bnl_set_ebic_gamma(e._ctx, self.handle, float(gamma)))

    def set_bge_prior(self, iss_mu=1.0, iss_w=0.0, nu=None):
        """BGe hyper-parameters (iss_w <= 0: ncol + 2; nu None: the column means)."""
        e = self.engine
        v = None if nu is None else np.ascontiguousarray(nu, dtype=np.float64)
        e._check(e._lib.bnl_set_bge_prior(e._ctx, self.handle, float(iss_mu), float(iss_w),
                                          None if v is None else _dp(v)))

    def moments(self):
        """(means, covariance matrix) of the numeric columns, registration order."""
        e = self.engine
        p = sum(1 for v in self.nlevels if v <= 0)
        mean, cov = np.zeros(p), np.zeros((p, p), order="F")
        e._check(e._lib.bnl_gaussian_moments(e._ctx, self.handle, _dp(mean), _dp(cov)))
        return mean, cov

    def set_moments(self, nobs_total, mean, cov):
        """Serve every Gaussian test / score of this data set from pooled moments
        (rows sharded over GPUs: replicate.pool_moments)."""
        e = self.engine
        mean = np.ascontiguousarray(mean, dtype=np.float64)
        cov = np.asfortranarray(cov, dtype=np.float64)
        e._check(e._lib.bnl_data_set_moments(e._ctx, self.handle, int(nobs_total), _dp(mean), _dp(cov)))

    # ---- generic dispatch on the test / score label -------------------------
    def ci_test(self, test, x, y, z):
        return self.ci_discrete(test, x, y, z) if test in DISCRETE_TESTS else self.ci_gaussian(test, x, y, z)

    def score(self, score, target, parents, arg=0.0):
        if score in DISCRETE_SCORES:
            return self.score_discrete(score, target, parents, arg)
        if score.endswith("-cg"):
            return self.score_cg(score, target, parents, arg)
        return self.score_gaussian(score, target, parents, arg)
