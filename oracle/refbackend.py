"""A learners backend that evaluates every statistic with the REFERENCE's own
.Call entry points (oracle/_ref/libbnref.so through oracle/refapi.py), for
end-to-end comparisons with the CUDA backend. Test infrastructure only."""
import numpy as np

from oracle import refapi as ra, refbind as rb


class _Counter:
    @property
    def value(self):
        return rb.test_counter()

    def reset(self):
        rb.reset_test_counter()


class RefBackend:
    def __init__(self, columns, nlevels=None, B=None, seed=None):
        self.B = B                             # permutation tests: permutations per test ...
        if seed is not None:                   # ... drawn from the shim's stream
            rb.lib().minir_rng_seed(int(seed))
        self.names = ["V%d" % i for i in range(len(columns))]
        self.data = ra.RefData(columns, nlevels, self.names)
        self.counter = _Counter()
        self.counter.reset()
        self._calls = 0

    def _gc(self):
        # the mini-R arena only grows; recycle it between calls.
        self._calls += 1
        if self._calls % 64 == 0:
            rb.reset()
            self.data.invalidate()

    def _n(self, idx):
        return [self.names[i] for i in idx]

    def indep_test(self, xs, y, sx, test, alpha=0.05):
        self._gc()
        return ra.indep_test(self.data, self._n(xs), self.names[y], self._n(sx), test, alpha, B=self.B)

    def allsubs(self, x, y, sx, fixed, test, alpha, minsize, maxsize):
        self._gc()
        p, pmin, pmax, ds = ra.allsubs_test(self.data, self.names[x], self.names[y],
                                            self._n(list(fixed) + list(sx)), self._n(fixed), test, alpha,
                                            minsize, min(maxsize, len(sx)), B=self.B)
        return p, pmin, pmax, None if ds is None else [self.names.index(d) for d in ds]

    def allsubs_batch(self, x, y, sx, nfixed, test, alpha, minsize, maxsize):
        return [self.allsubs(x[r], y[r], sx[r][nfixed[r]:], sx[r][:nfixed[r]], test, alpha, minsize, maxsize)
                for r in range(len(x))]

    def roundrobin(self, x, z, fixed, test, alpha):
        self._gc()
        return ra.roundrobin_test(self.data, self.names[x], self._n(z), self._n(fixed), test, alpha, B=self.B)

    def _kw(self, score, arg):
        if score in ("bde", "bds"):
            return {"iss": arg}
        if score in ("aic", "bic", "aic-g", "bic-g"):
            return {"k": arg}
        return {}

    def per_node_score(self, targets, parents, score, arg):
        self._gc()
        par = {self.names[t]: self._n(p) for t, p in zip(targets, parents)}
        return ra.per_node_score(self.data, par, self._n(targets), score, **self._kw(score, arg))

    def score_cache_fill(self, score, arg, amat, blmat, updated, equivalence, reference, cache):
        self._gc()
        n = amat.shape[0]
        par = {self.names[j]: self._n([i for i in range(n) if amat[i, j]]) for j in range(n)}
        return ra.score_cache_fill(self.data, par, score, reference, equivalence, updated, amat, cache,
                                   blmat, **self._kw(score, arg))
