"""Replicating one data frame over the GPUs of a job that runs one process per
GPU (torch.distributed, NCCL): the source rank registers the host data frame,
the packed columns travel to the other ranks over NVLink / NVSwitch with one
broadcast per block, and every rank ends up with its own committed data set.
SURVEY.md section 8e ("replicate the packed data set on every GPU: NCCL
broadcast from GPU 0"). torch is only the carrier of the collective here."""
import torch
import torch.distributed as dist


class _DevicePtr:
    """A raw device pointer dressed up for torch.as_tensor (no copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2}


def _as_tensor(ptr, nbytes):
    return torch.as_tensor(_DevicePtr(ptr, nbytes), device="cuda")


def broadcast_dataset(engine, columns, nlevels, nobs, src=0):
    """Every rank calls this; only `src` needs `columns` (host arrays, as for
    Engine.register). Returns this rank's DataSet."""
    rank = dist.get_rank()
    # checked on EVERY rank before the first collective: a rank that raised on its own
    # (alloc refuses what register accepts) would leave the others waiting in the broadcast.
    if nlevels is not None and any(int(l) > 127 for l in nlevels):
        raise ValueError("broadcast_dataset carries the uint8 factor block only: at most 127 levels per factor")
    if rank == src:
        ds = engine.register(columns, nlevels)
    else:
        ds = engine.alloc(nlevels if nlevels is not None else [0] * len(columns), nobs)
    u8, nu, f64, nf = ds.device_view(0)
    if nu:
        dist.broadcast(_as_tensor(u8, nu), src=src)
    if nf:
        dist.broadcast(_as_tensor(f64, nf), src=src)
    # the collective runs on NCCL's stream, ordered after torch's current stream:
    # wait for it before the library (its own streams) reads or recycles the blocks.
    torch.cuda.current_stream().synchronize()
    if rank != src:
        ds.commit()
    return ds


def pool_moments(nobs, mean, cov, device=None):
    """Pooled (nobs, mean, cov) of row blocks held by the ranks of the default
    process group, from each rank's own (nobs, mean, cov with divisor nobs - 1):
        n = sum n_r,  mean = sum n_r mean_r / n,
        (n - 1) cov = sum [(n_r - 1) cov_r + n_r (mean_r - mean)(mean_r - mean)'].
    Two sum all-reduces, over 1 + p and over p * p doubles (NCCL when `device` is a
    CUDA device, gloo on the CPU). Every rank gets the result and passes it to
    DataSet.set_moments ("for the Gram, one all-reduce")."""
    import numpy as np
    mean = np.asarray(mean, dtype=np.float64)
    cov = np.asarray(cov, dtype=np.float64)
    dev = device if device is not None else "cpu"
    first = torch.empty(1 + mean.size, dtype=torch.float64)
    first[0] = float(nobs)
    first[1:] = torch.from_numpy(mean * float(nobs))
    first = first.to(dev)
    dist.all_reduce(first, op=dist.ReduceOp.SUM)
    first = first.cpu().numpy()
    n = first[0]
    pooled_mean = first[1:] / n
    d = mean - pooled_mean
    second = torch.from_numpy((float(nobs) - 1.0) * cov + float(nobs) * np.outer(d, d)).to(dev)
    dist.all_reduce(second, op=dist.ReduceOp.SUM)
    return int(round(n)), pooled_mean, second.cpu().numpy() / (n - 1.0)


# ---------------------------------------------------------------------------
# Units sharded over the ranks of a job: every rank holds a
# replica of the data set, takes a contiguous block of each batch and the doubles
# are gathered in index order, so the result does not depend on the world size.

def block(n, world, rank):
    """Contiguous block [b, e) of n units for `rank` (the split bnl_block uses
    inside a multi-device context)."""
    return n * rank // world, n * (rank + 1) // world


def gather_rows(mine, n, device=None):
    """Every rank passes its block of rows (float64 array [e - b, width], blocks as
    block() cuts them); every rank gets the [n, width] array in index order. One
    all-gather of equal-size padded blocks (NCCL when `device` is a CUDA device)."""
    import numpy as np
    world = dist.get_world_size() if dist.is_initialized() else 1
    mine = np.ascontiguousarray(mine, dtype=np.float64)
    if world == 1:
        return mine
    width = mine.shape[1] if mine.ndim == 2 else 1
    per = (n + world - 1) // world + 1
    buf = torch.zeros((per, width), dtype=torch.float64)
    buf[:mine.shape[0]] = torch.from_numpy(mine.reshape(-1, width))
    if device is not None:
        buf = buf.to(device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty((n, width))
    for r, p in enumerate(parts):
        b, e = block(n, world, r)
        out[b:e] = p[:e - b].cpu().numpy()
    return out


def cost_blocks(costs, world, together=0):
    """Contiguous blocks of a unit list with about equal total cost: boundaries
    bounds[r] .. bounds[r + 1] for rank r (pc.stable's marginal tests come from one
    tensor-core contraction and cost a fraction of a conditional test, so equal
    COUNTS would leave the rank holding the conditional tests with most of the work).
    together: the first `together` units form ONE piece of work that does not get
    cheaper when cut (the contraction reads every row of every column whether it
    serves 100 pairs or 666); if their cost is at least a rank's share they all go to
    rank 0 and the rest is spread over the other ranks."""
    import numpy as np
    costs = np.asarray(costs, dtype=np.float64)
    if together and world > 1 and costs[:together].sum() >= 0.75 * costs.sum() / world:
        rest = cost_blocks(costs[together:], world - 1)
        return [0] + [together + b for b in rest]
    c = np.cumsum(costs)
    total = float(c[-1]) if len(c) else 0.0
    bounds = [0]
    for r in range(1, world):
        bounds.append(int(np.searchsorted(c, total * r / world, side="left")) if total > 0 else 0)
    bounds.append(len(c))
    for r in range(1, len(bounds)):
        bounds[r] = max(bounds[r], bounds[r - 1])
    return bounds


class Gatherer:
    """gather_rows with everything allocated once: pinned host buffers, device
    buffers and ONE all_gather_into_tensor per call (a bench step gathers three
    doubles per test; fresh tensors and a copy per rank cost more than the tests
    of a rank's block). bounds[r] .. bounds[r + 1] is rank r's block of rows."""

    def __init__(self, bounds, width, device):
        self.bounds, self.width, self.device = list(bounds), int(width), device
        self.world = len(bounds) - 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.per = max(b - a for a, b in zip(bounds[:-1], bounds[1:])) if self.world else 0
        if self.world > 1:
            self.h_in = torch.zeros((self.per, width), dtype=torch.float64).pin_memory()
            self.d_in = torch.zeros((self.per, width), dtype=torch.float64, device=device)
            self.d_out = torch.zeros((self.world * self.per, width), dtype=torch.float64, device=device)
            self.h_out = torch.zeros((self.world * self.per, width), dtype=torch.float64).pin_memory()

    def __call__(self, mine):
        import numpy as np
        if self.world <= 1:
            return np.asarray(mine, dtype=np.float64)
        m = np.asarray(mine, dtype=np.float64).reshape(-1, self.width)
        self.h_in.numpy()[:m.shape[0]] = m
        self.d_in.copy_(self.h_in, non_blocking=True)
        dist.all_gather_into_tensor(self.d_out, self.d_in)
        self.h_out.copy_(self.d_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        full = self.h_out.numpy()
        return np.concatenate([full[r * self.per: r * self.per + self.bounds[r + 1] - self.bounds[r]]
                               for r in range(self.world)])


class ShardedBackend:
    """A learners backend whose batched calls are split over the ranks: the pairs
    of one pc.stable level go to the workers in contiguous blocks, as
    smartSapply / parSapplyLB hands them to the cluster's slaves (R/pcalgo.R:24-28),
    and every rank receives all results, so the level loop runs identically
    everywhere. `backend` is this rank's own (learners.GpuBackend on its replica)."""

    def __init__(self, backend, device=None):
        self.backend = backend
        self.device = device
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.gathers = 0
        self.tests_by_size = {}                 # conditioning-set size -> tests this rank ran

    def allsubs_batch(self, x, y, sx, nfixed, test, alpha, minsize, maxsize):
        import numpy as np
        n = len(x)
        b, e = block(n, self.world, self.rank)
        before = self.backend.counter.value
        mine = self.backend.allsubs_batch(x[b:e], y[b:e], sx[b:e], nfixed[b:e], test, alpha, minsize, maxsize)
        if minsize == maxsize:
            self.tests_by_size[minsize] = self.tests_by_size.get(minsize, 0) + self.backend.counter.value - before
        if self.world == 1:                 # same types as the gathered form (python floats, int lists)
            return [(float(p), float(a), float(b), None if d is None else [int(v) for v in d]) for p, a, b, d in mine]
        width = 4 + max([len(s) for s in sx] + [0])
        rows = np.zeros((e - b, width))
        for k, (p, pmin, pmax, dsep) in enumerate(mine):
            rows[k, :3] = (p, pmin, pmax)
            rows[k, 3] = -1 if dsep is None else len(dsep)
            if dsep:
                rows[k, 4:4 + len(dsep)] = dsep
        full = gather_rows(rows, n, self.device)
        self.gathers += 1
        out = []
        for r in full:
            nd = int(r[3])
            out.append((float(r[0]), float(r[1]), float(r[2]), None if nd < 0 else [int(v) for v in r[4:4 + nd]]))
        return out

    def total_tests(self):
        """Sum of the ranks' test counters (R/learning-algorithms.R:163-165)."""
        t = torch.tensor([float(self.backend.counter.value)], dtype=torch.float64)
        if self.world > 1:
            if self.device is not None:
                t = t.to(self.device)
            dist.all_reduce(t)
        return float(t.cpu()[0])

    def __getattr__(self, name):            # indep_test, roundrobin, scores: this rank's own
        return getattr(self.backend, name)
