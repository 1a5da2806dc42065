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
