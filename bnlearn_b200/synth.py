"""Seeded synthetic data of the BASELINE.json shapes.

All generators use numpy.random.Generator(PCG64(seed)) and return data in the
reference's conventions: int32 1-based factor codes or float64 columns, one
contiguous array per column.

The alarm / insurance network STRUCTURES are the ones documented in the
reference's manual pages (man/alarm.Rd:115-121, man/insurance.Rd:102-114) and
the level counts are those of its bundled data frames; the conditional
probability tables are drawn here (Dirichlet(1) rows), so the data are
"alarm-shaped" / "insurance-shaped", not the original samples.
"""
import re

import numpy as np

ALARM_MODEL = (
    "[HIST|LVF][CVP|LVV][PCWP|LVV][HYP][LVV|HYP:LVF][LVF]"
    "[STKV|HYP:LVF][ERLO][HRBP|ERLO:HR][HREK|ERCA:HR][ERCA][HRSA|ERCA:HR][ANES]"
    "[APL][TPR|APL][ECO2|ACO2:VLNG][KINK][MINV|INT:VLNG][FIO2][PVS|FIO2:VALV]"
    "[SAO2|PVS:SHNT][PAP|PMB][PMB][SHNT|INT:PMB][INT][PRSS|INT:KINK:VTUB][DISC]"
    "[MVS][VMCH|MVS][VTUB|DISC:VMCH][VLNG|INT:KINK:VTUB][VALV|INT:VLNG]"
    "[ACO2|VALV][CCHL|ACO2:ANES:SAO2:TPR][HR|CCHL][CO|HR:STKV][BP|CO:TPR]")
# column order and level counts of data/alarm.rda
ALARM_NAMES = ["CVP", "PCWP", "HIST", "TPR", "BP", "CO", "HRBP", "HREK", "HRSA", "PAP", "SAO2",
               "FIO2", "PRSS", "ECO2", "MINV", "MVS", "HYP", "LVF", "APL", "ANES", "PMB", "INT",
               "KINK", "DISC", "LVV", "STKV", "CCHL", "ERLO", "HR", "ERCA", "SHNT", "PVS", "ACO2",
               "VALV", "VLNG", "VTUB", "VMCH"]
ALARM_LEVELS = [3, 3, 2, 3, 3, 3, 3, 3, 3, 3, 3, 2, 4, 4, 4, 3, 2, 2, 2, 2, 2, 3, 2, 2, 3, 3, 2,
                2, 3, 2, 2, 3, 3, 4, 4, 4, 4]

INSURANCE_MODEL = (
    "[Age][Mileage][SocioEcon|Age][GoodStudent|Age:SocioEcon]"
    "[RiskAversion|Age:SocioEcon][OtherCar|SocioEcon][VehicleYear|SocioEcon:RiskAversion]"
    "[MakeModel|SocioEcon:RiskAversion][SeniorTrain|Age:RiskAversion]"
    "[HomeBase|SocioEcon:RiskAversion][AntiTheft|SocioEcon:RiskAversion]"
    "[RuggedAuto|VehicleYear:MakeModel][Antilock|VehicleYear:MakeModel]"
    "[DrivingSkill|Age:SeniorTrain][CarValue|VehicleYear:MakeModel:Mileage]"
    "[Airbag|VehicleYear:MakeModel][DrivQuality|RiskAversion:DrivingSkill]"
    "[Theft|CarValue:HomeBase:AntiTheft][Cushioning|RuggedAuto:Airbag]"
    "[DrivHist|RiskAversion:DrivingSkill][Accident|DrivQuality:Mileage:Antilock]"
    "[ThisCarDam|RuggedAuto:Accident][OtherCarCost|RuggedAuto:Accident]"
    "[MedCost|Age:Accident:Cushioning][ILiCost|Accident]"
    "[ThisCarCost|ThisCarDam:Theft:CarValue][PropCost|ThisCarCost:OtherCarCost]")
INSURANCE_NAMES = ["GoodStudent", "Age", "SocioEcon", "RiskAversion", "VehicleYear", "ThisCarDam",
                   "RuggedAuto", "Accident", "MakeModel", "DrivQuality", "Mileage", "Antilock",
                   "DrivingSkill", "SeniorTrain", "ThisCarCost", "Theft", "CarValue", "HomeBase",
                   "AntiTheft", "PropCost", "OtherCarCost", "OtherCar", "MedCost", "Cushioning",
                   "Airbag", "ILiCost", "DrivHist"]
INSURANCE_LEVELS = [2, 3, 4, 4, 2, 4, 3, 4, 5, 3, 4, 2, 3, 2, 4, 2, 5, 4, 2, 4, 3, 2, 4, 4, 2, 4, 3]


def parse_modelstring(model):
    """'[A][B|A:C]...' -> {node: [parents]} (bnlearn's model string format)."""
    out = {}
    for tok in re.findall(r"\[([^\]]+)\]", model):
        node, _, pa = tok.partition("|")
        out[node] = pa.split(":") if pa else []
    return out


def topological_order(parents):
    order, seen = [], set()

    def visit(n):
        if n in seen:
            return
        for p in parents[n]:
            visit(p)
        seen.add(n)
        order.append(n)

    for n in parents:
        visit(n)
    return order


def sample_discrete_bn(names, levels, parents, n, rng):
    """Forward-sample a discrete BN whose CPT rows are Dirichlet(1,...,1).
    Returns int32 1-based columns in `names` order."""
    lev = dict(zip(names, levels))
    data = {}
    for node in topological_order(parents):
        pa = parents[node]
        cfg = np.zeros(n, dtype=np.int64)
        mul = 1
        for p in pa:
            cfg += (data[p] - 1).astype(np.int64) * mul
            mul *= lev[p]
        cpt = rng.dirichlet(np.ones(lev[node]), size=mul)          # (configs, levels)
        cdf = np.cumsum(cpt, axis=1)
        u = rng.random(n)
        col = (u[:, None] > cdf[cfg]).sum(axis=1)
        np.minimum(col, lev[node] - 1, out=col)
        data[node] = (col + 1).astype(np.int32)
    return [np.ascontiguousarray(data[nm]) for nm in names]


def alarm_shaped(n, seed=20261106):
    """Config B: 37 factors with the alarm DAG and level counts."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return list(ALARM_NAMES), sample_discrete_bn(ALARM_NAMES, ALARM_LEVELS,
                                                 parse_modelstring(ALARM_MODEL), n, rng), list(ALARM_LEVELS)


def insurance_shaped(n, seed=20261109):
    """Config E: 27 factors with the insurance DAG and level counts."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return list(INSURANCE_NAMES), sample_discrete_bn(INSURANCE_NAMES, INSURANCE_LEVELS,
                                                     parse_modelstring(INSURANCE_MODEL), n, rng), \
        list(INSURANCE_LEVELS)


def random_discrete_bn(nnodes, n, seed=20261107, max_in=4, mean_in=1.5):
    """Config C: random DAG (topological order = index order), in-degree
    min(Poisson(mean_in), max_in), levels uniform on {2, 3, 4}."""
    rng = np.random.Generator(np.random.PCG64(seed))
    names = ["X%d" % i for i in range(nnodes)]
    levels = [int(v) for v in rng.integers(2, 5, nnodes)]
    parents = {}
    for i, nm in enumerate(names):
        k = min(int(rng.poisson(mean_in)), max_in, i)
        pa = rng.choice(i, size=k, replace=False) if k > 0 else []
        parents[nm] = [names[int(j)] for j in sorted(pa)]
    return names, sample_discrete_bn(names, levels, parents, n, rng), levels, parents


def random_gaussian_bn(nnodes, n, seed=20261108, max_in=5, mean_in=1.5):
    """Config D: random linear-Gaussian BN; coefficients U(+-[0.3, 1]), intercept 0,
    noise sd U(0.5, 1.5)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    names = ["G%d" % i for i in range(nnodes)]
    cols, parents = [], {}
    for i, nm in enumerate(names):
        k = min(int(rng.poisson(mean_in)), max_in, i)
        pa = sorted(int(j) for j in rng.choice(i, size=k, replace=False)) if k > 0 else []
        x = rng.normal(size=n) * rng.uniform(0.5, 1.5)
        for j in pa:
            x += rng.uniform(0.3, 1.0) * rng.choice([-1.0, 1.0]) * cols[j]
        cols.append(np.ascontiguousarray(x))
        parents[nm] = [names[j] for j in pa]
    return names, cols, parents


# ---------------------------------------------------------------------------
# The two large shapes generated ON THE DEVICE with torch (seeded): numpy needs
# minutes for 10^9 categorical / 2 x 10^9 normal draws, the GPU seconds. The DAG,
# levels and coefficients come from numpy PCG64(seed) as above; only the uniform
# / normal draws are torch's. Harness code for tests and bench.py (needs CUDA);
# both return host arrays in the reference's conventions.

def random_discrete_bn_device(nnodes=1000, rows=1_000_000, seed=20261107, max_in=4, mean_in=1.5):
    """Config C shape: -> (int32 array [nnodes, rows] of 1-based codes, levels, parents)."""
    import torch
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    cpu = np.random.Generator(np.random.PCG64(seed))
    levels = [int(v) for v in cpu.integers(2, 5, nnodes)]
    cols = torch.empty((nnodes, rows), dtype=torch.int32, device="cuda")
    parents = []
    for i in range(nnodes):
        k = min(int(cpu.poisson(mean_in)), max_in, i)
        pa = sorted(int(j) for j in cpu.choice(i, size=k, replace=False)) if k else []
        parents.append(pa)
        cfg = torch.zeros(rows, dtype=torch.int64, device="cuda")
        mul = 1
        for j in pa:
            cfg += (cols[j] - 1).to(torch.int64) * mul
            mul *= levels[j]
        cpt = torch.from_numpy(cpu.dirichlet(np.ones(levels[i]), size=mul)).to("cuda")
        cdf = torch.cumsum(cpt, dim=1)
        u = torch.rand(rows, generator=g, device="cuda", dtype=torch.float64)
        lvl = (u[:, None] > cdf[cfg]).sum(dim=1).clamp_(max=levels[i] - 1)
        cols[i] = (lvl + 1).to(torch.int32)
    host = cols.cpu().numpy()
    del cols
    torch.cuda.empty_cache()
    return host, levels, parents


def random_gaussian_bn_device(nvars=2000, rows=1_000_000, seed=20261108, max_in=5, mean_in=1.5):
    """Config D shape: -> (float64 array [nvars, rows], parents)."""
    import torch
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    cpu = np.random.Generator(np.random.PCG64(seed))
    x = torch.empty((nvars, rows), dtype=torch.float64, device="cuda")
    parents = []
    for i in range(nvars):
        k = min(int(cpu.poisson(mean_in)), max_in, i)
        pa = [int(j) for j in cpu.choice(i, size=k, replace=False)] if k else []
        parents.append(pa)
        x[i] = torch.randn(rows, generator=g, device="cuda", dtype=torch.float64) * float(cpu.uniform(0.5, 1.5))
        for j in pa:
            x[i] += float(cpu.uniform(0.3, 1.0) * cpu.choice([-1.0, 1.0])) * x[j]
    host = x.cpu().numpy()
    del x
    torch.cuda.empty_cache()
    return host, parents


def alarm_shaped_cached(rows, seed=20261106):
    """alarm_shaped() as an int8 array [37, rows], cached on local disk between
    invocations (the numpy sampler needs ~30 s for 10 M rows)."""
    import os
    cache = os.path.join("/tmp", "bnl_bench_cache")
    os.makedirs(cache, exist_ok=True)
    path = os.path.join(cache, "alarm_%d_%d.npy" % (rows, seed))
    if os.path.exists(path):
        codes = np.load(path)
    else:
        _, cols, _ = alarm_shaped(rows, seed)
        codes = np.stack(cols).astype(np.int8)
        tmp = path + ".%d.tmp.npy" % os.getpid()
        np.save(tmp, codes)
        os.replace(tmp, path)
    return list(ALARM_NAMES), codes, list(ALARM_LEVELS)
