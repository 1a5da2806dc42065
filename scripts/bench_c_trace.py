"""bench.py's config C section alone, under BNL_TRACE=1 (why hc inside the bench differs from hc alone)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("BNL_TRACE", "1")
import torch
import bench
from bnlearn_b200 import Engine
with Engine() as eng:
    out, classes = bench.config_c_extras(eng, 1, 0, torch.device("cuda:0"), cpu_cores=int(os.environ.get("CORES", "0")), reps=5)
    print(json.dumps(out["hc_to_convergence"]), out["first_iteration"]["ms"])
