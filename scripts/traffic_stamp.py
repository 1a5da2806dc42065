#!/usr/bin/env python
"""Turn one `ncu --set full` raw CSV of a bench.py step into profiles/traffic.json.

    ncu --set full --clock-control none -k regex:"count_small|pair_onehot|epilogue" -s 6 -c 6 \
        -o gpurun_out/step python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline
    ncu -i gpurun_out/step.ncu-rep --page raw --csv > gpurun_out/step_raw.csv
    python scripts/traffic_stamp.py gpurun_out/step_raw.csv

The file is stamped with a hash of the kernel SOURCES (bnlearn_b200/csrc): bench.py
fills `roofline.traffic` from it only while that hash is the one of the tree it runs
from, so a number measured on another build is never carried into a bench line.
"""
import csv
import glob
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def source_stamp():
    h = hashlib.sha256()
    pats = ("*.cu", "*.cuh", "*.h", "*.cpp", "Makefile")
    files = sorted(f for p in pats for f in glob.glob(os.path.join(ROOT, "bnlearn_b200", "csrc", p)))
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units = rows[hdr], rows[hdr + 1]
    col = {n: names.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
                                        "dram__bytes_write.sum")}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = {}
    for r in rows[hdr + 2:]:
        k = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
        b = sum(float(r[col[m]]) * scale[units[col[m]]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        e = per.setdefault(k, {"launches": 0, "dram_bytes": 0.0, "ms": 0.0})
        e["launches"] += 1
        e["dram_bytes"] += b
        e["ms"] += float(r[col["gpu__time_duration.sum"]]) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(
            units[col["gpu__time_duration.sum"]], 1.0)
    kernels = {k: {"launches_captured": v["launches"], "dram_bytes_per_launch": int(v["dram_bytes"] / v["launches"]),
                   "ms_per_launch_under_ncu": round(v["ms"] / v["launches"], 5)} for k, v in per.items()}
    out = {"source_stamp": source_stamp(), "from": os.path.basename(path),
           "metric": "dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full --clock-control none, per launch",
           "command": "python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline",
           "kernels": kernels}
    json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
