"""Writes profiles/<tag>_sass_hot_loops.txt: excerpts of `cuobjdump -sass` for the
kernels DESIGN.md discusses, cut around the instructions that prove the design --
PRMT + ATOMS.POPC.INC (packed-index counting), DMMA + UBLKCP (fp64 tensor cores of the Gram,
stages filled by the bulk-copy engine), UTCIMMA / LDTM / UTCBAR (tcgen05 pair kernel) -- with the instruction histogram of
each kernel. Runs here (no GPU): nvcc cross-compiles and cuobjdump reads the .so.

    python scripts/sass_excerpts.py r2
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "bnlearn_b200", "libbnl_b200.so")
WANT = [  # (substring of the mangled name, mnemonics to centre the excerpt on, lines of context)
    ("count_small_kernelILb0EE", ("ATOMS.POPC.INC",), 14),
    ("count_pair32_kernel", ("ATOMS.POPC.INC",), 12),
    ("count_group_kernelILi512ELi16384EE", ("ATOMS.POPC.INC",), 10),
    ("count_based_kernelILi5ELi1ELi512ELb1EE", ("ATOMS.POPC.INC",), 12),
    ("gram_dmma_kernelILb1EE", ("DMMA", "UBLKCP", "SYNCS"), 10),
    ("pair_onehot_kernelILi128ELi1ELi2ELb0ELb0ELi11ELi4ELi8EE", ("UTCIMMA", "LDTM", "UTCBAR", "PRMT", "STS.128"), 8),
    ("pair_onehot_kernelILi256ELi3ELi2ELb0ELb0ELi19ELi2E", ("UTCIMMA", "LDTM", "UTCBAR", "STS.128"), 8),
]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    funcs = re.split(r"\n\s*Function : ", sass)
    out = ["SASS excerpts of bnlearn_b200/libbnl_b200.so (sm_100a), git %s, `cuobjdump -sass`" % head, ""]
    for key, marks, ctx in WANT:
        body = next((f for f in funcs if f.startswith("_Z") and key in f.split("\n", 1)[0]), None)
        if body is None:
            out.append("!! %s not found" % key)
            continue
        name, _, text = body.partition("\n")
        lines = [l.rstrip() for l in text.split("\n") if re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", l)]
        ops = collections.Counter()
        for l in lines:
            m = re.search(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", l)
            if m:
                ops[m.group(1).split(".")[0]] += 1
        out.append("=" * 100)
        out.append(name)
        out.append("instructions: %d; histogram: %s" % (len(lines), ", ".join("%s %d" % kv for kv in ops.most_common(14))))
        for mk in marks:
            hits = [i for i, l in enumerate(lines) if mk in l]
            out.append("-- %s: %d occurrences%s" % (mk, len(hits), "" if hits else " (none)"))
            if hits:
                i = hits[len(hits) // 2]
                for l in lines[max(0, i - ctx): i + ctx + 1]:
                    out.append("    " + re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).strip())
        out.append("")
    path = os.path.join(ROOT, "profiles", "%s_sass_hot_loops.txt" % tag)
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    print(path, len(out), "lines")


if __name__ == "__main__":
    main()
