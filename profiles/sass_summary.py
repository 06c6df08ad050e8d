"""Opcode histogram per kernel of libbe_b200.so: `cuobjdump -sass` → profiles/sass_summary.txt.

usage: python profiles/sass_summary.py > profiles/sass_summary.txt
Lists, per kernel, the instruction count and the counts of the opcodes that tell how a kernel moves and computes:
UBLKCP (cp.async.bulk, the TMA bulk-copy engine), UTMALDG / UTMASTG (tensor-map TMA: the gather fetches whole stages of rows as 2-D boxes),
SYNCS (mbarrier), LDGSTS (cp.async), LDG / STG / LDS / STS / ATOMS / ATOMG / RED, F2F (fp32 <-> fp64 conversions), DADD /
DMUL / DFMA (fp64 arithmetic), MUFU, REDUX / VOTE / SHFL / MATCH (warp collectives), ACQBULK / griddepcontrol shows as
ACQBULK-free `BAR`-less waits; the tensor-core opcodes (UTC*MMA, HMMA, ...) are listed to show there are none."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "bioengine-worker_b200", "libbe_b200.so")
WATCH = ["UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG", "RED", "F2F",
         "DADD", "DMUL", "DFMA", "MUFU", "REDUX", "VOTE", "SHFL", "MATCH", "BAR", "HMMA", "UTCHMMA", "LDTM"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            name = re.sub(r"\(.*", "", name).replace("void ", "")
            cur = kernels.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_total"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + ".") or (w in ("UTCHMMA",) and op.startswith("UTC")):
                    cur[w] += 1
    print(f"# {os.path.relpath(SO, ROOT)}: SASS opcode counts per kernel (static instruction counts, sm_100a)")
    print("# " + " ".join(f"{w:>7s}" for w in ["total"] + WATCH))
    for name, c in kernels.items():
        print(name)
        print("  " + " ".join(f"{c.get(w, 0):7d}" for w in ["_total"] + WATCH))
    tot = collections.Counter()
    for c in kernels.values():
        tot.update(c)
    print("ALL KERNELS")
    print("  " + " ".join(f"{tot.get(w, 0):7d}" for w in ["_total"] + WATCH))


if __name__ == "__main__":
    main()
