#!/usr/bin/env python
"""Per-kernel census of the Blackwell-specific SASS in the built library (the .so itself is git-ignored):
tcgen05.mma (UTCHMMA), TMEM loads (LDTM), bulk async copies (UBLKCP / UBLKPF), tcgen05.commit (UTCBAR), cp.async (LDGSTS),
register re-allocation (USETMAXREG), and registers / spills from the ptxas logs.
    python scripts/sass_census.py > profiles/<tag>_sass_census.txt"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch-motion-planner_b200", "nfopp_b200", "libnfopp.so")
PAT = ["UTCHMMA", "LDTM", "UBLKCP", "UBLKPF", "UTCBAR", "LDGSTS", "USETMAXREG", "SYNCS", "HMMA", "R2UR"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
            counts.setdefault(cur, collections.Counter())
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["instructions"] += 1
            for p in PAT:
                if op.startswith(p):
                    counts[cur][p] += 1
    regs = {}
    for log in glob.glob(os.path.join(ROOT, "pytorch-motion-planner_b200", "csrc", "build", "*.ptxas.log")):
        txt = open(log).read()
        for m in re.finditer(r"Compiling entry function '(\S+)'.*?\n.*?\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers",
                             txt):
            name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
            regs[name] = (int(m.group(5)), int(m.group(3)), int(m.group(4)))
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# SASS census of libnfopp.so (sm_100a), built from commit {commit} (+ working tree); cuobjdump -sass")
    print(f"# {'kernel':58s} {'instr':>7s} {'regs':>5s} {'spill st/ld':>11s}  " + " ".join(f"{p:>10s}" for p in PAT))
    for k, c in counts.items():
        r = regs.get(k, ("?", "?", "?"))
        print(f"{k[:60]:60s} {c['instructions']:7d} {str(r[0]):>5s} {str(r[1]) + '/' + str(r[2]):>11s}  " + " ".join(f"{c[p]:10d}" for p in PAT))


if __name__ == "__main__":
    main()
