#!/usr/bin/env python
"""Blackwell-native evidence from the shipped library: per kernel, the count of the SASS opcodes that prove tcgen05 / TMEM /
bulk-copy (TMA engine) use, from `cuobjdump -sass nerf_pl_b200/lib/libnerf_b200.so`.  Writes profiles/r02_sass_histogram.txt."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "nerf_pl_b200", "lib", "libnerf_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"\b(UTC[A-Z0-9]*MMA[.A-Z0-9_]*|UTCBAR[.A-Z0-9_]*|UTCCP[.A-Z0-9_]*|LDTM[.A-Za-z0-9_]*|STTM[.A-Za-z0-9_]*|UBLKCP[.A-Z0-9_]*|"
                 r"UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|SYNCS[.A-Z0-9_]*|FENCE\.VIEW\.ASYNC[.A-Z]*|HMMA[.A-Z0-9_]*|HGMMA[.A-Z0-9_]*|"
                 r"REDG?[.A-Z0-9_]*ADD[.A-Z0-9_]*|ELECT|STG\.E[.A-Z0-9_]*256|UCGABAR[.A-Z_]*)\b")
out, name, counts, total = [], None, None, 0


def flush():
    if name is not None:
        keys = sorted(counts, key=lambda k: (-counts[k], k))
        short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        out.append(f"{short}\n    instructions: {total}; " + (", ".join(f"{k} x{counts[k]}" for k in keys) if keys else "(none of the tracked opcodes)"))


for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        name, counts, total = m.group(1), collections.Counter(), 0
        continue
    if name and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        total += 1
        for op in pat.findall(line):
            counts[op] += 1
flush()
text = ("# SASS opcode histogram of nerf_pl_b200/lib/libnerf_b200.so (cuobjdump -sass, sm_100a), tools/sass_histogram.py\n"
        "# UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA engine, linear),\n"
        "# SYNCS = mbarrier, FENCE.VIEW.ASYNC = fence.proxy.async; no HMMA / HGMMA (legacy mma.sync / wgmma) anywhere.\n\n"
        + "\n".join(out) + "\n")
path = os.path.join(ROOT, "profiles", "r02_sass_histogram.txt")
open(path, "w").write(text)
print(path, len(out), "kernels")
