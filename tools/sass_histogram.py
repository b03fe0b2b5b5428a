#!/usr/bin/env python
"""SASS opcode histogram per kernel of the product library (static instruction counts), written to profiles/.
usage: python tools/sass_histogram.py [lib.so] > profiles/r02_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "blitzdg_b200", "csrc", "libblitzdg_b200.so")
KEYS = ["DMMA", "DFMA", "DMUL", "DADD", "DSETP", "MUFU", "UBLKCP", "UBLKPF", "LDGSTS", "SYNCS", "USETMAXREG", "LDS", "STS", "LDG", "STG", "ATOMG", "BAR", "SHFL"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names, hist, cur = [], {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", cur)
        cur = re.sub(r"\(.*$", "", cur)
        names.append(cur)
        hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        hist[cur]["total"] += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                hist[cur][k] += 1
print("# SASS opcode histogram per kernel of %s (cuobjdump -sass, sm_100a; static instruction counts; tools/sass_histogram.py)" % os.path.relpath(lib, ROOT))
print("# DMMA = fp64 tensor pipe (mma.sync.m8n8k4.f64 -- tcgen05/TMEM have no fp64 form); UBLKCP = cp.async.bulk (TMA 1-D bulk copy);")
print("# UBLKPF = cp.async.bulk.prefetch.L2; LDGSTS = cp.async; SYNCS = mbarrier ops; USETMAXREG = setmaxnreg (warp-specialised register split)\n")
for n in names:
    h = hist[n]
    print("%-72s total %6d | %s" % (n[:72], h["total"], " ".join("%s=%d" % (k, h[k]) for k in KEYS if h[k])))
