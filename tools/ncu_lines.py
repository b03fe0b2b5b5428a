#!/usr/bin/env python
"""Per-CUDA-source-line samples / instructions / shared-memory conflicts from an .ncu-rep captured with
--import-source on (kernels compiled with -lineinfo).  usage: ncu_lines.py rep [kernel-regex] [launch-index] [min-pct]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else ""
which = int(sys.argv[3]) if len(sys.argv) > 3 else -1
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
args = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if kre:
    args += ["--kernel-name", "regex:" + kre]
out = subprocess.run(args, capture_output=True, text=True).stdout
launches, cur, hdr, fpath, seen = [], None, None, None, set()
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1]
        if cur is None or fpath in seen:  # a file repeats -> next launch
            cur, seen = {}, set()
            launches.append(cur)
        seen.add(fpath)
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if cur is None or hdr is None or len(r) < 8 or r[2] != "-" or not r[0].isdigit():
        continue
    if not fpath.endswith("sw2d_fused.cu") and not fpath.endswith(".cu"):
        continue
    cur.setdefault((fpath.split("/")[-1], int(r[0])), [r[1], r])
L = launches[which]
i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
i_x, i_w = hdr.index("L1 Wavefronts Shared Excessive"), hdr.index("L1 Wavefronts Shared")
ts = sum(int(v[1][i_s] or 0) for v in L.values())
ti = sum(int(v[1][i_i] or 0) for v in L.values())
print("samples %d  warp-instructions %d" % (ts, ti))
for ln in sorted(L):
    src, r = L[ln]
    if len(r) <= max(i_s, i_i, i_x, i_w):
        continue
    s, i = int(r[i_s] or 0), int(r[i_i] or 0)
    if 100.0 * s / ts >= minpct or 100.0 * i / ti >= minpct:
        print("%s %5.1f%%smp %5.1f%%ins  shw %9s exc %9s | %s" % (ln[1], 100.0 * s / ts, 100.0 * i / ti, r[i_w], r[i_x], src.strip()[:105]))
