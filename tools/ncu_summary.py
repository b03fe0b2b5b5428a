#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics + stall reasons + hottest SASS lines.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [kernel-regex] [launch-index]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else ""
which = int(sys.argv[3]) if len(sys.argv) > 3 else -1
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread ", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ", "lts__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum ", "launch__grid_size", "launch__shared_mem_per_block_dynamic", "sm__inst_executed.sum ",
        "smsp__inst_executed.avg.per_cycle_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for i, h in enumerate(hdr):
    if any((h + " ").startswith(k) for k in KEYS) or h == "Kernel Name":
        print("%-95s %-12s %s" % (h, units[i], [r[i][:40] for r in rows[2:]]))
args = ["ncu", "-i", rep, "--page", "source", "--csv"]
if kre:
    args += ["--kernel-name", "regex:" + kre]
src = subprocess.run(args, capture_output=True, text=True).stdout
blocks, cur, h2 = [], [], None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Kernel Name":
        if cur:
            blocks.append(cur)
        cur = []
        continue
    if r and r[0] == "Address":
        h2 = r
        continue
    if r:
        cur.append(r)
if cur:
    blocks.append(cur)
b = blocks[which]
idx = {h: i for i, h in enumerate(h2)}
stalls = [h for h in h2 if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: 0 for s in stalls}
for r in b:
    for s in stalls:
        try:
            tot[s] += int(r[idx[s]])
        except ValueError:
            pass
S = sum(tot.values())
print("\nstall samples: total", S, " instructions", len(b))
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:9]:
    print("  %-24s %7d %5.1f%%" % (s, v, 100.0 * v / S))
print("\nhottest SASS:")
for r in sorted(b, key=lambda r: -int(r[idx["# Samples"]] or 0))[:28]:
    st = sorted([(s[6:], int(r[idx[s]])) for s in stalls if r[idx[s]] not in ("", "0")], key=lambda x: -x[1])
    print("  %6s  %-58s %s" % (r[idx["# Samples"]], r[idx["Source"]].strip()[:58], st[:3]))
ex = sum(int(r[idx["Instructions Executed"]] or 0) for r in b)
print("\nwarp instructions executed:", ex)
from collections import Counter
cnt = Counter()
for r in b:
    op = r[idx["Source"]].strip().split()
    if not op:
        continue
    o = op[1] if op[0].startswith("@") and len(op) > 1 else op[0]
    cnt[o.split(".")[0]] += int(r[idx["Instructions Executed"]] or 0)
print("by opcode:", [(k, v, "%.1f%%" % (100.0 * v / ex)) for k, v in cnt.most_common(16)])
