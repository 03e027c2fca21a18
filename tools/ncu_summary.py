#!/usr/bin/env python
"""Condense an .ncu-rep into the few numbers DESIGN.md / profiles/ quote.

usage: tools/ncu_summary.py report.ncu-rep libsagpu.so 'kernel-mangled-prefix' > profiles/xxx.txt
Joins ncu's per-instruction execution counts (--page source) with nvdisasm -g line info of the same cubin.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, so, kern = sys.argv[1], sys.argv[2], sys.argv[3]
KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_per_inst_issued.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
print("== raw metrics (%s)" % os.path.basename(rep))
for h, u, v in zip(hdr, units, vals):
    if h in KEYS or h.startswith("smsp__average_warps_issue_stalled") and float(v or 0) > 0.05:
        print("%-75s %s %s" % (h, v, u))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
sh = srows[1]
ia, isrc = sh.index("Instructions Executed"), sh.index("Source")
data = srows[2:]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin") and "host" not in f][0]
dis = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines, cur, on = [], None, False
for l in dis.splitlines():
    if l.startswith(".text."):
        on = l.startswith(".text." + kern)
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        lines.append(cur)
if len(lines) != len(data):
    print("!! instruction count mismatch: nvdisasm %d vs ncu %d (different build?)" % (len(lines), len(data)))
agg, opagg, tot = collections.Counter(), collections.Counter(), 0
for r, srcl in zip(data, lines):
    n = int(r[ia])
    tot += n
    agg[srcl] += n
    t = r[isrc].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    opagg[op] += n
print("\n== warp-instructions executed: %d; by source line (top 40)" % tot)
for k, v in agg.most_common(40):
    print("%-28s %14d %5.1f%%" % ("%s:%s" % k if k else "?", v, 100.0 * v / tot))
print("\n== by opcode (top 30)")
for k, v in opagg.most_common(30):
    print("%-12s %14d %5.1f%%" % (k, v, 100.0 * v / tot))

if len(sys.argv) > 4:   # per-line listing of one source file in line order: tools/ncu_summary.py rep so kern file.cuh [unit]
    want = sys.argv[4]
    unit = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
    print("\n== %s in line order (count / %g)" % (want, unit))
    for k in sorted(k for k in agg if k and k[0] == want):
        print("%5d %10.2f" % (k[1], agg[k] / unit))
