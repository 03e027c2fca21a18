#!/usr/bin/env python
"""Per-instruction listing of one kernel from an .ncu-rep: executed count, stall samples, main stall reason, SASS.
usage: tools/ncu_sass.py report.ncu-rep > listing.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ia, isrc, ist = h.index("Instructions Executed"), h.index("Source"), h.index("Warp Stall Sampling (All Samples)")
stalls = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[ia]) for r in rows[2:])
tots = sum(int(r[ist]) for r in rows[2:])
print("# total executed %d, total stall samples %d" % (tot, tots))
for k, r in enumerate(rows[2:]):
    n, st = int(r[ia]), int(r[ist])
    best = max(stalls, key=lambda i: int(r[i] or 0))
    print("%5d %9.2fM %6.2f%% %6.2f%% %-18s %s" % (k, n / 1e6, 100.0 * n / tot, 100.0 * st / tots,
                                                 h[best][6:] if int(r[best] or 0) else "", r[isrc].strip()))
