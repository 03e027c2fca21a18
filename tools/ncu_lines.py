#!/usr/bin/env python
"""Per-instruction execution counts of one kernel from an .ncu-rep, in SASS order, each with its source line (nvdisasm -g of
the same library), plus a by-line summary.  usage: tools/ncu_lines.py report.ncu-rep libsagpu.so mangled-kernel-prefix out.txt [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, so, kern, outp = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 50
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
sh = srows[1]
ia, isrc = sh.index("Instructions Executed"), sh.index("Source")
data = srows[2:]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = sorted(f for f in os.listdir(tmp) if f.endswith(".cubin"))[-1] if False else [f for f in os.listdir(tmp) if f.endswith(".cubin") and "sa_gpu.sm" in f][0]
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
    if re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l):
        lines.append(cur)
assert len(lines) == len(data), (len(lines), len(data))
tot = sum(int(r[ia]) for r in data)
by = collections.Counter()
with open(outp, "w") as f:
    for r, s in zip(data, lines):
        n = int(r[ia])
        by[s] += n
        f.write("%-24s %14d  %s\n" % ("%s:%d" % s if s else "?", n, r[isrc]))
here = os.path.dirname(os.path.abspath(so))
cache = {}
print("total warp-instructions %d" % tot)
# stall samples by source line (all samples), with the four big reasons split out
cols = {k: sh.index(k) for k in ("# Samples", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb", "stall_math", "stall_branch_resolving", "stall_no_inst") if k in sh}
if os.environ.get("STALLS") and cols:
    st = collections.defaultdict(lambda: collections.Counter())
    for r, s_ in zip(data, lines):
        for k, i in cols.items():
            st[s_][k] += int(r[i] or 0)
    allS = sum(v["# Samples"] for v in st.values())
    print("samples %d: " % allS + ", ".join("%s %.1f%%" % (k, 100.0 * sum(v[k] for v in st.values()) / allS) for k in cols if k != "# Samples"))
    for k, v in sorted(st.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
        text = ""
        if k:
            p = os.path.join(os.path.dirname(os.path.abspath(so)), "csrc", k[0])
            if os.path.exists(p):
                L = open(p).read().split("\n")
                text = L[k[1] - 1].strip()[:80] if k[1] - 1 < len(L) else ""
        print("%-22s %5.1f%%  lsb %4.1f bar %4.1f wait %4.1f ssb %4.1f | %s" % ("%s:%d" % k if k else "?", 100.0 * v["# Samples"] / allS,
              100.0 * v["stall_long_sb"] / allS, 100.0 * v["stall_barrier"] / allS, 100.0 * v["stall_wait"] / allS, 100.0 * v["stall_short_sb"] / allS, text))
    print()
for k, v in by.most_common(top):
    text = ""
    if k:
        p = os.path.join(here, "csrc", k[0])
        if os.path.exists(p):
            cache.setdefault(p, open(p).read().split("\n"))
            text = cache[p][k[1] - 1].strip()[:100] if k[1] - 1 < len(cache[p]) else ""
    print("%-22s %5.1f%%  %s" % ("%s:%d" % k if k else "?", 100.0 * v / tot, text))
