#!/usr/bin/env python
"""SASS opcode histogram of the dominant kernels of libsagpu.so (static instruction counts).
usage: tools/sass_histogram.py [libsagpu.so] > profiles/<round>_sass_histogram.txt"""
import collections, os, re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "signalalign_b200", "libsagpu.so")
KERNELS = ["_Z16sa_wide32_kernelILb1ELi2048ELi256ELi3EEv7FParams", "_Z16sa_wide32_kernelILb0ELi1024ELi256ELi3EEv7FParams",
           "_Z14sa_wide_kernelILb0ELb0ELb1ELi2048ELi512ELi2EEv7FParams", "_Z14sa_wide_kernelILb1ELb0ELb1ELi2048ELi512ELi2EEv7FParams",
           "_Z14sa_wide_kernelILb0ELb1ELb1ELi2048ELi512ELi2EEv7FParams"]
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, hist = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
GROUPS = [("FP64", r"^(DADD|DMUL|DFMA|DSETP|DMNMX)"), ("FP32", r"^(FADD|FMUL|FFMA|FMNMX|FSETP|FSEL|FMNMX3)"), ("conversion", r"^(F2F|F2I|I2F|I2FP|F2FP)"),
          ("shared ld/st", r"^(LDS|STS|LDSM)"), ("global ld/st", r"^(LDG|STG|LD\.|ST\.|LD$|ST$|ATOMG|RED)"), ("integer / logic", r"^(IADD|IMAD|IADD3|LOP3|SHF|LEA|VIADD|VIMNMX|VIADDMNMX|ISETP|SEL|IABS|POPC|FLO|BREV|PRMT|IMNMX|SGXT|LOP|MOV|UMOV|CS2R|S2R|S2UR|R2UR|P2R|R2P|PLOP3)"),
          ("uniform datapath", r"^U[A-Z]"), ("control / barrier", r"^(BRA|BSSY|BSYNC|BAR|EXIT|WARPSYNC|NOP|CALL|RET|BRX|JMP|YIELD|DEPBAR|MEMBAR|ERRBAR|BMOV|BPT)"),
          ("warp collectives", r"^(SHFL|VOTE|MATCH|REDUX)"), ("tensor core / TMA", r"^(UTC|UTMA|HMMA|IMMA|QMMA|UBLKCP|SYNCS)")]
for k in KERNELS:
    h = hist.get(k)
    if not h:
        print("%s: not in %s" % (k, so))
        continue
    tot = sum(h.values())
    print("%s  (%d instructions)" % (subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip(), tot))
    left = dict(h)
    for name, rx in GROUPS:
        n = sum(v for op, v in h.items() if re.match(rx, op))
        for op in [op for op in left if re.match(rx, op)]:
            left.pop(op)
        print("  %-20s %5d  %4.1f%%" % (name, n, 100.0 * n / tot))
    print("  %-20s %5d  %4.1f%%   %s" % ("other", sum(left.values()), 100.0 * sum(left.values()) / tot, " ".join(sorted(left))[:120]))
    print("  top opcodes: " + ", ".join("%s %d" % kv for kv in h.most_common(14)))
    print()
