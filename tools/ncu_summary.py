#!/usr/bin/env python
"""text summary of one kernel of an .ncu-rep (ncu --set full): headline counters, stall reasons, pipe mix, memory, and the
per-source-line hot spots (tools/ncu_lines.py).  usage: tools/ncu_summary.py report.ncu-rep mangled-kernel-prefix > profiles/x.txt"""
import csv, io, os, subprocess, sys
rep, kern = sys.argv[1:3]
here = os.path.dirname(os.path.abspath(__file__))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, u, v = r[0], r[1], r[2]
m = {k: (v[i], u[i]) for i, k in enumerate(h)}
def show(title, keys):
    print(title)
    for k in keys:
        if k in m:
            print("  %-78s %s %s" % (k, m[k][0], m[k][1]))
print("report: %s\nkernel: %s" % (os.path.basename(rep), m.get("Kernel Name", ("?",))[0]))
show("launch", ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
                "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
                "sm__maximum_warps_per_active_cycle_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"])
show("time and issue", ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
                        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
                        "smsp__thread_inst_executed_per_inst_executed.ratio"])
show("pipe utilisation (instructions issued to the pipe, % of its peak while the SM is active)",
     [k for k in h if k.startswith("sm__inst_executed_pipe_") and k.endswith(".avg.pct_of_peak_sustained_active") and float(m[k][0] or 0) > 0])
show("stalls per issued instruction", sorted([k for k in h if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k],
                                             key=lambda k: -float(m[k][0] or 0)))
show("memory", ["dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
                "l1tex__t_sector_hit_rate.pct", "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct",
                "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
                "smsp__sass_inst_executed_op_shared_ld.sum", "smsp__sass_inst_executed_op_shared_st.sum",
                "smsp__sass_inst_executed_op_global_ld.sum", "smsp__sass_inst_executed_op_global_st.sum"])
print()
env = dict(os.environ, STALLS="1")
so = os.path.join(os.path.dirname(here), "signalalign_b200", "libsagpu.so")
out = subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), rep, so, kern, "/tmp/ncu_lines_%d.txt" % os.getpid(), "30"],
                     capture_output=True, text=True, env=env)
print("per source line (stall samples first, then executed warp instructions):")
print(out.stdout)
# dynamic opcode histogram from the per-instruction listing
import collections, re
hist = collections.Counter()
for line in open("/tmp/ncu_lines_%d.txt" % os.getpid()):
    p = line.split(None, 2)
    if len(p) == 3:
        mm = re.match(r"(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", p[2].strip())
        if mm:
            hist[mm.group(1).split(".")[0]] += int(p[1])
tot = sum(hist.values())
print("executed warp instructions by opcode (dynamic):")
print("  " + ", ".join("%s %.1f%%" % (k, 100.0 * n / tot) for k, n in hist.most_common(28)))
