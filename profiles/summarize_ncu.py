#!/usr/bin/env python3
"""Turns an .ncu-rep (one `ncu --set full --import-source on` capture of k_trace_pool) into the
text summary committed next to it: headline counters, SASS opcode mix, stall reasons.
usage: python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread ",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__sass_average_branch_targets_threads_uniform.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum ", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "smsp__cycles_active.avg ", "sm__cycles_elapsed.max "]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
print("# %s" % rep)
for k_row in rows[2:]:
    print("## kernel: %s" % k_row[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "## kernel")
    for i, h in enumerate(hdr):
        if any((h + " ").startswith(k) or h == k.strip() for k in KEYS):
            print("%-75s %-14s %s" % (h, units[i], k_row[i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
ops, smp, stalls = collections.Counter(), collections.Counter(), collections.Counter()
tot = thr = n_sass = 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    toks = r[idx["Source"]].split()
    m = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
    n = int(r[idx["Instructions Executed"]])
    ops[m] += n
    smp[m] += int(r[idx["# Samples"]])
    tot += n
    n_sass += 1
    thr += int(r[idx["Thread Instructions Executed"]])
    for c in hdr:
        if c.startswith("stall_") and "Not Issued" not in c:
            try:
                stalls[c] += int(r[idx[c]])
            except ValueError:
                pass
print("\n## SASS: %d instructions in the kernel image, %d warp-instructions executed, %.2f active threads / warp-instruction"
      % (n_sass, tot, thr / max(1, tot)))
ts = max(1, sum(smp.values()))
for m, n in ops.most_common(18):
    print("%-8s %5.1f%% of executed  %5.1f%% of stall samples" % (m, 100.0 * n / tot, 100.0 * smp[m] / ts))
print("\n## warp stall reasons (all samples)")
s = max(1, sum(stalls.values()))
for c, n in stalls.most_common(8):
    print("%-26s %5.1f%%" % (c, 100.0 * n / s))
