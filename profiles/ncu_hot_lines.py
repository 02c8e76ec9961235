#!/usr/bin/env python3
"""Dynamic profile of one `ncu --set full --import-source on` capture by SOURCE LINE:
stall samples, executed warp-instructions and active lanes per instruction.
usage: python profiles/ncu_hot_lines.py capture.ncu-rep [top]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
cur_file, cur_line, cur_src, hdr = "?", "", "", None
acc = collections.OrderedDict()
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if len(r) >= 2 and r[0] == "Function Name":
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = {n: i for i, n in enumerate(r)}
        i_s, i_i, i_t = r.index("# Samples"), r.index("Instructions Executed"), r.index("Thread Instructions Executed")
        continue
    if hdr is None or len(r) <= i_t:
        continue
    if r[0] != "":
        cur_line, cur_src = r[0], r[1]
        continue
    try:
        v = (int(r[i_s]), int(r[i_i]), int(r[i_t]))
    except ValueError:
        continue
    k = (cur_file, cur_line, cur_src.strip())
    a = acc.setdefault(k, [0, 0, 0, 0])
    a[0] += v[0]; a[1] += v[1]; a[2] += v[2]; a[3] += 1
ts = sum(a[0] for a in acc.values()) or 1
ti = sum(a[1] for a in acc.values()) or 1
tt = sum(a[2] for a in acc.values())
print("# %s: %d stall samples, %d warp-instructions executed, %.2f active lanes / instruction" % (rep, ts, ti, tt / ti))
byf = collections.defaultdict(lambda: [0, 0, 0, 0])
for (f, _, _), a in acc.items():
    for j in range(4):
        byf[f][j] += a[j]
print("## by file: samples, executed, lanes, static SASS")
for f, a in sorted(byf.items(), key=lambda x: -x[1][0]):
    print("%-22s %5.1f%% %5.1f%% %5.1f %6d" % (f, 100.0 * a[0] / ts, 100.0 * a[1] / ti, a[2] / max(1, a[1]), a[3]))
print("## top source lines by stall samples: file:line samples executed lanes static | source")
for (f, ln, src), a in sorted(acc.items(), key=lambda x: -x[1][0])[:top]:
    print("%-18s %5s %5.1f%% %5.1f%% %5.1f %5d | %s" % (f, ln, 100.0 * a[0] / ts, 100.0 * a[1] / ti, a[2] / max(1, a[1]), a[3], src[:100]))
