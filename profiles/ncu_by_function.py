#!/usr/bin/env python3
"""Attribute an ncu capture to out-of-line device functions and source lines.
usage: ncu_by_function.py <report.ncu-rep> <nvdisasm -g -c listing of the same build> <kernel substring> [top]
The SASS page of the report gives per-instruction executed counts and stall samples by address;
the listing gives, for the same kernel, the function label and file:line of every offset."""
import csv, io, re, subprocess, sys, collections
rep, dis, key = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
# ---- listing: offset -> (function, file:line) for the kernel's section
fn_of, line_of = {}, {}
insec = False
cur_fn, cur_line = None, None
for ln in open(dis):
    if ln.startswith(".text."):
        insec = key in ln
        cur_fn = "(kernel body)"
        continue
    if not insec:
        continue
    m = re.match(r"^\$(\S+):\s*$", ln)
    if m:
        parts = [p for p in m.group(1).split("$") if p]
        name = parts[-1]
        try:
            name = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0]
        except Exception:
            pass
        cur_fn = name
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur_line = "%s:%s" % (m.group(1).split("/")[-1], m.group(2))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m:
        off = int(m.group(1), 16)
        fn_of[off] = cur_fn
        line_of[off] = cur_line
# ---- report
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ia, ie, is_, it = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
base = None
F = collections.defaultdict(lambda: [0, 0, 0])
Lc = collections.defaultdict(lambda: [0, 0, 0])
for r in rows[hdr_i + 1:]:
    if len(r) <= it or not r[ia].startswith("0x"):
        continue
    a = int(r[ia], 16)
    if base is None:
        base = a
    off = a - base
    ex, sm, th = int(r[ie] or 0), int(r[is_] or 0), int(r[it] or 0)
    for tab, k in ((F, fn_of.get(off, "?")), (Lc, line_of.get(off, "?"))):
        tab[k][0] += ex
        tab[k][1] += sm
        tab[k][2] += th
tot_e = sum(v[0] for v in F.values()) or 1
tot_s = sum(v[1] for v in F.values()) or 1
print("# %s  kernel~%s: %d warp-instructions, %d stall samples" % (rep, key, tot_e, tot_s))
print("## by function: %executed  %samples  threads/inst")
for k, v in sorted(F.items(), key=lambda kv: -kv[1][1]):
    print("%6.1f%% %6.1f%% %5.1f  %s" % (100.0 * v[0] / tot_e, 100.0 * v[1] / tot_s, v[2] / max(v[0], 1), k))
print("## by source line (top %d by samples)" % top)
for k, v in sorted(Lc.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%6.1f%% %6.1f%% %5.1f  %s" % (100.0 * v[0] / tot_e, 100.0 * v[1] / tot_s, v[2] / max(v[0], 1), k))
