#!/usr/bin/env python3
"""Static SASS size of every kernel and every out-of-line device function in it.
usage: sass_sizes.py <nvdisasm -c output>   (cuobjdump -xelf all lib.so; nvdisasm -g -c x.cubin > dis.txt)"""
import re, sys, subprocess
cur = None
sizes = {}
order = []
for ln in open(sys.argv[1]):
    m = re.match(r"^(\$?[_A-Za-z][^\s:]*):\s*$", ln)
    if m and not m.group(1).startswith(".L"):
        cur = m.group(1)
        if cur not in sizes:
            sizes[cur] = 0
            order.append(cur)
        continue
    if cur and re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        sizes[cur] += 1
def dem(n):
    parts = [p for p in n.split("$") if p]
    out = []
    for p in parts:
        try:
            out.append(subprocess.run(["c++filt", p], capture_output=True, text=True).stdout.strip().split("(")[0])
        except Exception:
            out.append(p)
    return " :: ".join(out)
for n in order:
    if sizes[n]:
        print(f"{sizes[n]:7d}  {dem(n)}")
