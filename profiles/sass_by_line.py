#!/usr/bin/env python3
"""Static SASS size per source line / file of one kernel (nvdisasm -g -c output).
usage: sass_by_line.py dis.txt <kernel-substring> [top]"""
import re, sys, collections
path, key = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cur = None
infn = False
by_line = collections.Counter()
by_file = collections.Counter()
total = 0
for ln in open(path):
    if ln.startswith(".text."):
        infn = key in ln
        continue
    if infn and ln.startswith("$"):  # out-of-line device functions follow the kernel body
        infn = False
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        total += 1
        by_line[cur] += 1
        by_file[cur[0] if cur else None] += 1
print("total instructions", total)
for f, n in by_file.most_common():
    print(f"{n:7d} {100.0*n/total:5.1f}%  {f}")
print()
for (k, n) in by_line.most_common(top):
    print(f"{n:7d}  {k}")
