#!/usr/bin/env python
"""Attribute an ncu SASS source page (--page source --csv) to CUDA source lines.
usage: ncu_lines.py <src.csv> <cubin> <kernel-substring> [top]
Uses nvdisasm -g line info of the cubin; instruction order in both listings is the same."""
import collections
import csv
import re
import subprocess
import sys

srccsv, cubin, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
lines, infn, cur = [], False, None
for ln in dis.splitlines():
    if ln.startswith(".text."):
        infn = kname in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
rows = list(csv.reader(open(srccsv)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
ci, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
body = rows[h + 1:]
print(f"# sass rows {len(body)}  disasm instrs {len(lines)}")
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for i, r in enumerate(body):
    key = lines[i] if i < len(lines) else None
    a = agg[key]
    a[0] += int(r[ci]); a[1] += int(r[ti]); a[2] += int(r[si]); a[3] += 1
tot = sum(a[0] for a in agg.values()); tots = sum(a[2] for a in agg.values())
print(f"# total warp instr {tot}  samples {tots}")
for k, a in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
    print(f"{k}  sass={a[3]}  inst={a[0]} ({a[0]/tot:.3f})  samples={a[2]} ({a[2]/tots:.3f})")
