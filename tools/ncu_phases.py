#!/usr/bin/env python
"""ncu source page (SASS view, --csv) -> instructions executed / stall samples per code region between block barriers
usage: tools/ncu_phases.py X_source.csv [extra split mnemonics, comma separated]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {k: i for i, k in enumerate(hdr)}
split = ["BAR.SYNC"] + (sys.argv[2].split(",") if len(sys.argv) > 2 else [])
tot_i = sum(int(r[ix["Instructions Executed"]]) for r in rows[2:] if len(r) > 10)
tot_s = sum(int(r[ix["# Samples"]]) for r in rows[2:] if len(r) > 10)
reg, acc = [], dict(n=0, inst=0, thr=0, samp=0, first=None, stalls={})
stall_cols = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
for r in rows[2:]:
    if len(r) < 10:
        continue
    src = r[ix["Source"]].strip()
    if acc["first"] is None:
        acc["first"] = r[ix["Address"]][-5:]
    acc["n"] += 1
    acc["inst"] += int(r[ix["Instructions Executed"]])
    acc["thr"] += int(r[ix["Thread Instructions Executed"]])
    acc["samp"] += int(r[ix["# Samples"]])
    for k in stall_cols:
        acc["stalls"][k] = acc["stalls"].get(k, 0) + int(r[ix[k]] or 0)
    if any(s in src for s in split):
        acc["end"] = src[:40]
        reg.append(acc)
        acc = dict(n=0, inst=0, thr=0, samp=0, first=None, stalls={})
reg.append(acc)
print(f"total instructions {tot_i}, samples {tot_s}")
for i, a in enumerate(reg):
    if a["inst"] == 0:
        continue
    top = sorted(a["stalls"].items(), key=lambda kv: -kv[1])[:3]
    print(f"{i:3d} @{a['first']} sass {a['n']:5d}  inst {100*a['inst']/tot_i:5.1f}%  samples {100*a['samp']/max(tot_s,1):5.1f}%  lanes {a['thr']/max(a['inst'],1):4.1f}  "
          + " ".join(f"{k[6:]}={100*v/max(a['samp'],1):.0f}%" for k, v in top) + f"   ..{a.get('end','')}")
