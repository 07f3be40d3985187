#!/usr/bin/env python
"""summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total ns, share"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[h + 1:]:
    if len(r) <= mv:
        continue
    name = r[kn].split("(")[0]
    agg[name][0] += 1
    agg[name][1] += float(r[mv].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print(f"# unit=ns total={tot:.0f}")
print("kernel,launches,total,share")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k},{v[0]},{v[1]:.0f},{v[1] / tot:.4f}")
