#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python profiles/launch_summary.py gpurun_out/launches.csv > profiles/<name>.txt"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        hdr, start = r, i + 1
        break
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = {}, collections.Counter()
for r in rows[start:]:
    if len(r) <= mv:
        continue
    name = re.sub(r"\(.*", "", r[kn]).replace("void ", "").replace("mb200::", "").replace("<unnamed>::", "")
    v = float(r[mv].replace(",", ""))
    v = v / 1e3 if r[mu] == "ns" else (v * 1e3 if r[mu] == "ms" else v)
    tot[name] = tot.get(name, 0.0) + v
    cnt[name] += 1
T = sum(tot.values())
print(f"launches {sum(cnt.values())}  total {T:.1f} us (per-launch times are cold-cache and serialised: compare shares)")
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    print(f"{v:12.1f} us {100 * v / T:5.1f}%  n={cnt[k]:5d}  avg={v / cnt[k]:9.1f} us  {k}")
