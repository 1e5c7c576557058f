#!/usr/bin/env python3
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total ms and share.
usage: tools/ncu_summary.py gpurun_out/launches.csv [substring-filter]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
flt = sys.argv[2] if len(sys.argv) > 2 else ""
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0]
    if flt and flt not in name:
        continue
    v = float(r[vi].replace(",", ""))
    v = v / 1e6 if r[ui] == "ns" else v / 1e3 if r[ui] == "us" else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}  filter={flt!r}  total {tot:.3f} ms over {sum(a[0] for a in agg.values())} launches")
print(f"{'total ms':>10} {'n':>5} {'avg ms':>9} {'share':>7}  kernel")
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1]:10.3f} {a[0]:5d} {a[1] / a[0]:9.4f} {100 * a[1] / tot:6.1f}%  {n[:100]}")
