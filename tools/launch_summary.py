#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) < 15:
        continue
    name = r[4].split("(")[0]
    agg[name][0] += 1
    agg[name][1] += float(r[-1].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print("%-28s %6s %12s %7s %11s" % ("kernel", "n", "total_us", "share", "avg_us"))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-28s %6d %12.1f %6.1f%% %11.1f" % (k, v[0], v[1] / 1e3, 100 * v[1] / tot, v[1] / 1e3 / v[0]))
print("%-28s %6d %12.1f" % ("TOTAL", sum(v[0] for v in agg.values()), tot / 1e3))
