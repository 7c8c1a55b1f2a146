#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total time, share."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    agg[r[ki][:100]][0] += 1
    agg[r[ki][:100]][1] += v
tot = sum(v[1] for v in agg.values())
print("launches   total_ms  share  kernel")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%8d %10.3f %5.1f%%  %s" % (v[0], v[1] / 1e6, 100 * v[1] / tot, k))
