#!/usr/bin/env python
"""launch_summary.py <launches.csv> [first_id last_id] -- per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` log."""
import collections
import csv
import sys

rows, hdr = [], None
with open(sys.argv[1]) as f:
    for line in f:
        if line.startswith('"ID"'):
            hdr = next(csv.reader([line]))
            break
    for r in csv.reader(f):
        if hdr and len(r) == len(hdr):
            rows.append(dict(zip(hdr, r)))
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else len(rows)
rows = [r for r in rows if lo <= int(r["ID"]) < hi]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    k = r["Kernel Name"].split("(")[0].replace("void ", "").replace("mgb::", "").replace("<unnamed>::", "")[:60]
    agg[k][0] += 1
    agg[k][1] += float(r["Metric Value"])
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / 1e3:.1f} us total")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{v[1] / 1e3:10.1f} us {v[0]:6d} {100 * v[1] / tot:5.1f}%  {k}")
