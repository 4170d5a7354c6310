#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if 'Kernel Name' in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        try:
            v = float(d['Metric Value'].replace(',', ''))
        except ValueError:
            continue
        k = d['Kernel Name'][:64]
        agg[k][0] += 1
        agg[k][1] += v
tot = sum(t for _, t in agg.values())
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print('%-66s %5d launches %10.1f us total %8.1f us avg %5.1f%%' % (k, c, t / 1e3, t / 1e3 / c, 100 * t / tot))
