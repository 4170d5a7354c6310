#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total and share per kernel.
usage: python tools/launch_summary.py launches.csv > summary.txt"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ik, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    if len(r) <= iv or 'gpu__time_duration' not in ','.join(r):
        continue
    v = float(r[iv].replace(',', ''))
    if v != v:                                   # a launch ncu could not time (the hot path under a programmatic launch)
        continue
    v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(r[iu], 1.0)
    tot[r[ik]] += v
    cnt[r[ik]] += 1
s = sum(tot.values())
for k in sorted(tot, key=lambda k: -tot[k]):
    print('%-70s %6d launches %11.1f us total %9.1f us avg %5.1f%%' % (k[:70], cnt[k], tot[k], tot[k] / cnt[k], 100 * tot[k] / s))
