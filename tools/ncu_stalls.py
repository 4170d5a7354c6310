#!/usr/bin/env python
"""Summarise the per-instruction warp-stall samples of an ncu report (needs -lineinfo / --import-source on):
usage: ncu -i X.ncu-rep --page source --csv > src.csv; python tools/ncu_stalls.py src.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr)]
reasons = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]


def f(r, k):
    try:
        return float(r[ix[k]])
    except ValueError:
        return 0.0


tot = sum(f(r, '# Samples') for r in data)
print(rows[0][1] if len(rows[0]) > 1 else '')
print('instructions: %d, samples: %d' % (len(data), tot))
print('by reason: ' + ', '.join('%s=%.1f%%' % (k[6:], 100 * v / tot) for k, v in
                                sorted(((k, sum(f(r, k) for r in data)) for k in reasons), key=lambda kv: -kv[1]) if v > 0.01 * tot))
for reason in ('stall_no_inst', 'stall_long_sb', 'stall_wait'):
    by = collections.Counter()
    for r in data:
        src = r[ix['Source']].split()
        op = (src[1] if src and src[0].startswith('@') and len(src) > 1 else (src[0] if src else '')).split('.')[0]
        by[op] += f(r, reason)
    print('%s by opcode: ' % reason[6:] + ', '.join('%s %d' % kv for kv in by.most_common(8)))
for r in sorted(data, key=lambda r: -f(r, '# Samples'))[:top]:
    why = {k[6:]: int(f(r, k)) for k in reasons if f(r, k) > 0.15 * f(r, '# Samples')}
    print('%6d (%4.1f%%) %-60s %s' % (f(r, '# Samples'), 100 * f(r, '# Samples') / tot, r[ix['Source']][:60], why))
