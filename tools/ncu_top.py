#!/usr/bin/env python
"""Summarise an ncu `--page source --csv` dump: top SASS instructions by warp-stall samples, with the dominant
stall reason.  usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/ncu_top.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
S = ix['# Samples']
tot = sum(float(r[S] or 0) for r in body)
print('instructions: %d, total samples: %d' % (len(body), tot))
agg = {c: sum(float(r[ix[c]] or 0) for r in body) for c in stall_cols}
print('by reason:', ', '.join('%s=%.1f%%' % (c[6:], 100 * v / max(tot, 1)) for c, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
order = sorted(range(len(body)), key=lambda i: -float(body[i][S] or 0))[:N]
for i in sorted(order):
    r = body[i]
    top = max(stall_cols, key=lambda c: float(r[ix[c]] or 0))
    print('%5d %6s (%4.1f%%) %-14s exec=%-8s %s' % (i, r[S], 100 * float(r[S] or 0) / max(tot, 1), top[6:], r[ix['Instructions Executed']], r[ix['Source']].strip()[:100]))
