#!/usr/bin/env python
"""Summarise a tools/tc_timeline.py dump: per-stage epilogue durations and the super-tile period."""
import re, sys
ev = []
for line in open(sys.argv[1]):
    m = re.match(r'\s*(\d+)\s+(MMA|WG0|WG1|LOAD)\s+tag\s+(\d+)', line)
    if m:
        ev.append((int(m.group(1)), m.group(2), int(m.group(3))))
for role in ('WG0', 'WG1'):
    st, wait, dur = {}, {}, {}
    for c, r, t in ev:
        if r != role:
            continue
        if 100 <= t < 200: st[('w', t - 100)] = c
        if 200 <= t < 300:
            st[t - 200] = c
            if ('w', t - 200) in st: wait.setdefault(t - 200, []).append(c - st[('w', t - 200)])
        if 300 <= t < 400 and (t - 300) in st: dur.setdefault(t - 300, []).append(c - st[t - 300])
    med = lambda v: sorted(v)[len(v) // 2]
    print(role, 'epilogue cycles per stage (10*layer+tile): ', {k: med(v) for k, v in sorted(dur.items())})
    print(role, 'wait-for-accumulator cycles per stage:      ', {k: med(v) for k, v in sorted(wait.items())})
t100 = [c for c, r, t in ev if r == 'WG0' and t == 100]
print('super-tile period (cycles):', [b - a for a, b in zip(t100, t100[1:])])
