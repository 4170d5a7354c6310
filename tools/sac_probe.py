#!/usr/bin/env python
"""Time of one SAC update (mopo_sac_update_from_pools: one cooperative kernel per update) on HalfCheetah shapes with a
100k + 250k pool pair.   usage: python tools/sac_probe.py [n_updates] [--trace]
--trace runs on the probe build of the library and prints the kernel's stage timeline (CTA 0's %globaltimer after every
stage barrier, median over the updates)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
trace = '--trace' in sys.argv
argv = [a for a in sys.argv[1:] if not a.startswith('--')]
if trace:                                    # the product calls and the probe must go through the same library instance
    from mopo_b200 import build as _b
    os.environ['MOPO_B200_LIB'] = _b.build(probes=True)
import numpy as np
import torch
from mopo_b200 import _lib
from mopo_b200.replay_pools import SimpleReplayPool
from mopo_b200.algorithms.sac import SACUpdate

n_upd = int(argv[0]) if argv else 1000
od, ad = 17, 6
rng = np.random.RandomState(0)
pools = []
for n in (100000, 250000):
    p = SimpleReplayPool(od, ad, n, device=0)
    p.add_samples({'observations': rng.standard_normal((n, od)).astype(np.float32), 'actions': np.tanh(rng.standard_normal((n, ad))).astype(np.float32),
                   'next_observations': rng.standard_normal((n, od)).astype(np.float32), 'rewards': rng.standard_normal((n, 1)).astype(np.float32),
                   'terminals': np.zeros((n, 1), bool)})
    pools.append(p)
sac = SACUpdate(od, ad, device=0, seed=7)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    sac.train_from_pools(pools[0], pools[1], n_updates=50)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(side)
    sac.train_from_pools(pools[0], pools[1], n_updates=n_upd)
    e1.record(side)
    torch.cuda.synchronize()
out = {'ms_per_sac_update': e0.elapsed_time(e1) / n_upd, 'n_updates': n_upd}
if trace:
    dbg = _lib.load_probes()
    buf = torch.zeros(64, dtype=torch.int64, device='cuda')
    _lib.check(dbg.mopo_debug_sac_trace(sac._handle, buf.data_ptr()), dbg)
    rows = []
    for _ in range(50):
        sac.train_from_pools(pools[0], pools[1], n_updates=1)
        torch.cuda.synchronize()
        rows.append(buf.cpu().numpy().copy())
    _lib.check(dbg.mopo_debug_sac_trace(sac._handle, None), dbg)
    t = np.stack(rows)
    n_marks = int((t[0] > 0).sum())
    d = np.median(np.diff(t[:, :n_marks], axis=1), axis=0) * 1e-3
    out['stage_us'] = [round(float(x), 2) for x in d]          # the last entry is the Adam / Polyak tail (no barrier after it)
    out['kernel_us'] = round(float(np.median(t[:, n_marks - 1] - t[:, 0])) * 1e-3, 2)
print(json.dumps(out))
