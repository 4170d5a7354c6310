#!/usr/bin/env python
"""Time of one SAC update (mopo_sac_update_from_pools, graph replay) on HalfCheetah shapes with a 100k + 250k pool pair.
usage: python tools/sac_probe.py [n_updates]"""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mopo_b200.replay_pools import SimpleReplayPool
from mopo_b200.algorithms.sac import SACUpdate

n_upd = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
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
    logs = sac.train_from_pools(pools[0], pools[1], n_updates=n_upd, return_losses=True)
    e1.record(side)
    torch.cuda.synchronize()
print(json.dumps({'ms_per_sac_update': e0.elapsed_time(e1) / n_upd, 'losses': {k: float(v) for k, v in logs.items()}}))
