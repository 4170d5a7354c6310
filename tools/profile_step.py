#!/usr/bin/env python
"""Run the fused FakeEnv.step (tensor-core path, HalfCheetah shapes) a few times: for ncu / quick timing.
usage: python tools/profile_step.py [B] [reps]"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from mopo_b200 import _lib
from mopo_b200.models.constructor import construct_model
from mopo_b200.models.fake_env import FakeEnv
import mopo_b200.static as static
from oracle import ensemble as ens

B = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
so, sys.stdout = sys.stdout, open(os.devnull, 'w')
p, obs, act = ens.synthetic_case('halfcheetah', B, seed=0)
model = construct_model(obs_dim=17, act_dim=6, separate_mean_var=True, name='prof', device=0, precision=_lib.PREC_BF16)
model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
model._model_inds = [0, 1, 2, 3, 4]
sys.stdout = so
env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
d_obs, d_act = torch.as_tensor(obs).cuda(), torch.as_tensor(act).cuda()
st = torch.cuda.current_stream()
evs = []
for i in range(reps + 3):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(st)
    env.step(d_obs, d_act, diagnostics=False)
    e.record(st)
    evs.append((s, e))
torch.cuda.synchronize()
ms = sorted(s.elapsed_time(e) for s, e in evs[3:])
print('fused step B=%d: median %.4f ms, min %.4f ms' % (B, ms[len(ms) // 2], ms[0]))
