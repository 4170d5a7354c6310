#!/usr/bin/env python
"""Where the end-to-end (host buffers) rollout time goes: H2D of starts + actions, rollout, D2H read-back."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from mopo_b200 import _lib
from mopo_b200.models.constructor import construct_model
from mopo_b200.models.fake_env import FakeEnv
from mopo_b200.replay_pools import SimpleReplayPool
from mopo_b200.algorithms.mopo import ModelRollout
import mopo_b200.static as static
from oracle import ensemble as ens

B, h = 50000, 5
so, sys.stdout = sys.stdout, open(os.devnull, 'w')
p, obs, act = ens.synthetic_case('halfcheetah', B, seed=0)
model = construct_model(obs_dim=17, act_dim=6, separate_mean_var=True, name='prof', device=0, precision=_lib.PREC_BF16)
model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
model._model_inds = [0, 1, 2, 3, 4]
env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
ro = ModelRollout(env, SimpleReplayPool(17, 6, 16, device=0), rollout_batch_size=B, rollout_schedule=(20, 100, h, h), seed=1)
ro._set_rollout_length(); ro._reallocate_model_pool()
sys.stdout = so
dev = torch.device('cuda', 0)
h_obs = torch.as_tensor(obs).pin_memory()
h_acts = [torch.as_tensor(act).pin_memory() for _ in range(h)]
h_rew = torch.empty(B, 1).pin_memory()
st = torch.cuda.current_stream()

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(st); return e

for it in range(6):
    t0 = time.perf_counter(); e0 = ev()
    o = h_obs.to(dev, non_blocking=True); a = [x.to(dev, non_blocking=True) for x in h_acts]
    t1 = time.perf_counter(); e1 = ev()
    ro._rollout_model(B, start_obs=o, actions=a, quiet=True)
    t2 = time.perf_counter(); e2 = ev()
    pool = ro._model_pool
    last = (pool._pointer - B) % pool._max_size
    h_rew.copy_(pool.fields['rewards'][last:last + B], non_blocking=True)
    t3 = time.perf_counter(); e3 = ev()
    st.synchronize()
    t4 = time.perf_counter()
    print('it %d host ms: h2d %.3f rollout %.3f d2h %.3f sync %.3f | device ms: h2d %.3f rollout %.3f d2h %.3f' % (
        it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3)))
