#!/usr/bin/env python
"""Run a few device-resident HalfCheetah-shaped rollouts (B x h, bf16 tensor-core path): the ncu / timing target.
usage: python tools/profile_rollout.py [B] [h] [reps] [policy]   (policy: act = policy(obs) on the device each step)"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from mopo_b200 import _lib
if os.environ.get('MOPO_LIB'):            # A/B runs: another build of the library (same ABI)
    _lib.LIB_PATH = os.path.abspath(os.environ['MOPO_LIB'])
from mopo_b200.models.constructor import construct_model
from mopo_b200.models.fake_env import FakeEnv
from mopo_b200.replay_pools import SimpleReplayPool
from mopo_b200.algorithms.mopo import ModelRollout
import mopo_b200.static as static
from oracle import ensemble as ens

B = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
h = int(sys.argv[2]) if len(sys.argv) > 2 else 5
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
use_policy = len(sys.argv) > 4 and sys.argv[4] == 'policy'
so, sys.stdout = sys.stdout, open(os.devnull, 'w')
p, obs, act = ens.synthetic_case('halfcheetah', B, seed=0)
model = construct_model(obs_dim=17, act_dim=6, separate_mean_var=True, name='prof', device=0, precision=_lib.PREC_BF16)
model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
model._model_inds = [0, 1, 2, 3, 4]
env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
pol = None
if use_policy:
    from mopo_b200.policies import GaussianPolicy
    from oracle import policy as opol
    pol = GaussianPolicy(17, 6, device=0, precision=_lib.PREC_BF16, seed=1)
    pol.set_params(*opol.init_params(17, 6, seed=0))
ro = ModelRollout(env, SimpleReplayPool(17, 6, 16, device=0), rollout_batch_size=B, rollout_schedule=(20, 100, h, h), seed=1, policy=pol)
ro._set_rollout_length()
ro._reallocate_model_pool()
sys.stdout = so
d_obs = torch.as_tensor(obs).cuda()
d_acts = [torch.as_tensor(act).cuda() for _ in range(h)]
st = torch.cuda.current_stream()
evs = []
for i in range(reps):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(st)
    ro._rollout_model(B, start_obs=d_obs, actions=None if use_policy else d_acts, quiet=True)
    e.record(st)
    evs.append((s, e))
torch.cuda.synchronize()
print('rollout B=%d h=%d%s: ms per rollout %s' % (B, h, ' (device policy)' if use_policy else '', ['%.4f' % s.elapsed_time(e) for s, e in evs]))
