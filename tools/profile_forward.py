#!/usr/bin/env python
"""Run the ensemble-forward kernel (mopo_ens_forward) a few times on synthetic HalfCheetah-shaped input.
Used under ncu (profiles/) and for quick CUDA-event timing: python tools/profile_forward.py [B] [reps] [prec]"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from mopo_b200 import _lib
from mopo_b200.models.constructor import construct_model
from oracle import ensemble as ens

B = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
prec = {'bf16': _lib.PREC_BF16, 'fp32': _lib.PREC_FP32}[sys.argv[3] if len(sys.argv) > 3 else 'bf16']
obs_dim, act_dim = 17, 6
sys.stdout, so = open(os.devnull, 'w'), sys.stdout
p = ens.init_params(obs_dim, act_dim, seed=0, separate_mean_var=True)
model = construct_model(obs_dim=obs_dim, act_dim=act_dim, separate_mean_var=True, name='prof', device=0, precision=prec)
model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
sys.stdout = so
lib = _lib.load()
dev = torch.device('cuda', 0)
xn = torch.randn(B, obs_dim + act_dim, device=dev)
mean = torch.empty(7, B, obs_dim + 1, device=dev)
lv = torch.empty_like(mean)
st = torch.cuda.current_stream(dev)
evs = []
for i in range(reps + 3):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(st)
    _lib.check(lib.mopo_ens_forward(model._handle, xn.data_ptr(), B, mean.data_ptr(), lv.data_ptr(), prec, st.cuda_stream))
    e.record(st)
    evs.append((s, e))
torch.cuda.synchronize()
ms = sorted(s.elapsed_time(e) for s, e in evs[3:])
fl = 2 * 7 * (23 * 200 + 3 * 200 * 200 + 2 * 200 * 18) * B
print('forward B=%d: median %.4f ms, min %.4f ms, %.1f TFLOP/s (min)' % (B, ms[len(ms) // 2], ms[0], fl / ms[0] / 1e9))
