#!/usr/bin/env python
"""Decode the tcgen05 forward timeline (mopo_debug_tc_trace) and run the pipe-rate microbenchmarks.
usage (on a GPU box): python tools/tc_timeline.py [B] > gpurun_out/timeline.txt"""
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
lib = _lib.load_probes()
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream(dev)

names = {0: 'tanh.f32', 1: 'tanh.bf16x2', 2: 'ex2.f32', 3: 'tanh.f16x2', 4: 'fma.bf16x2', 5: 'fma.f32', 6: 'rcp.f32', 7: 'swish(ex2+rcp)', 8: 'cvt.f16x2.f32', 9: 'cvt.f32.f16+sub', 10: 'ex2.f16x2', 11: 'ex2.bf16x2'}
for kind in (range(12) if os.environ.get('TL_PIPES', '1') == '1' else ()):
    nb, iters = 148, 2000
    cyc = torch.zeros(nb + 1, dtype=torch.int64, device=dev)
    for _ in range(2):
        _lib.check(lib.mopo_debug_pipe_bench(kind, iters, cyc.data_ptr(), nb, st.cuda_stream))
    torch.cuda.synchronize()
    c = cyc[:nb].cpu().numpy().astype(np.float64)
    per_clk = 512 * 8 * iters / np.median(c)
    print('pipe bench %-12s: %.2f instr-results/clk/SM (x2 elements if packed)' % (names[kind], per_clk))

obs_dim, act_dim = 17, 6
so, sys.stdout = sys.stdout, open(os.devnull, 'w')
p = ens.init_params(obs_dim, act_dim, seed=0, separate_mean_var=True)
model = construct_model(obs_dim=obs_dim, act_dim=act_dim, separate_mean_var=True, name='prof', device=0, precision=_lib.PREC_BF16)
model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
sys.stdout = so
xn = torch.randn(B, obs_dim + act_dim, device=dev)
mean = torch.empty(7, B, obs_dim + 1, device=dev)
lv = torch.empty_like(mean)
trace = torch.zeros(5 * 2048, dtype=torch.int64, device=dev)
fused = os.environ.get('TL_FUSED', '0') == '1'
print('mode:', 'fused head' if fused else 'full outputs')
tc3 = int(os.environ.get('TL_TC3', '0'))          # 3 / 1: the streamed-weights kernel (ens_tc3.cu) in fp32-accurate / bf16 mode
for _ in range(int(os.environ.get('TL_REPS', '2'))):
    if tc3:
        _lib.check(lib.mopo_debug_tc3_trace(model._handle, xn.data_ptr(), B, None if fused else mean.data_ptr(), None if fused else lv.data_ptr(), tc3, trace.data_ptr(), st.cuda_stream), lib)
    else:
        _lib.check(lib.mopo_debug_tc_trace(model._handle, xn.data_ptr(), B, None if fused else mean.data_ptr(), None if fused else lv.data_ptr(), trace.data_ptr(), st.cuda_stream), lib)
torch.cuda.synchronize()
t = trace.cpu().numpy().astype(np.uint64).reshape(5, 2048)
if tc3 and t[4, -1]:
    cyc, ns = int(t[4, -2]) - int(t[4, -4]), int(t[4, -1]) - int(t[4, -3])
    print('issuer loop of pair 0: %d cycles in %d ns -> SM clock %.0f MHz under this load' % (cyc, ns, cyc / max(ns, 1) * 1e3))
    t[4, -4:] = 0
    pp = t[4, 1024:1024 + 512].astype(np.int64).reshape(-1, 2)
    pp = pp[pp[:, 1] > 0]
    if len(pp):
        base = pp[:, 0].min()
        print('per-pair issuer loops (ns from the first start): start min/median/max %d/%d/%d, end min/median/max %d/%d/%d, duration min/median/max %d/%d/%d' % (
            (pp[:, 0] - base).min(), np.median(pp[:, 0] - base), (pp[:, 0] - base).max(), (pp[:, 1] - base).min(), np.median(pp[:, 1] - base), (pp[:, 1] - base).max(),
            (pp[:, 1] - pp[:, 0]).min(), np.median(pp[:, 1] - pp[:, 0]), (pp[:, 1] - pp[:, 0]).max()))
        print('per-pair durations (us):', ' '.join('%.0f' % (d / 1e3) for d in (pp[:, 1] - pp[:, 0])))
    t[4, 1024:1024 + 512] = 0
ev = []
for role in range(5):
    for v in t[role]:
        if v == 0:
            break
        ev.append((int(v & np.uint64(0xFFFFFFFFFFFF)), role, int(v >> np.uint64(48))))
ev.sort()
if ev:
    t0 = ev[0][0]
    print('events: %d, span %d cycles' % (len(ev), ev[-1][0] - t0))
    role_names = ['MMA ', 'WG0 ', 'WG1 ', 'LOAD', 'HEAD']
    last = {}
    lim = int(os.environ.get('TL_MAX', '400'))
    for c, role, tag in ev[:lim]:
        d = c - last.get(role, c)
        last[role] = c
        print('%9d  %s%s tag %3d  (+%d)' % (c - t0, '    ' * role * 3, role_names[role], tag, d))
