"""Developer probe: the fp32-accurate tcgen05 forward (MOPO_PREC_FP32, ens_tc3.cu) against the float64 oracle and the
FFMA checker kernel (MOPO_PREC_FP32_REF), error and time per shape / batch.  Not on any product path."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ensemble as ens          # noqa: E402  (test infrastructure: the checker)
from mopo_b200 import _lib                  # noqa: E402
from mopo_b200.models.constructor import construct_model   # noqa: E402


def make(p, prec):
    m = construct_model(obs_dim=p['obs_dim'], act_dim=p['act_dim'], hidden_dim=p['hidden_dim'], num_networks=p['num_networks'],
                        separate_mean_var=p['separate_mean_var'], name='t', device=0, precision=prec)
    m.set_params(p['weights'], p['biases'], p.get('var_weights'), p.get('var_biases'), p['mu'], p['sigma'],
                 p['max_logvar'], p['min_logvar'])
    m._model_inds = list(range(5))
    return m


def main():
    cases = [('halfcheetah', 300, True), ('hopper', 777, True), ('halfcheetah', 1000, False), ('humanoid', 300, True),
             ('ant', 515, True), ('halfcheetah', 20000, True)]
    names = {_lib.PREC_FP32: 'fp32(tc3)', _lib.PREC_FP32_REF: 'fp32_ref(ffma)', _lib.PREC_BF16: 'bf16'}
    for domain, B, sep in cases:
        p, obs, act = ens.synthetic_case(domain, B, seed=21, separate_mean_var=sep)
        x = np.concatenate([obs, act], -1)
        m64, v64 = ens.forward(p, x, np.float64)
        for prec in (_lib.PREC_FP32_REF, _lib.PREC_FP32, _lib.PREC_BF16):
            try:
                m = make(p, prec)
                mean, var = m.predict(x, factored=True)
            except NotImplementedError as e:
                print(json.dumps({'case': [domain, B, sep], 'prec': names[prec], 'unsupported': str(e)[:60]}))
                continue
            em = np.abs(mean - m64)
            bad = int((em > 2e-5 + 1e-3 * np.abs(m64)).sum())
            print(json.dumps({'case': [domain, B, sep], 'prec': names[prec], 'mean_max_abs_err': float(em.max()),
                              'mean_scale': float(np.abs(m64).max()), 'var_max_rel_err': float((np.abs(var - v64) / v64).max()),
                              'elements_outside_rtol1e-3': bad, 'nan': bool(np.isnan(mean).any())}))
    # time: the contraction alone, HalfCheetah B = 50 000
    p, obs, act = ens.synthetic_case('halfcheetah', 50000, seed=3)
    x = torch.as_tensor(np.concatenate([obs, act], -1)).cuda()
    lib = _lib.load()
    for prec in (_lib.PREC_FP32, _lib.PREC_BF16, _lib.PREC_FP32_REF):
        m = make(p, prec)
        m._ensure_handle()
        mean = torch.empty((7, 50000, 18), device='cuda'); lv = torch.empty_like(mean)
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            _lib.check(lib.mopo_ens_forward(m._handle, x.data_ptr(), 50000, mean.data_ptr(), lv.data_ptr(), prec, st))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10 if prec != _lib.PREC_FP32_REF else 2
        e0.record()
        for _ in range(n):
            _lib.check(lib.mopo_ens_forward(m._handle, x.data_ptr(), 50000, mean.data_ptr(), lv.data_ptr(), prec, st))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(json.dumps({'forward_B50000': names[prec], 'ms': ms, 'tflops_algorithmic': 50000 * 1845200 / ms / 1e9}))


if __name__ == '__main__':
    t0 = time.time()
    main()
    print('done in %.1f s' % (time.time() - t0))
