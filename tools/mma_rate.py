#!/usr/bin/env python
"""tcgen05.mma rate probe sweep (mopo_debug_mma_rate). usage: python tools/mma_rate.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from mopo_b200 import _lib
lib = _lib.load_probes()
st = torch.cuda.current_stream().cuda_stream
out = torch.zeros(2, dtype=torch.int64, device='cuda')
print('cg  N   nacc a_tmem ksteps | cyc/kstep(total) cyc/kstep(issue) | ideal N/2')
for cg in (2,):
    for a_tmem in (1, 0):
        for N, nacc in ((256, 1), (208, 1), (112, 1), (96, 1), (48, 1), (112, 2), (96, 4), (48, 4)):
            ks = 13 * nacc * (8 // nacc)
            for _ in range(2):
                _lib.check(lib.mopo_debug_mma_rate(cg, N, nacc, a_tmem, ks, out.data_ptr(), st))
            torch.cuda.synchronize()
            t = out.cpu().tolist()
            print('%d  %3d  %d    %d     %3d   | %8.1f        %8.1f        | %5.1f' % (cg, N, nacc, a_tmem, ks, t[0] / ks, t[1] / ks, N / 2))
