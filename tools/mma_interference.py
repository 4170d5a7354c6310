#!/usr/bin/env python
"""tcgen05.mma vs epilogue interference probe (mopo_debug_mma_interference): does the tensor pipe slow down while other
warps of the SM run MUFU chains / tcgen05.ld+st / the whole fp32-accurate epilogue body, and vice versa?
usage (GPU box): python tools/mma_interference.py > gpurun_out/mma_interference.txt"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mopo_b200 import _lib

lib = _lib.load_probes()
out = torch.zeros(8, dtype=torch.int64, device='cuda')
st = torch.cuda.current_stream().cuda_stream
KS = 13 * 400
names = {0: 'idle', 1: 'ex2+rcp (registers)', 2: 'tcgen05.ld/st x16', 3: 'full group body', 4: 'kernel-like rounds'}
print('N  imode(interferer)        warps icol | MMA cyc/kstep  (ideal) | interferer groups/1000clk/SMSP  (alone)')
for N in (112, 96):
    for imode in (0, 1, 2, 3, 4):
        for nw in ((0,) if imode == 0 else (4, 8, 12)):
            for icol in ((208,) if imode < 3 or nw < 12 else (N, 128, 208, 256, 352) if imode == 3 else (208,)):
                res = []
                for with_mma in (1, 0):
                    for _ in range(2):
                        out.zero_()
                        _lib.check(lib.mopo_debug_mma_interference(N, 1, KS, imode, nw, with_mma, icol, out.data_ptr(), st), lib)
                        torch.cuda.synchronize()
                    o = out.cpu().tolist()
                    res.append((o[0] / KS, o[2] / max(o[0], 1) * 1000 / 4))
                print('%3d  %-24s %2d  %3d  | %6.1f        (%5.1f) | %6.2f   (%6.2f)' % (N, names[imode], nw, icol, res[0][0], N / 2, res[0][1], res[1][1]))
