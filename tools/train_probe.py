"""Time the device optimiser loop of BNN.train (mopo_ens_train_batches) on a D4RL-shaped problem and, beside it, the
float64 oracle port on the host cores (a bounded sample).  usage: train_probe.py [obs] [act] [rows] [batch] [epochs]"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mopo_b200 import _lib                                   # noqa: E402
from oracle import ensemble as ens, train as otr             # noqa: E402
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
from test_train import _handle, _decays                      # noqa: E402


def main():
    od, ad, n_rows, bs, epochs = [int(a) for a in sys.argv[1:6]] + [17, 6, 100000, 256, 3][len(sys.argv) - 1:]
    E = 7
    p = ens.init_params(od, ad, 1, 200, E, 4, separate_mean_var=True, seed=0)
    rng = np.random.RandomState(0)
    inputs = rng.standard_normal((n_rows, od + ad)).astype(np.float32)
    targets = (0.1 * rng.standard_normal((n_rows, od + 1))).astype(np.float32)
    ens.fit_scaler(p, inputs)
    idx = rng.randint(n_rows, size=(E, n_rows)).astype(np.int32)
    lib, h = _handle(p)
    d_in, d_tg, d_idx = (torch.as_tensor(a).cuda() for a in (inputs, targets, idx))
    _lib.check(lib.mopo_ens_train_begin(h, 1e-3, _decays(p)))
    side = torch.cuda.Stream()          # not the legacy default stream: the batch step is graph-captured
    torch.cuda.synchronize()
    torch.cuda.set_stream(side)
    st = side.cuda_stream
    nb = int(np.ceil(n_rows / bs))

    def epoch():
        _lib.check(lib.mopo_ens_train_batches(h, d_in.data_ptr(), d_tg.data_ptr(), n_rows, d_idx.data_ptr(), n_rows, n_rows, bs, None, st))

    l0 = lib.mopo_launch_count()
    epoch()
    torch.cuda.synchronize()
    launches = (lib.mopo_launch_count() - l0) / nb
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(epochs):
        epoch()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (epochs * nb)
    # host baseline: the float64 NumPy port, a few batches
    v = otr.trainable(p)
    sta = otr.adam_init(v)
    k = 4
    t0 = time.perf_counter()
    otr.train_batches(p, v, sta, inputs, targets, idx[:, :k * bs], bs)
    cpu_ms = (time.perf_counter() - t0) / k * 1e3
    flop = 3 * 2 * E * bs * ((od + ad) * 200 + 3 * 200 * 200 + 200 * 2 * (od + 1))
    print(json.dumps({'workload': 'bnn_train obs%d act%d E%d batch%d rows%d' % (od, ad, E, bs, n_rows), 'ms_per_adam_step': ms,
                      'net_rows_per_s': E * bs / ms * 1e3, 'launches_per_step': launches, 'gflop_per_step': flop / 1e9,
                      'tflops': flop / ms / 1e9, 'cpu_port_ms_per_step': cpu_ms, 'cpu_cores': os.cpu_count()}))
    _lib.check(lib.mopo_ens_train_end(h, st))


if __name__ == '__main__':
    main()
