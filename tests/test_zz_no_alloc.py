"""mopo_ens_predict_noalloc (no counterpart in the reference): the FFMA predict with
its scratch in a static device array instead of cudaMalloc'd buffers.  Collected last on purpose.
Called through the C ABI with caller-owned device buffers, compared with the existing entry point mopo_ens_predict on two calls
in a row (the array is shared and keeps the previous call's contents), plus the size limit and the no-allocation claim."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import ensemble as ens

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-3, 2e-5     # the fp32 path's tolerance in tests/test_gpu_parity.py


@pytest.fixture(scope='module', autouse=True)
def _built():
    import __graft_entry__ as ge
    ge.build()


def _model(p):
    from mopo_b200.models.constructor import construct_model
    from mopo_b200 import _lib
    m = construct_model(obs_dim=p['obs_dim'], act_dim=p['act_dim'], hidden_dim=p['hidden_dim'], num_networks=p['num_networks'],
                        separate_mean_var=p['separate_mean_var'], name='noalloc', device=0, precision=_lib.PREC_FP32)
    m.set_params(p['weights'], p['biases'], p.get('var_weights'), p.get('var_biases'), p['mu'], p['sigma'],
                 p['max_logvar'], p['min_logvar'])
    return m


def _predict(lib, fn_name, handle, x, E, D1, precision=None):
    from mopo_b200 import _lib
    B = x.shape[0]
    mean = torch.full((E, B, D1), float('nan'), device='cuda')
    var = torch.full((E, B, D1), float('nan'), device='cuda')
    stream = _lib.stream_ptr(0)
    fn = getattr(lib, fn_name)
    if precision is None:
        rc = fn(handle, x.data_ptr(), B, mean.data_ptr(), var.data_ptr(), stream)
    else:
        rc = fn(handle, x.data_ptr(), B, mean.data_ptr(), var.data_ptr(), precision, stream)
    _lib.check(rc)
    torch.cuda.synchronize()
    return mean.cpu().numpy(), var.cpu().numpy()


def test_predict_noalloc_matches_predict_on_two_calls_in_a_row():
    from mopo_b200 import _lib
    lib = _lib.load()
    p, obs, act = ens.synthetic_case('halfcheetah', 1100, seed=77)
    m = _model(p)
    E, D1 = p['num_networks'], p['obs_dim'] + 1
    x_all = torch.as_tensor(np.concatenate([obs, act], -1)).cuda().contiguous()
    # second call smaller and on other rows: whatever the first call left in the shared array must not show
    for rows in (slice(0, 1100), slice(300, 557)):
        x = x_all[rows].contiguous()
        mean, var = _predict(lib, 'mopo_ens_predict_noalloc', m._handle, x, E, D1)
        assert np.isfinite(mean).all() and np.isfinite(var).all()
        ref_mean, ref_var = _predict(lib, 'mopo_ens_predict', m._handle, x, E, D1, _lib.PREC_FP32_REF)     # same kernels, own buffers
        assert np.array_equal(mean, ref_mean) and np.array_equal(var, ref_var)
        tc_mean, tc_var = _predict(lib, 'mopo_ens_predict', m._handle, x, E, D1, _lib.PREC_FP32)            # the product's fp32 path
        np.testing.assert_allclose(mean, tc_mean, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(var, tc_var, rtol=RTOL, atol=1e-12)
        m64, v64 = ens.forward(p, x.cpu().numpy(), np.float64)
        np.testing.assert_allclose(mean, m64, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(var, v64, rtol=RTOL)


def test_predict_noalloc_allocates_nothing_and_rejects_oversize():
    from mopo_b200 import _lib
    lib = _lib.load()
    p, obs, act = ens.synthetic_case('hopper', 512, seed=78)
    m = _model(p)
    E, D1 = p['num_networks'], p['obs_dim'] + 1
    x = torch.as_tensor(np.concatenate([obs, act], -1)).cuda().contiguous()
    mean = torch.empty((E, 512, D1), device='cuda')
    var = torch.empty((E, 512, D1), device='cuda')
    stream = _lib.stream_ptr(0)
    _lib.check(lib.mopo_ens_predict_noalloc(m._handle, x.data_ptr(), 512, mean.data_ptr(), var.data_ptr(), stream))   # (set-up happens here)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _lib.check(lib.mopo_ens_predict_noalloc(m._handle, x.data_ptr(), 512, mean.data_ptr(), var.data_ptr(), stream))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0                                   # the driver handed out no device memory during the calls
    # beyond the static array: a clear error, nothing launched (the pointers are never touched)
    too_many = 200000
    rc = lib.mopo_ens_predict_noalloc(m._handle, x.data_ptr(), too_many, mean.data_ptr(), var.data_ptr(), stream)
    assert rc == _lib.EINVAL
    msg = lib.mopo_last_error().decode()
    assert 'MOPO_NOALLOC_SCRATCH_BYTES' in msg and str(too_many) in msg
    with pytest.raises(ValueError):
        _lib.check(rc)
    # and the call after a rejected one still works
    _lib.check(lib.mopo_ens_predict_noalloc(m._handle, x.data_ptr(), 512, mean.data_ptr(), var.data_ptr(), stream))
    torch.cuda.synchronize()
    m64, _ = ens.forward(p, x.cpu().numpy(), np.float64)
    np.testing.assert_allclose(mean.cpu().numpy(), m64, rtol=RTOL, atol=ATOL)
