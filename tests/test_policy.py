"""The SAC actor inside the rollout: oracle sanity on CPU; device parity of
mopo_policy_act (fp32 FFMA path rtol 1e-3, tcgen05 bf16 path reported separately) and of the policy-driven
device-resident rollout against the literal get_action_meta / FakeEnv.step loop."""
import numpy as np
import pytest

from oracle import ensemble as ens
from oracle import philox
from oracle import policy as opol


def test_oracle_policy_shapes_and_squash():
    W, b = opol.init_params(17, 6, seed=0)
    obs = np.random.RandomState(1).standard_normal((32, 17))
    noise = np.random.RandomState(2).standard_normal((32, 6))
    act, mu, std = opol.forward(W, b, obs, noise)
    assert act.shape == (32, 6) and np.all(np.abs(act) < 1)
    assert np.all(std >= np.exp(-20)) and np.all(std <= np.exp(2) * (1 + 1e-12))
    det, mu2, _ = opol.forward(W, b, obs, deterministic=True)
    np.testing.assert_allclose(det, np.tanh(mu2))
    # zero weights: action = tanh(b_mu + noise * exp(clip(b_log_std)))  (closed form)
    Wz = [np.zeros_like(w) for w in W]
    bz = [np.zeros_like(x) for x in b]
    bz[2][:] = 0.3
    bz[3][:] = 5.0                                       # clipped to LOG_STD_MAX = 2
    act, _, std = opol.forward(Wz, bz, obs, noise)
    np.testing.assert_allclose(std, np.exp(2.0))
    np.testing.assert_allclose(act, np.tanh(0.3 + noise * np.exp(2.0)), rtol=1e-5, atol=1e-6)


def test_oracle_policy_uniform_stream_is_distinct():
    u = philox.uniforms_pm1(np.arange(1000), 3, 77, 6)
    assert u.shape == (1000, 6) and np.all(np.abs(u) < 1) and abs(u.mean()) < 0.05
    z_pol = philox.normals(np.arange(64), 3, 77, 6, stream=philox.STREAM_POLICY)
    z_env = philox.normals(np.arange(64), 3, 77, 6)
    assert not np.allclose(z_pol, z_env)


def _policy(obs_dim, act_dim, precision, seed=0, scale=1.0):
    from mopo_b200.policies import GaussianPolicy
    W, b = opol.init_params(obs_dim, act_dim, seed=seed)
    W = [w * scale for w in W]
    pol = GaussianPolicy(obs_dim, act_dim, device=0, precision=precision, seed=1234)
    pol.set_params(W, b)
    return pol, W, b


@pytest.mark.gpu
@pytest.mark.parametrize('dims', [(17, 6), (11, 3), (111, 8)])
@pytest.mark.parametrize('B', [1, 300, 4097])
def test_policy_fp32_matches_oracle(dims, B):
    from mopo_b200 import _lib
    od, ad = dims
    pol, W, b = _policy(od, ad, _lib.PREC_FP32)
    rng = np.random.RandomState(B)
    obs = rng.standard_normal((B, od)).astype(np.float32)
    noise = rng.standard_normal((B, ad)).astype(np.float32)
    act, _, (mu, std) = pol.get_action_meta(obs, None, noise=noise, return_stats=True)
    want, wmu, wstd = opol.forward(W, b, obs, noise)
    np.testing.assert_allclose(mu, wmu, rtol=1e-3, atol=2e-5)
    np.testing.assert_allclose(std, wstd, rtol=1e-3)
    np.testing.assert_allclose(act, want, rtol=1e-3, atol=2e-5)
    det, _ = pol.get_action_meta(obs, None, deterministic=True)
    np.testing.assert_allclose(det, np.tanh(wmu), rtol=1e-3, atol=2e-5)
    one, _ = pol.get_action_meta(obs[0], None, deterministic=True)          # 1-D in -> 1-D out
    assert one.shape == (ad,)


@pytest.mark.gpu
@pytest.mark.parametrize('dims', [(17, 6), (11, 3)])
@pytest.mark.parametrize('B', [100, 128, 5000, 50000])
def test_policy_bf16_tensor_core_close_to_oracle(dims, B):
    """tcgen05 path (bf16 operands, fp32 accumulation): reported separately from the rtol-1e-3 path."""
    from mopo_b200 import _lib
    od, ad = dims
    pol, W, b = _policy(od, ad, _lib.PREC_BF16)
    rng = np.random.RandomState(B + 1)
    obs = rng.standard_normal((B, od)).astype(np.float32)
    noise = rng.standard_normal((B, ad)).astype(np.float32)
    act, _, (mu, std) = pol.get_action_meta(obs, None, noise=noise, return_stats=True)
    want, wmu, wstd = opol.forward(W, b, obs, noise)
    scale = np.abs(wmu).max()
    assert np.abs(mu - wmu).max() <= 0.02 * scale + 0.02, np.abs(mu - wmu).max()
    np.testing.assert_allclose(std, wstd, rtol=0.08)
    assert np.abs(act - want).max() <= 0.06


@pytest.mark.gpu
@pytest.mark.parametrize('precision_name', ['fp32', 'bf16'])
def test_policy_philox_noise_and_uniform_actions(precision_name):
    from mopo_b200 import _lib
    prec = {'fp32': _lib.PREC_FP32, 'bf16': _lib.PREC_BF16}[precision_name]
    pol, W, b = _policy(17, 6, prec)
    B = 2000
    obs = np.random.RandomState(9).standard_normal((B, 17)).astype(np.float32)
    act, _, (mu, std) = pol.get_action_meta(obs, None, return_stats=True, step_id=5)
    z = philox.normals(np.arange(B), 5, 1234, 6, stream=philox.STREAM_POLICY)
    np.testing.assert_allclose(act, np.tanh(mu + z * std), atol=2e-3)


@pytest.mark.gpu
@pytest.mark.parametrize('precision_name', ['fp32', 'bf16'])
@pytest.mark.parametrize('domain', ['halfcheetah', 'hopper'])
def test_rollout_with_device_policy_equals_literal_loop(precision_name, domain):
    """mopo.py:730-758 with act = policy(obs) on the device: the fused call must equal the step-by-step loop
    (same kernels, same Philox counters for the policy noise; env noise / picks injected)."""
    import torch
    from mopo_b200 import _lib
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    prec = {'fp32': _lib.PREC_FP32, 'bf16': _lib.PREC_BF16}[precision_name]
    od, ad = (17, 6) if domain == 'halfcheetah' else (11, 3)
    B, h = 700, 3
    p, obs, _ = ens.synthetic_case(domain, B, seed=21)
    if domain == 'hopper':
        p['max_logvar'] = np.full_like(p['max_logvar'], -6.0)       # low noise: most rows survive a few steps
    m = construct_model(obs_dim=od, act_dim=ad, separate_mean_var=True, name='t', device=0, precision=prec)
    m.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    m._model_inds = [0, 1, 2, 3, 4]
    pol, W, b = _policy(od, ad, prec, scale=0.5)
    rng = np.random.RandomState(4)
    noise = [rng.standard_normal((B, od + 1)).astype(np.float32) for _ in range(h)]
    inds = [rng.choice(5, size=B).astype(np.int32) for _ in range(h)]
    env = FakeEnv(m, static[domain], penalty_coeff=1.0, penalty_learned_var=True, seed=1234)
    ro = ModelRollout(env, SimpleReplayPool(od, ad, 10), rollout_batch_size=B, rollout_schedule=(0, 1, h, h), seed=1234, policy=pol)
    ro._set_rollout_length()
    ro._model_pool = SimpleReplayPool(od, ad, B * h)
    rid = ro._rollout_count
    ro._rollout_model(B, start_obs=obs, noise=noise, model_inds=inds, precision=prec, quiet=True)
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()}
    steps = ro._last_steps_added
    # literal loop (origin ids: the surviving rows keep their start index as the Philox counter)
    o, origin, lo, row = obs, np.arange(B), 0, 0
    for i in range(len(steps)):
        n = len(o)
        assert steps[i] == n
        a_t = torch.empty(n, ad, device='cuda')
        o_t = torch.as_tensor(o).cuda().contiguous()
        org = torch.as_tensor(origin.astype(np.int32)).cuda()
        import ctypes as C
        args = _lib.PolicyArgs(obs=o_t.data_ptr(), B=n, deterministic=0, random_uniform=0, precision=prec, noise=None, seed=1234,
                               step_id=(rid << 16) ^ i, origin=org.data_ptr(), origin_base=0, actions=a_t.data_ptr(), mu_out=None, std_out=None)
        _lib.check(_lib.load().mopo_policy_act(pol._handle, C.byref(args), _lib.stream_ptr(torch.device('cuda', 0))))
        act = a_t.cpu().numpy()
        nobs, rew, term, info = env.step(o, act, noise=noise[i][:n], model_inds=inds[i][:n], diagnostics=False, precision=prec)
        sl = slice(row, row + n)
        np.testing.assert_array_equal(got['actions'][sl], act, err_msg='actions step %d' % i)
        np.testing.assert_array_equal(got['observations'][sl], o.astype(np.float32))
        np.testing.assert_array_equal(got['next_observations'][sl], nobs.astype(np.float32))
        np.testing.assert_array_equal(got['terminals'][sl], term)
        row += n
        keep = ~term.squeeze(-1)
        if keep.sum() == 0:
            break
        o, origin = nobs[keep].astype(np.float32), origin[keep]
    assert row == sum(steps)
