"""Edge cases of the path through the public shims: empty batches, a rollout whose rows all terminate at once, call-order
and argument errors mapped to the reference's exception types."""
import numpy as np
import pytest

from oracle import ensemble as ens

pytestmark = pytest.mark.gpu


def _model(p, prec=None, elites=True):
    from mopo_b200 import _lib
    from mopo_b200.models.constructor import construct_model
    m = construct_model(obs_dim=p['obs_dim'], act_dim=p['act_dim'], separate_mean_var=True, name='t', device=0,
                        precision=_lib.PREC_FP32 if prec is None else prec)
    m.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    if elites:
        m._model_inds = [0, 1, 2, 3, 4]
    return m


@pytest.mark.parametrize('precision_name', ['fp32', 'bf16'])
def test_empty_batch_is_a_no_op(precision_name):
    from mopo_b200 import _lib
    from mopo_b200.models.fake_env import FakeEnv
    import mopo_b200.static as static
    prec = {'bf16': _lib.PREC_BF16, 'fp32': _lib.PREC_FP32}[precision_name]
    p, obs, act = ens.synthetic_case('hopper', 4, seed=1)
    m = _model(p, prec)
    mean, var = m.predict(np.zeros((0, 14), np.float32), factored=True)
    assert mean.shape == (7, 0, 12) and var.shape == (7, 0, 12)
    env = FakeEnv(m, static['hopper'], penalty_coeff=1.0, penalty_learned_var=True)
    nobs, rew, term, info = env.step(obs[:0], act[:0])
    assert nobs.shape == (0, 11) and rew.shape == (0, 1) and term.shape == (0, 1) and info['penalty'].shape == (0, 1)


def test_rollout_stops_when_every_row_terminates():
    """mopo.py:753-756: nothing survives step 0 -> the loop breaks; one step's rows are in the pool."""
    import torch
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    B, h = 300, 4
    p, obs, _ = ens.synthetic_case('hopper', B, seed=2)
    p['biases'][-1][:, :, 1] = -5.0                      # predicted height delta -5: every next state is below 0.7 (hopper.py:13)
    p['max_logvar'][:] = -9.0
    m = _model(p)
    env = FakeEnv(m, static['hopper'], penalty_coeff=1.0, penalty_learned_var=True)
    ro = ModelRollout(env, SimpleReplayPool(11, 3, 10), rollout_batch_size=B, rollout_schedule=(0, 1, h, h))
    ro._set_rollout_length()
    ro._model_pool = SimpleReplayPool(11, 3, B * h)
    rng = np.random.RandomState(0)
    acts = [np.tanh(rng.standard_normal((B, 3))).astype(np.float32) for _ in range(h)]
    stats = ro._rollout_model(B, start_obs=obs, actions=acts, quiet=True)
    torch.cuda.synchronize()
    assert ro._last_steps_added == [B] and stats['mean_rollout_length'] == 1.0
    assert ro._model_pool.size == B and bool(ro._model_pool.fields['terminals'][:B].all())


def test_errors_follow_the_reference_exception_types():
    from mopo_b200 import _lib
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    p, obs, act = ens.synthetic_case('halfcheetah', 8, seed=3)
    m = _model(p, elites=False)
    with pytest.raises(ValueError):                       # fc.py:103 "Invalid input dimension."
        m.predict(np.zeros((4, 22), np.float32))
    env = FakeEnv(m, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True)
    with pytest.raises(RuntimeError):                     # _model_inds only exists after training (bnn.py:337-340)
        env.step(obs, act)
    with pytest.raises(AttributeError):
        m._model_inds
    with pytest.raises(ValueError):
        m._model_inds = [0, 1, 7]                         # elite index outside the ensemble
    m._model_inds = [0, 1, 2, 3, 4]
    with pytest.raises(ValueError):
        env.step(obs, act[:, :5])
    with pytest.raises(TypeError):                        # host static fns cannot be compiled into the step kernel
        FakeEnv(m, object())
    ro = ModelRollout(env, SimpleReplayPool(17, 6, 10), rollout_batch_size=8, rollout_schedule=(0, 1, 2, 2))
    ro._set_rollout_length()
    ro._model_pool = SimpleReplayPool(11, 3, 64)          # wrong dims for this model
    with pytest.raises(ValueError):
        ro._rollout_model(8, start_obs=obs, actions=[act, act], quiet=True)
    with pytest.raises(NotImplementedError):              # tensor-core path: hidden_dim 200 x 4 only
        from mopo_b200.models.constructor import construct_model
        small = construct_model(obs_dim=17, act_dim=6, hidden_dim=64, separate_mean_var=True, device=0, precision=_lib.PREC_BF16)
        small._model_inds = [0, 1, 2, 3, 4]
        FakeEnv(small, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True).step(obs, act)


def test_identity_model_type_bypasses_the_model():
    """model_type == 'identity' (mopo.py:740-744): next_obs = obs, zero rewards, terminals = identity_terminal; the pool
    appends, steps_added and the early break are the reference loop's."""
    from mopo_b200.algorithms.mopo import ModelRollout
    from mopo_b200.replay_pools import SimpleReplayPool
    rng = np.random.RandomState(0)
    obs = rng.standard_normal((50, 17)).astype(np.float32)
    acts = [np.tanh(rng.standard_normal((50, 6))).astype(np.float32) for _ in range(3)]
    for ident_term, want_steps in ((0, [50, 50, 50]), (1, [50])):
        ro = ModelRollout(None, SimpleReplayPool(17, 6, 10), rollout_batch_size=50, rollout_schedule=(0, 1, 3, 3),
                          model_type='identity', identity_terminal=ident_term)
        ro._set_rollout_length()
        ro._model_pool = SimpleReplayPool(17, 6, 200)
        stats = ro._rollout_model(50, start_obs=obs, actions=acts, quiet=True)
        assert ro._last_steps_added == want_steps and stats['mean_rollout_length'] == len(want_steps)
        f = {k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()}
        n = 50 * len(want_steps)
        assert ro._model_pool.size == n
        assert np.array_equal(f['next_observations'][:n], f['observations'][:n]) and np.array_equal(f['observations'][:50], obs)
        assert not f['rewards'][:n].any() and f['terminals'][:n].all() == bool(ident_term)
        assert np.array_equal(f['actions'][:50], acts[0])


def test_host_rollouts_back_to_back_with_changing_sizes():
    """mopo_rollout_host on the streamed-weights path keeps per-handle upload state (staging buffers, flag words, a flag
    epoch that wraps after 255 calls): 300 consecutive host rollouts of changing size, nothing synchronised in between,
    and the checked ones equal the device-input route bit for bit."""
    import torch
    from mopo_b200 import _lib
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    h = 2
    p, obs_all, _ = ens.synthetic_case('halfcheetah', 2000, seed=12)
    m = _model(p)
    rng = np.random.RandomState(4)
    acts_all = [np.tanh(rng.standard_normal((2000, 6))).astype(np.float32) for _ in range(h)]

    def runner():
        env = FakeEnv(m, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
        ro = ModelRollout(env, SimpleReplayPool(17, 6, 10), rollout_batch_size=2000, rollout_schedule=(0, 1, h, h), seed=3)
        ro._set_rollout_length()
        return ro

    host, dev = runner(), runner()
    sizes = [257, 1, 2000, 300, 513, 64]
    for it in range(300):
        B = sizes[it % len(sizes)]
        check = it in (0, 1, 2, 254, 255, 256, 299)
        host._model_pool = SimpleReplayPool(17, 6, B * h)
        o = torch.as_tensor(obs_all[:B]).pin_memory()
        a = [torch.as_tensor(x[:B]).pin_memory() for x in acts_all]
        rid = host._rollout_count                                   # (the rollout id feeds the Philox step id)
        host._rollout_model(B, start_obs=o, actions=a, quiet=True)
        if not check:
            continue
        torch.cuda.synchronize()
        dev._model_pool = SimpleReplayPool(17, 6, B * h)
        dev._rollout_count = rid
        dev._rollout_model(B, start_obs=torch.as_tensor(obs_all[:B]).cuda(), actions=[torch.as_tensor(x[:B]).cuda() for x in acts_all], quiet=True)
        torch.cuda.synchronize()
        for k, v in host._model_pool.fields.items():
            assert np.array_equal(v.cpu().numpy(), dev._model_pool.fields[k].cpu().numpy()), (it, B, k)
