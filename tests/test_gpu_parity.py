"""GPU parity tests: the CUDA path (through the C ABI) vs the oracle and the golden fixtures that
were generated from the reference's own FakeEnv / static fns / FlexibleReplayPool.

Bars (BASELINE.json): elite indices, termination masks (on identical next_obs), pool indices /
pointer / size bit-exact; next_obs / reward / penalty rtol 1e-3 on the fp32 path.
"""
import numpy as np
import pytest
import torch
from conftest import golden_files

from oracle import ensemble as ens
from oracle.fake_env import FakeEnv as OracleEnv
from oracle.static_fns import StaticFns as OracleStatic
from oracle.pool import SimpleReplayPool as OraclePool
from oracle.rollout import rollout_model

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-3, 2e-5     # fp32 path tolerance (north_star: rtol 1e-3); atol covers values near 0
SHAPES = {'halfcheetah': (17, 6), 'hopper': (11, 3), 'walker2d': (17, 6), 'ant': (27, 8), 'humanoid': (45, 17)}


@pytest.fixture(scope='module', autouse=True)
def _built():
    import __graft_entry__ as ge
    ge.build()


def make_model(p, elites=None, precision=None):
    from mopo_b200.models.constructor import construct_model
    from mopo_b200 import _lib
    obs_dim, act_dim = p['obs_dim'], p['act_dim']
    m = construct_model(obs_dim=obs_dim, act_dim=act_dim, hidden_dim=p['hidden_dim'], num_networks=p['num_networks'],
                        separate_mean_var=p['separate_mean_var'], name='t', device=0,
                        precision=_lib.PREC_FP32 if precision is None else precision)
    m.set_params(p['weights'], p['biases'], p.get('var_weights'), p.get('var_biases'), p['mu'], p['sigma'],
                 p['max_logvar'], p['min_logvar'])
    m._model_inds = list(range(5)) if elites is None else list(elites)
    return m


@pytest.mark.parametrize('domain,sep', [('halfcheetah', True), ('hopper', True), ('halfcheetah', False), ('humanoid', True)])
def test_predict_matches_oracle(domain, sep):
    p, obs, act = ens.synthetic_case(domain, 300, seed=21, separate_mean_var=sep)
    x = np.concatenate([obs, act], -1)
    m = make_model(p)
    mean, var = m.predict(x, factored=True)
    m64, v64 = ens.forward(p, x, np.float64)
    assert mean.shape == m64.shape and mean.dtype == np.float32
    np.testing.assert_allclose(mean, m64, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(var, v64, rtol=RTOL)
    # non-factored aggregation (bnn.py:261-263)
    mm, vv = m.predict(x)
    np.testing.assert_allclose(mm, m64.mean(0), rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(vv, v64.mean(0) + np.square(m64 - m64.mean(0)).mean(0), rtol=RTOL)


@pytest.mark.parametrize('path', golden_files('fwd_'))
def test_predict_matches_reference_bnn_fixture(path):
    """The device forward against fixtures produced by the reference's own BNN.predict (tests/golden/make_golden.py runs
    constructor.py / bnn.py / fc.py / utils.py unmodified over a NumPy TF stand-in): 2-D factored and aggregated."""
    g = np.load(path)
    p, obs, act = ens.synthetic_case(str(g['domain']), int(g['B']), seed=int(g['seed']), separate_mean_var=bool(g['separate']))
    x = np.concatenate([obs, act], -1)
    m = make_model(p)
    mean, var = m.predict(x, factored=True)
    np.testing.assert_allclose(mean, g['mean_fac'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(var, g['var_fac'], rtol=RTOL)
    mm, vv = m.predict(x)
    np.testing.assert_allclose(mm, g['mean'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(vv, g['var'], rtol=RTOL)
    # 3-D inputs: net e sees x3[e] (bnn.py:541-545, fc.py:100-101) -- on the fp32-accurate tensor-core path and on the FFMA checker
    from mopo_b200 import _lib
    for prec in (_lib.PREC_FP32, _lib.PREC_FP32_REF):
        m3, v3 = m.predict(g['x3'], factored=True, precision=prec)
        np.testing.assert_allclose(m3, g['mean3'], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(v3, g['var3'], rtol=RTOL)
    with pytest.raises(ValueError):
        m.predict(g['x3'][:3])                          # leading dimension != ensemble size (fc.py:100-103)


@pytest.mark.parametrize('E,B', [(1, 300), (3, 1000), (12, 257)])
def test_predict_other_ensemble_sizes_on_the_tensor_core_path(E, B):
    """(net, tile) units are range-partitioned over the CTA pairs for any ensemble size: 1 net (fewer units than pairs), 3, 12."""
    from mopo_b200.models.constructor import construct_model
    p = ens.init_params(17, 6, num_networks=E, seed=E, bias_std=0.1)
    ens.scale_weights(p, 3.0)
    rng = np.random.RandomState(E)
    x = rng.standard_normal((B, 23)).astype(np.float32)
    ens.fit_scaler(p, x)
    m = construct_model(obs_dim=17, act_dim=6, num_networks=E, num_elites=1, separate_mean_var=True, name='e', device=0)
    m.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    mean, var = m.predict(x, factored=True)
    m64, v64 = ens.forward(p, x, np.float64)
    np.testing.assert_allclose(mean, m64, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(var, v64, rtol=RTOL)


def test_linear_model_predict_and_step():
    """construct_model(model_type='linear') (constructor.py:25-27): one FC whose output finalize() doubles into [mean | logvar]
    (bnn.py:185); predict and FakeEnv.step against the oracle."""
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    import mopo_b200.static as static
    od, ad, B = 11, 3, 257
    p = ens.init_params(od, ad, n_hidden=0, separate_mean_var=False, seed=4, bias_std=0.1)
    rng = np.random.RandomState(5)
    obs = (rng.standard_normal((B, od)) * 0.05).astype(np.float32); obs[:, 0] += 1.25
    act = np.tanh(rng.standard_normal((B, ad))).astype(np.float32)
    ens.fit_scaler(p, np.concatenate([obs, act], -1))
    m = construct_model(obs_dim=od, act_dim=ad, model_type='linear', separate_mean_var=False, name='lin', device=0)
    assert len(m.layers) == 1 and m.layers[0].get_output_dim() == 2 * (od + 1)
    m.set_params(p['weights'], p['biases'], None, None, p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    m._model_inds = [0, 1, 2, 3, 4]
    x = np.concatenate([obs, act], -1)
    mean, var = m.predict(x, factored=True)
    m64, v64 = ens.forward(p, x, np.float64)
    np.testing.assert_allclose(mean, m64, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(var, v64, rtol=RTOL)
    noise = rng.standard_normal((7, B, od + 1))
    inds = rng.choice(5, size=B)
    o_env = OracleEnv(ens.OracleModel(p), OracleStatic('hopper'), penalty_coeff=1.0, penalty_learned_var=True)
    r_nobs, r_rew, r_term, r_info = o_env.step(obs, act, noise=noise, model_inds=inds)
    nobs, rew, term, info = FakeEnv(m, static['hopper'], penalty_coeff=1.0, penalty_learned_var=True).step(obs, act, noise=noise, model_inds=inds)
    np.testing.assert_allclose(nobs, r_nobs, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(rew, r_rew, rtol=RTOL, atol=ATOL)
    assert np.array_equal(term, r_term)
    with pytest.raises(RuntimeError):                   # bnn.py:175-176: a separate variance head needs var layers, 'linear' adds none
        construct_model(obs_dim=od, act_dim=ad, model_type='linear', separate_mean_var=True, name='lin2', device=0)


def test_kat_zero_weights_device():
    p = ens.init_params(11, 3, seed=0)
    p['weights'] = [w * 0 for w in p['weights']]
    p['var_weights'] = p['var_weights'] * 0
    m = make_model(p)
    x = np.random.RandomState(0).randn(70, 14).astype(np.float32)
    mean, var = m.predict(x, factored=True)
    assert np.all(mean == 0)
    np.testing.assert_allclose(var, 0.6225047, rtol=2e-6)
    from mopo_b200.models.fake_env import FakeEnv
    import mopo_b200.static as static
    env = FakeEnv(m, static['hopper'], penalty_coeff=1.0, penalty_learned_var=True)
    _, _, _, info = env.step(x[:, :11], x[:, 11:], noise=np.zeros((70, 12)), model_inds=np.zeros(70, int))
    np.testing.assert_allclose(info['penalty'], 0.7889897 * np.sqrt(12), rtol=2e-6)


@pytest.mark.parametrize('path', golden_files('step_'))
def test_step_matches_reference_golden(path):
    from mopo_b200.models.fake_env import FakeEnv
    import mopo_b200.static as static
    g = np.load(path)
    domain = str(g['domain'])
    p, obs, act = ens.synthetic_case(domain, int(g['B']), seed=int(g['seed']), separate_mean_var=bool(g['separate']))
    m = make_model(p, elites=g['elites'].tolist())
    env = FakeEnv(m, static[domain], penalty_coeff=float(g['coeff']), penalty_learned_var=bool(g['learned_var']))
    det = bool(g['deterministic'])
    nobs, rew, term, info = env.step(obs, act, deterministic=det, noise=g['noise'], model_inds=g['model_inds'])
    assert nobs.shape == g['next_obs'].shape and rew.shape == (len(obs), 1) and term.dtype == bool
    np.testing.assert_allclose(nobs, g['next_obs'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(rew, g['rew'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(info['unpenalized_rewards'], g['unpenalized'], rtol=RTOL, atol=ATOL)
    if g['penalty'].size:
        np.testing.assert_allclose(info['penalty'], g['penalty'], rtol=RTOL, atol=ATOL)
    else:
        assert info['penalty'] is None
    if not det:
        assert np.array_equal(info['model_inds'], g['model_inds'])          # elite gather bit-exact
    np.testing.assert_allclose(info['mean'], g['info_mean'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(info['std'], g['info_std'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(info['dev'], g['info_dev'], rtol=RTOL, atol=ATOL)
    fin = np.isfinite(g['info_log_prob'])
    np.testing.assert_allclose(info['log_prob'][fin], g['info_log_prob'][fin], rtol=5e-3, atol=5e-3)
    # termination: bit-exact on identical next_obs ...
    t_same = static[domain].termination_fn(obs, act, g['next_obs'].astype(np.float32))
    t_ref32 = OracleStatic(domain).termination_fn(obs, act, g['next_obs'].astype(np.float32).astype(np.float64))
    assert np.array_equal(t_same, t_ref32)
    # ... and end to end (fixtures have no row within rounding distance of a threshold)
    assert np.array_equal(term, g['term'])


@pytest.mark.parametrize('path', golden_files('term_'))
def test_termination_bit_exact(path):
    import mopo_b200.static as static
    g = np.load(path)
    nobs = g['next_obs']
    z = np.zeros_like(nobs)
    term = static[str(g['domain'])].termination_fn(z, np.zeros((len(nobs), 1), np.float32), nobs)
    assert term.dtype == bool and term.shape == (len(nobs), 1)
    assert np.array_equal(term, g['term'])
    tt = static[str(g['domain'])].termination_fn(torch.as_tensor(z).cuda(), torch.zeros(len(nobs), 1).cuda(), torch.as_tensor(nobs).cuda())
    assert tt.is_cuda and np.array_equal(tt.cpu().numpy(), g['term'])


def test_single_row_step_returns_unbatched():
    from mopo_b200.models.fake_env import FakeEnv
    import mopo_b200.static as static
    p, obs, act = ens.synthetic_case('walker2d', 4, seed=2)
    env = FakeEnv(make_model(p), static['walker2d'], penalty_coeff=1.0, penalty_learned_var=True)
    nobs, rew, term, info = env.step(obs[0], act[0])                       # fake_env.py:39-44,121-127
    assert nobs.shape == (17,) and rew.shape == (1,) and term.shape == (1,) and info['mean'].shape == (19,)
    with pytest.raises(AssertionError):
        env.step(obs, act[0])
    with pytest.raises(ValueError):
        env.step(obs[:, :5], act)


def test_pool_matches_reference_golden():
    from mopo_b200.replay_pools import SimpleReplayPool
    g = np.load(golden_files('pool_')[0])
    pool = SimpleReplayPool(int(g['obs_dim']), int(g['act_dim']), int(g['max_size']))
    for k in pool.fields:
        assert not pool.fields[k].any()                                    # zero-initialised
    for i in range(int(g['n_chunks'])):
        pool.add_samples({k: g['chunk%d_%s' % (i, k)] for k in pool.fields})
        assert pool._pointer == g['ptrs'][i] and pool.size == g['sizes'][i]
    for k, v in pool.fields.items():
        ref = g['final_' + k]
        assert tuple(v.shape) == ref.shape
        assert np.array_equal(v.cpu().numpy(), ref), k
    allv = pool.return_all_samples()
    assert allv['observations'].shape[0] == pool.size
    idx = np.array([0, 5, 36, 5])
    b = pool.batch_by_indices(idx)
    assert np.array_equal(b['rewards'], g['final_rewards'][idx])
    part = SimpleReplayPool(int(g['obs_dim']), int(g['act_dim']), 37)
    part.add_samples({k: g['chunk0_%s' % k] for k in part.fields})          # 10 rows of 37
    with pytest.raises(ValueError):                                        # flexible_replay_pool.py:122-125
        part.batch_by_indices(np.array([3, 37 + 12]))
    assert part.batch_by_indices(np.array([37 + 2]))['rewards'].shape == (1, 1)
    d = pool.random_batch_device(1000, seed=3)
    ii = d['indices'].cpu().numpy()
    assert ii.min() >= 0 and ii.max() < pool.size and len(np.unique(ii)) > 20
    assert np.array_equal(d['observations'].cpu().numpy(), g['final_observations'][ii])


def _oracle_rollout(domain, p, obs, acts, noises, inds, h, max_size, coeff=1.0, learned=True):
    env = OracleEnv(ens.OracleModel(p), OracleStatic(domain), penalty_coeff=coeff, penalty_learned_var=learned)
    pool = OraclePool(obs.shape[1], acts[0].shape[1], max_size)
    # noise/inds are defined per ORIGINAL row; the oracle sees the surviving subset each step
    alive = [np.arange(len(obs))]

    def noise_fn(i, n): return noises[i][:, alive[-1]]
    def inds_fn(i, n): return inds[i][alive[-1]]

    orig_step = env.step

    def step(o, a, **kw):
        out = orig_step(o, a, **kw)
        alive.append(alive[-1][~out[2][:, 0]])
        return out
    env.step = step
    stats, steps = rollout_model(env, pool, obs, h, lambda i, o: acts[i][alive[-1]], noise_fn=noise_fn, inds_fn=inds_fn)
    return pool, stats, steps, alive


@pytest.mark.parametrize('domain,B,h,max_size', [('halfcheetah', 700, 3, 5000), ('hopper', 900, 5, 1000),
                                                  ('walker2d', 600, 4, 100000), ('hopper', 300, 3, 200)])
def test_rollout_matches_oracle(domain, B, h, max_size):
    """Device-resident h-step rollout (one C call) vs the literal reference loop: steps_added, pointer, size,
    terminals and row placement bit-exact; float fields rtol 1e-3.  Injected randomness is indexed by the
    ORIGINAL row, and each side gathers the surviving rows, so both consume identical draws."""
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    p, obs, _ = ens.synthetic_case(domain, B, seed=31)
    if domain != 'halfcheetah':
        # low-noise ensemble so that a good fraction survives several steps
        p['max_logvar'][:] = -9.0
        p['weights'][-1] = p['weights'][-1] * 0.02
        p['biases'][-1] = p['biases'][-1] * 0.0
        # spread heights / angles across the termination thresholds so a fraction dies at every step
        r0 = np.random.RandomState(77)
        obs[:, 0] = r0.uniform(0.6, 1.5, B).astype(np.float32)
        obs[:, 1] = r0.uniform(-0.25, 0.25, B).astype(np.float32)
    od, ad = SHAPES[domain]
    rng = np.random.RandomState(5)
    acts = [np.tanh(rng.standard_normal((B, ad))).astype(np.float32) for _ in range(h)]
    noises = [rng.standard_normal((7, B, od + 1)) for _ in range(h)]
    inds = [rng.choice(5, size=B) for _ in range(h)]
    o_pool, o_stats, o_steps, alive = _oracle_rollout(domain, p, obs, acts, noises, inds, h, max_size)

    m = make_model(p)
    env = FakeEnv(m, static[domain], penalty_coeff=1.0, penalty_learned_var=True)
    env_pool = SimpleReplayPool(od, ad, 10)
    ro = ModelRollout(env, env_pool, rollout_batch_size=B, rollout_schedule=(0, 1, h, h))
    ro._set_rollout_length()
    ro._model_pool = SimpleReplayPool(od, ad, max_size)
    # device side: per-step arrays in the compacted order of that step
    d_acts = [acts[i][alive[i]] if i < len(alive) else acts[i][:0] for i in range(h)]
    d_noise = [noises[i][inds[i][alive[i]], alive[i]] if i < len(alive) else np.zeros((0, od + 1)) for i in range(h)]
    d_inds = [inds[i][alive[i]] if i < len(alive) else inds[i][:0] for i in range(h)]
    # pad to B rows: the kernel reads only the first n_i rows
    pad = lambda a: np.concatenate([a, np.zeros((B - len(a),) + a.shape[1:], a.dtype)])
    stats = ro._rollout_model(B, start_obs=obs, actions=[pad(a) for a in d_acts], noise=[pad(a) for a in d_noise],
                              model_inds=[pad(a) for a in d_inds], quiet=True)
    assert ro._last_steps_added == o_steps
    assert stats['mean_rollout_length'] == o_stats['mean_rollout_length']
    pool = ro._model_pool
    assert pool._pointer == o_pool._pointer and pool.size == o_pool.size
    if domain != 'halfcheetah':
        assert len(o_steps) > 1 and o_steps[1] < o_steps[0]                 # compaction actually exercised
    got = {k: v.cpu().numpy() for k, v in pool.fields.items()}
    assert np.array_equal(got['terminals'], o_pool.fields['terminals'])
    for k in ('observations', 'actions', 'next_observations', 'rewards'):
        np.testing.assert_allclose(got[k], o_pool.fields[k], rtol=RTOL, atol=5e-5, err_msg=k)


def test_rollout_philox_is_reproducible_and_shard_invariant():
    """Device RNG: same (seed, rollout id) -> identical pool; a 2-way batch shard with origin_base produces the
    same rows as the unsharded run (Philox counter = global row id)."""
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    B, h = 512, 3
    p, obs, _ = ens.synthetic_case('halfcheetah', B, seed=41)
    m = make_model(p, elites=[1, 3, 4, 5, 6])
    rng = np.random.RandomState(1)
    acts = [np.tanh(rng.standard_normal((B, 6))).astype(np.float32) for _ in range(h)]

    def run(lo, hi):
        env = FakeEnv(m, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True)
        ro = ModelRollout(env, SimpleReplayPool(17, 6, 10), rollout_batch_size=hi - lo, rollout_schedule=(0, 1, h, h), seed=77)
        ro._set_rollout_length()
        ro._model_pool = SimpleReplayPool(17, 6, 4096)
        ro._rollout_model(hi - lo, start_obs=obs[lo:hi], actions=[a[lo:hi] for a in acts], origin_base=lo, quiet=True)
        return {k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()}, ro._model_pool.size

    full, n = run(0, B)
    again, _ = run(0, B)
    assert n == B * h
    for k in full:
        assert np.array_equal(full[k], again[k])
    a, na = run(0, B // 2)
    b, nb = run(B // 2, B)
    for i in range(h):                                   # step-major rows: shard rows interleave per step
        for k in ('next_observations', 'rewards'):
            np.testing.assert_array_equal(full[k][i * B:i * B + B // 2], a[k][i * (B // 2):(i + 1) * (B // 2)])
            np.testing.assert_array_equal(full[k][i * B + B // 2:(i + 1) * B], b[k][i * (B // 2):(i + 1) * (B // 2)])
    # noise sanity: (next_obs - mean)/std ~ N(0,1) is covered by the oracle Philox restatement test
    assert np.isfinite(full['next_observations'][:n]).all()
    # elite picks only from the elite set and roughly uniform
    env = FakeEnv(m, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True)
    _, _, _, info = env.step(obs, acts[0])
    picks = info['model_inds']
    assert set(np.unique(picks)) <= {1, 3, 4, 5, 6} and len(np.unique(picks)) == 5


def test_callback_rollout_matches_fused():
    """The reference-style per-step loop with a host policy callback gives the same pool as the fused call."""
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    B, h = 256, 3
    p, obs, _ = ens.synthetic_case('halfcheetah', B, seed=51)
    m = make_model(p)
    rng = np.random.RandomState(2)
    acts = [np.tanh(rng.standard_normal((B, 6))).astype(np.float32) for _ in range(h)]
    pools = []
    for mode in ('fused', 'callback'):
        env = FakeEnv(m, static['halfcheetah'], penalty_coeff=0.5, penalty_learned_var=False, seed=9)
        calls = []
        ro = ModelRollout(env, SimpleReplayPool(17, 6, 10), rollout_batch_size=B, rollout_schedule=(0, 1, h, h), seed=9,
                          get_action_meta=lambda o, hid: (acts[len(calls)], calls.append(1))[0:1] + (None,))
        ro._set_rollout_length(); ro._reallocate_model_pool()
        if mode == 'fused':
            ro._rollout_model(B, start_obs=obs, actions=acts, quiet=True)
        else:
            ro._rollout_model(B, start_obs=obs, quiet=True)
        pools.append({k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()})
        assert ro._model_pool.size == B * h
    # different Philox step ids between the two drivers -> compare the deterministic parts only
    assert np.array_equal(pools[0]['actions'], pools[1]['actions'])
    assert np.array_equal(pools[0]['observations'][:B], pools[1]['observations'][:B])


@pytest.mark.gpu
@pytest.mark.parametrize('precision_name,B', [('bf16', 3000), ('fp32', 3000), ('fp32', 50000), ('fp32', 300)])
def test_rollout_host_actions_overlapped_upload_equals_device_actions(precision_name, B):
    """Pinned host start states / action arrays are uploaded while the rollout runs -- bf16: behind per-step events
    (mopo_rollout_args.action_events); fp32 (streamed-weights kernel): in chunks behind flags the forward polls, units
    tile-major -- and the pool must equal the one produced from device-resident inputs, bit for bit."""
    import torch
    from mopo_b200 import _lib
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    h = 4
    p, obs, _ = ens.synthetic_case('halfcheetah', B, seed=61)
    prec = {'bf16': _lib.PREC_BF16, 'fp32': _lib.PREC_FP32}[precision_name]
    m = make_model(p, precision=prec)
    rng = np.random.RandomState(3)
    acts = [np.tanh(rng.standard_normal((B, 6))).astype(np.float32) for _ in range(h)]
    pools = []
    for mode in ('device', 'host'):
        env = FakeEnv(m, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=5)
        ro = ModelRollout(env, SimpleReplayPool(17, 6, 10), rollout_batch_size=B, rollout_schedule=(0, 1, h, h), seed=5)
        ro._set_rollout_length()
        ro._model_pool = SimpleReplayPool(17, 6, B * h)
        a = [torch.as_tensor(x).cuda() for x in acts] if mode == 'device' else [torch.as_tensor(x).pin_memory() for x in acts]
        o = torch.as_tensor(obs).cuda() if mode == 'device' else torch.as_tensor(obs).pin_memory()
        ro._rollout_model(B, start_obs=o, actions=a, precision=prec, quiet=True)
        torch.cuda.synchronize()
        assert ro._last_steps_added == [B] * h
        pools.append({k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()})
    for k in pools[0]:
        assert np.array_equal(pools[0][k], pools[1][k]), k


@pytest.mark.gpu
@pytest.mark.parametrize('domain,B', [('hopper', 5000), ('walker2d', 20000)])
def test_rollout_host_inputs_with_terminating_rows_equal_device_inputs(domain, B):
    """Terminating environments (compaction between the steps) with pinned HOST inputs: the chunked, flag-guarded uploads of
    the fp32-accurate path must give the pool of the device-input route bit for bit, rows dying at every step."""
    import torch
    from mopo_b200 import _lib
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    h = 4
    od, ad = SHAPES[domain]
    p, obs, _ = ens.synthetic_case(domain, B, seed=63)
    p['max_logvar'][:] = -9.0
    p['weights'][-1] = p['weights'][-1] * 0.02
    p['biases'][-1] = p['biases'][-1] * 0.0
    r0 = np.random.RandomState(7)
    obs[:, 0] = r0.uniform(0.75, 1.5, B).astype(np.float32)
    obs[:, 1] = r0.uniform(-0.19, 0.19, B).astype(np.float32)
    m = make_model(p)
    rng = np.random.RandomState(3)
    acts = [np.tanh(rng.standard_normal((B, ad))).astype(np.float32) for _ in range(h)]
    pools, steps = [], []
    for mode in ('device', 'host'):
        env = FakeEnv(m, static[domain], penalty_coeff=1.0, penalty_learned_var=True, seed=5)
        ro = ModelRollout(env, SimpleReplayPool(od, ad, 10), rollout_batch_size=B, rollout_schedule=(0, 1, h, h), seed=5)
        ro._set_rollout_length()
        ro._model_pool = SimpleReplayPool(od, ad, B * h)
        a = [torch.as_tensor(x).cuda() for x in acts] if mode == 'device' else [torch.as_tensor(x).pin_memory() for x in acts]
        o = torch.as_tensor(obs).cuda() if mode == 'device' else torch.as_tensor(obs).pin_memory()
        ro._rollout_model(B, start_obs=o, actions=a, quiet=True)
        torch.cuda.synchronize()
        steps.append(list(ro._last_steps_added))
        pools.append({k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()})
    assert steps[0] == steps[1] and len(steps[0]) > 1 and 0 < steps[0][1] < steps[0][0]
    for k in pools[0]:
        assert np.array_equal(pools[0][k], pools[1][k]), k


@pytest.mark.gpu
def test_offline_pool_restore_contiguous_and_d4rl_dict(tmp_path):
    """loader.py:102-126 / :23-29: the offline dataset lands in the device pool exactly as FlexibleReplayPool.add_samples
    would place it (oracle pool), incl. the f64 -> f32 cast and the bool terminals; random_batch then feeds the rollout."""
    from mopo_b200.off_policy.loader import restore_pool, restore_pool_d4rl
    from mopo_b200.replay_pools import SimpleReplayPool
    from oracle.pool import SimpleReplayPool as OraclePool
    od, ad, n = 11, 3, 1000
    rng = np.random.RandomState(0)
    data = np.concatenate([rng.standard_normal((n, od)), np.tanh(rng.standard_normal((n, ad))), rng.standard_normal((n, od)),
                           rng.standard_normal((n, 1)), (rng.uniform(size=(n, 1)) < 0.1).astype(np.float64)], axis=1)
    path = str(tmp_path / 'pool.npy')
    np.save(path, data)
    pool = SimpleReplayPool(od, ad, 1500)
    restore_pool(pool, path, max_size=1500)
    want = OraclePool(od, ad, 1500)
    want.add_samples({'observations': data[:, :od], 'actions': data[:, od:od + ad], 'next_observations': data[:, od + ad:2 * od + ad],
                      'rewards': data[:, -2:-1], 'terminals': data[:, -1:].astype(bool)})
    assert pool.size == want.size == n and pool._pointer == want._pointer
    for k, v in pool.fields.items():
        assert np.array_equal(v.cpu().numpy(), want.fields[k]), k
    # wrong width -> the reference's assertion (then the BEAR fallback fails on the .npy: an exception either way)
    np.save(path, data[:, :-1])
    with pytest.raises(Exception):
        restore_pool(SimpleReplayPool(od, ad, 1500), path, max_size=1500)
    # D4RL-style dict with 1-D rewards / terminals (loader.py:26-28)
    pool2 = SimpleReplayPool(od, ad, 1500)
    restore_pool_d4rl(pool2, 'hopper-medium-v0', dataset={'observations': data[:, :od], 'actions': data[:, od:od + ad],
                                                       'next_observations': data[:, od + ad:2 * od + ad],
                                                       'rewards': data[:, -2], 'terminals': data[:, -1].astype(bool)})
    for k, v in pool2.fields.items():
        assert np.array_equal(v.cpu().numpy(), want.fields[k]), k
    batch = pool2.random_batch(64)
    assert batch['observations'].shape == (64, od)
