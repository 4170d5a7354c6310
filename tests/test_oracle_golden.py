"""Oracle vs fixtures generated from the REFERENCE's own FakeEnv / static fns / FlexibleReplayPool
(tests/golden/make_golden.py), plus the closed-form KATs of SURVEY.md section 4."""
import numpy as np
import pytest
from conftest import golden_files
from oracle import ensemble as ens
from oracle.fake_env import FakeEnv
from oracle.static_fns import StaticFns
from oracle.pool import SimpleReplayPool
from oracle.rollout import rollout_model


@pytest.mark.parametrize('path', golden_files('step_'))
def test_step_matches_reference(path):
    g = np.load(path)
    domain = str(g['domain'])
    p, obs, act = ens.synthetic_case(domain, int(g['B']), seed=int(g['seed']), separate_mean_var=bool(g['separate']))
    model = ens.OracleModel(p, model_inds=g['elites'].tolist())
    env = FakeEnv(model, StaticFns(domain), penalty_coeff=float(g['coeff']), penalty_learned_var=bool(g['learned_var']))
    next_obs, rew, term, info = env.step(obs, act, deterministic=bool(g['deterministic']),
                                         noise=g['noise'], model_inds=g['model_inds'])
    # same NumPy ops in the same order -> bit-identical
    assert np.array_equal(next_obs, g['next_obs'])
    assert np.array_equal(rew, g['rew'])
    assert np.array_equal(term, g['term']) and term.dtype == bool
    assert np.array_equal(info['mean'], g['info_mean'])
    assert np.array_equal(info['std'], g['info_std'])
    assert np.array_equal(info['log_prob'], g['info_log_prob'], equal_nan=True)
    assert np.array_equal(info['dev'], g['info_dev'])
    if g['penalty'].size:
        assert np.array_equal(info['penalty'], g['penalty'])
    else:
        assert info['penalty'] is None


@pytest.mark.parametrize('path', golden_files('fwd_'))
def test_forward_matches_reference(path):
    """oracle.ensemble.forward vs the reference's own BNN.predict (constructor.py -> bnn.py finalize / _compile_outputs,
    fc.py compute_output_tensor, utils.py transform) run unmodified over the NumPy TF stand-in (tests/golden/make_golden.py):
    2-D factored, 2-D aggregated (bnn.py:261-263), 3-D per-net inputs (fc.py:100-101), separate and combined heads."""
    g = np.load(path)
    p, obs, act = ens.synthetic_case(str(g['domain']), int(g['B']), seed=int(g['seed']), separate_mean_var=bool(g['separate']))
    x = np.concatenate([obs, act], -1)
    mean, var = ens.forward(p, x, np.float32)
    np.testing.assert_allclose(mean, g['mean_fac'], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(var, g['var_fac'], rtol=1e-6)
    m, v = ens.OracleModel(p).predict(x, factored=False)
    np.testing.assert_allclose(m, g['mean'], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(v, g['var'], rtol=1e-6)
    mean3, var3 = ens.forward(p, g['x3'], np.float32)
    np.testing.assert_allclose(mean3, g['mean3'], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(var3, g['var3'], rtol=1e-6)
    # and the float64 "truth" the device paths are held to stays within fp32 rounding of the reference's float32 run
    m64, v64 = ens.forward(p, x, np.float64)
    np.testing.assert_allclose(g['mean_fac'], m64, rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize('path', golden_files('term_'))
def test_termination_matches_reference(path):
    g = np.load(path)
    nobs = g['next_obs'].astype(np.float64)      # the pipeline dtype (see make_golden.termination_case)
    fns = StaticFns(str(g['domain']))
    term = fns.termination_fn(np.zeros_like(nobs), np.zeros((len(nobs), 1), np.float32), nobs)
    assert term.dtype == bool and term.shape == (len(nobs), 1)
    assert np.array_equal(term, g['term'])
    assert 0 < term.sum() < term.size or str(g['domain']) in ('halfcheetah', 'point2denv')


def test_pool_matches_reference():
    g = np.load(golden_files('pool_')[0])
    pool = SimpleReplayPool(int(g['obs_dim']), int(g['act_dim']), int(g['max_size']))
    for i in range(int(g['n_chunks'])):
        pool.add_samples({k: g['chunk%d_%s' % (i, k)] for k in pool.fields})
        assert pool._pointer == g['ptrs'][i] and pool.size == g['sizes'][i]
    for k, v in pool.fields.items():
        assert v.dtype == g['final_' + k].dtype
        assert np.array_equal(v, g['final_' + k])


def test_kat_zero_weights():
    # zero weights/biases, mu=0 sigma=1, max_lv=.5, min_lv=-10
    p = ens.init_params(11, 3, seed=0)
    p['weights'] = [w * 0 for w in p['weights']]
    p['var_weights'] *= 0
    x = np.random.RandomState(0).randn(9, 14).astype(np.float32)
    mean, var = ens.forward(p, x)
    assert np.all(mean == 0)
    np.testing.assert_allclose(var, 0.6225047, rtol=2e-6)
    mean, lv = ens.forward(p, x, ret_log_var=True)
    np.testing.assert_allclose(lv, -0.4740041, rtol=2e-6)
    env = FakeEnv(ens.OracleModel(p), StaticFns('hopper'), penalty_coeff=1.0, penalty_learned_var=True)
    _, _, _, info = env.step(x[:, :11], x[:, 11:], noise=np.zeros((7, 9, 12)), model_inds=np.zeros(9, int))
    np.testing.assert_allclose(info['penalty'], 0.7889897 * np.sqrt(12), rtol=2e-6)


def test_kat_hopper_abs_of_bool_quirk():
    # SURVEY.md section 4 item 6 (hopper.py:12)
    nobs = np.zeros((2, 11), np.float32); nobs[:, 0] = 1.25
    nobs[0, 2] = -500.0; nobs[1, 2] = 500.0
    term = StaticFns('hopper').termination_fn(nobs, np.zeros((2, 3), np.float32), nobs)
    assert term[:, 0].tolist() == [False, True]


def test_forward_f32_vs_f64_and_modes():
    for sep in (True, False):
        p, obs, act = ens.synthetic_case('halfcheetah', 64, seed=3, separate_mean_var=sep)
        x = np.concatenate([obs, act], -1)
        m32, v32 = ens.forward(p, x, np.float32)
        m64, v64 = ens.forward(p, x, np.float64)
        assert m32.shape == (7, 64, 18) and m32.dtype == np.float32
        np.testing.assert_allclose(m32, m64, rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(v32, v64, rtol=1e-4)
        # 3-D input path == 2-D path when every net gets the same input (fc.py:98-101)
        m3, v3 = ens.forward(p, np.broadcast_to(x, (7,) + x.shape), np.float32)
        np.testing.assert_allclose(m3, m32, rtol=1e-5, atol=1e-6)


def test_rollout_invariants():
    # SURVEY.md section 4 items 2 and 4
    p, obs, act = ens.synthetic_case('hopper', 200, seed=5)
    model = ens.OracleModel(p)
    env = FakeEnv(model, StaticFns('hopper'), penalty_coeff=1.0, penalty_learned_var=True)
    pool = SimpleReplayPool(11, 3, 700)
    rng = np.random.RandomState(0)
    acts = [np.tanh(rng.standard_normal((200, 3))).astype(np.float32) for _ in range(5)]
    np.random.seed(0)
    stats, steps = rollout_model(env, pool, obs, 5, lambda i, o: acts[i][:len(o)])
    assert pool.size == min(sum(steps), 700) and pool._pointer == sum(steps) % 700
    assert stats['mean_rollout_length'] == sum(steps) / 200
    assert steps[0] == 200 and all(a >= b for a, b in zip(steps, steps[1:]))
    if len(steps) > 1 and sum(steps) <= 700:
        nxt = pool.fields['next_observations'][:steps[0]]
        keep = ~pool.fields['terminals'][:steps[0], 0]
        assert np.array_equal(pool.fields['observations'][steps[0]:steps[0] + steps[1]], nxt[keep])


def test_philox_known_answers():
    """Philox4x32-10 known-answer vectors (Random123 kat_vectors): pins oracle/philox.py, which pins the device RNG."""
    from oracle import philox
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = philox.philox4x32_10(*[np.array([c], np.uint32) for c in ctr], key[0], key[1])
        assert tuple(int(g[0]) for g in got) == want


def test_philox_streams_statistics():
    from oracle import philox
    rows = np.arange(200000)
    z = philox.normals(rows, 3, 12345, 18)
    assert abs(z.mean()) < 5e-3 and abs(z.std() - 1) < 5e-3
    e = philox.elite_index(rows, 3, 12345, 5)
    assert e.min() == 0 and e.max() == 4
    assert np.abs(np.bincount(e, minlength=5) / len(rows) - 0.2).max() < 5e-3


def test_pool_draws_with_different_offsets_are_independent():
    """The replay-pool sampler (mopo_pool_sample, the SAC sampler): draws with different ids share no Philox counter, so
    consecutive minibatches share only what chance gives (the reference draws np.random.randint per batch,
    flexible_replay_pool.py:85-87; mopo.py:801-821).  Guards the round-1 counter layout where update u+1 replayed update
    u's rows shifted by two."""
    from oracle import philox
    size, n = 1_000_000, 244
    draws = [philox.pool_indices(n, size, seed=99, offset=o) for o in range(8)]
    for a in range(8):
        assert draws[a].min() >= 0 and draws[a].max() < size
        for b in range(a + 1, 8):
            assert len(np.intersect1d(draws[a], draws[b])) <= 2            # expectation n^2 / size = 0.06
            assert not np.array_equal(draws[a][2:], draws[b][:-2])
    # uniformity over a small range (floor(u * size / 2^64))
    small = philox.pool_indices(70000, 7, seed=1, offset=5)
    assert np.all(np.abs(np.bincount(small, minlength=7) - 10000) < 400)
