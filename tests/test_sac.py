"""The SAC update (reference mopo/algorithms/mopo.py:271-457, 801-843).

CPU: the oracle's analytic gradients against central finite differences of the restated losses; the batch mix of
_training_batch.  GPU: mopo_sac_update against the float64 oracle (losses and every variable after a few updates, with
supplied and with Philox policy noise), the pool-sampling path, and the actor export into the rollout policy."""
import numpy as np
import pytest

from oracle import sac as osac
from oracle import philox


def _batch(rng, n, od, ad):
    return {'observations': rng.standard_normal((n, od)).astype(np.float32),
            'actions': np.tanh(rng.standard_normal((n, ad))).astype(np.float32),
            'next_observations': rng.standard_normal((n, od)).astype(np.float32),
            'rewards': rng.standard_normal((n, 1)).astype(np.float32),
            'terminals': rng.rand(n, 1) < 0.3}


def _f64(batch):
    return {k: (v.astype(np.float64) if v.dtype != bool else v) for k, v in batch.items()}


def test_oracle_sac_gradients_match_finite_differences():
    rng = np.random.RandomState(0)
    od, ad, n = 5, 3, 9
    st = osac.init_state(od, ad, hidden=16, seed=1, bias_std=0.2)
    st['log_alpha'] = np.array(-0.3)
    for k in ('q1_targ', 'q2_targ'):
        for w in st[k]['W']:
            w += rng.standard_normal(w.shape) * 0.05
    batch = _f64(_batch(rng, n, od, ad))
    e1, e2 = rng.standard_normal((n, ad)), rng.standard_normal((n, ad))
    hp = dict(osac.DEFAULT_HP, target_entropy=-3.0)
    out, cache = osac.losses(st, batch, e1, e2, hp)
    g = osac.grads(st, batch, e1, hp, cache)
    for net, lk in (('q1', 'q1_loss'), ('q2', 'q2_loss'), ('pi', 'policy_loss')):
        for k in ('W', 'b'):
            for i, a in enumerate(st[net][k]):
                for _ in range(6):
                    j = tuple(rng.randint(s) for s in a.shape)
                    old, h = a[j], 1e-6
                    a[j] = old + h; lp = osac.losses(st, batch, e1, e2, hp)[0][lk]
                    a[j] = old - h; lm = osac.losses(st, batch, e1, e2, hp)[0][lk]
                    a[j] = old
                    fd = (lp - lm) / (2 * h)
                    assert abs(fd - g[net][k][i][j]) <= 1e-5 * abs(fd) + 1e-9, (net, k, i, j)
    # alpha: d/d log_alpha of -mean(log_alpha (logp + H)) with logp held fixed
    np.testing.assert_allclose(g['log_alpha'], -(cache['logp_pi'] + hp['target_entropy']).mean())


def test_oracle_sac_update_known_answers():
    rng = np.random.RandomState(1)
    st = osac.init_state(4, 2, hidden=8, seed=2)
    opt = osac.adam_init(st)
    before_t = [w.copy() for w in st['q1_targ']['W']]
    before_m = [w.copy() for w in st['q1']['W']]
    batch = _f64(_batch(rng, 16, 4, 2))
    hp = dict(osac.DEFAULT_HP, target_entropy=-2.0)
    out = osac.update(st, opt, batch, rng.standard_normal((16, 2)), rng.standard_normal((16, 2)), hp)
    assert opt['t'] == 1 and all(np.isfinite(v) for v in out.values())
    # first Adam step moves log_alpha by lr * sign(grad); Polyak mixes the UPDATED main weights into the target
    assert abs(abs(float(st['log_alpha'])) - hp['policy_lr']) < 1e-6
    for bt, bm, t, m in zip(before_t, before_m, st['q1_targ']['W'], st['q1']['W']):
        assert np.array_equal(bt, bm)                                   # target_init
        np.testing.assert_allclose(t, (1 - hp['tau']) * bt + hp['tau'] * m, rtol=1e-12)


def test_training_batch_mix():
    """mopo.py:803-804: env_batch_size = int(batch_size * real_ratio); the rest comes from the model pool."""
    for bs, rr, want in ((256, 0.05, (12, 244)), (256, 0.5, (128, 128)), (100, 0.0, (0, 100)), (256, 1.0, (256, 0))):
        env = int(bs * rr)
        assert (env, bs - env) == want


# ------------------------------------------------------------------------------------------------------------------
def _state_to_device(sac, st):
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):          # the device holds float32 variables: start the oracle from the same values
        for k in ('W', 'b'):
            st[name][k] = [a.astype(np.float32).astype(np.float64) for a in st[name][k]]
    st['log_alpha'] = np.array(np.float64(np.float32(st['log_alpha'])))
    sac.set_params(st['pi']['W'], st['pi']['b'], st['q1']['W'], st['q1']['b'], st['q2']['W'], st['q2']['b'], float(st['log_alpha']))


def _compare(got, st, atol):
    worst = 0.0
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
        for k in ('W', 'b'):
            for a, b in zip(got[name][k], st[name][k]):
                worst = max(worst, float(np.abs(a - b.reshape(a.shape)).max()))
    worst = max(worst, abs(got['log_alpha'] - float(st['log_alpha'])))
    assert worst <= atol, worst


@pytest.mark.gpu
@pytest.mark.parametrize('shape', [(17, 6, 256, 256), (11, 3, 64, 100), (5, 2, 32, 7), (45, 17, 30, 33)])   # last: odd pitches -> scalar-load fallbacks, 5 Philox blocks
def test_sac_update_matches_oracle(shape):
    from mopo_b200.algorithms.sac import SACUpdate, LOSS_NAMES
    od, ad, H, n = shape
    rng = np.random.RandomState(od)
    st = osac.init_state(od, ad, hidden=H, seed=3, bias_std=0.1)
    st['log_alpha'] = np.array(-0.2)
    hp = dict(osac.DEFAULT_HP, target_entropy=-float(ad))
    sac = SACUpdate(od, ad, hidden_sizes=(H, H), device=0)
    _state_to_device(sac, st)
    opt = osac.adam_init(st)
    K = 3
    for it in range(K):
        batch = _batch(rng, n, od, ad)
        eps = rng.standard_normal((2 * n, ad)).astype(np.float32)
        got = sac._do_training(it, batch, eps=eps)
        want = osac.update(st, opt, _f64(batch), eps[:n].astype(np.float64), eps[n:].astype(np.float64), hp)
        np.testing.assert_allclose([got[k] for k in LOSS_NAMES],
                                   [want['q1_loss'], want['q2_loss'], want['policy_loss'], want['alpha_loss']], rtol=2e-4, atol=1e-6)
    # an Adam step moves an entry by ~lr = 3e-4; fp32 vs float64 gradients keep the variables far closer than one step
    _compare(sac.get_params(), st, 1e-5 * K)


@pytest.mark.gpu
def test_sac_update_philox_noise_and_target_interval():
    """eps = NULL: the policy noise is Philox stream 4 keyed (seed, update index 1.., row); target_update_interval = 2
    applies the Polyak update on updates 0, 2, ... (iteration % interval == 0, mopo.py:839)."""
    from mopo_b200.algorithms.sac import SACUpdate
    od, ad, H, n = 11, 3, 64, 96
    rng = np.random.RandomState(5)
    st = osac.init_state(od, ad, hidden=H, seed=7)
    hp = dict(osac.DEFAULT_HP, target_entropy=-3.0, tau=0.1)
    sac = SACUpdate(od, ad, hidden_sizes=(H, H), device=0, tau=0.1, target_update_interval=2, seed=4242)
    _state_to_device(sac, st)
    opt = osac.adam_init(st)
    for it in range(3):
        batch = _batch(rng, n, od, ad)
        sac._do_training(it, batch, return_losses=False)
        eps = philox.normals(np.arange(2 * n), it + 1, 4242, ad, stream=philox.STREAM_SAC)
        targ_before = {k: {kk: [a.copy() for a in st[k][kk]] for kk in ('W', 'b')} for k in ('q1_targ', 'q2_targ')}
        osac.update(st, opt, _f64(batch), eps[:n], eps[n:], hp)
        if it % 2 != 0:
            for k in targ_before:
                st[k] = targ_before[k]
    _compare(sac.get_params(), st, 5e-5)


@pytest.mark.gpu
def test_sac_train_from_pools_equals_explicit_batches_and_exports_actor():
    """mopo_sac_update_from_pools == drawing the same Philox-indexed rows through the pool API (env rows first, then
    model rows, mopo.py:808-815) and calling mopo_sac_update: bit for bit.  Then the trained actor is exported into the
    rollout policy handle and acts like the oracle's forward of the trained variables."""
    import torch
    from mopo_b200.algorithms.sac import SACUpdate
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.policies import GaussianPolicy
    from oracle import policy as opol
    od, ad = 17, 6
    rng = np.random.RandomState(2)
    env_pool, model_pool = SimpleReplayPool(od, ad, 3000), SimpleReplayPool(od, ad, 20000)
    env_pool.add_samples(_batch(rng, 3000, od, ad))
    model_pool.add_samples(_batch(rng, 15000, od, ad))
    a = SACUpdate(od, ad, device=0, seed=99)
    b = SACUpdate(od, ad, device=0, seed=99)                       # same seed -> same glorot init
    assert a.batch_sizes() == (12, 244)
    K = 10
    side = torch.cuda.Stream()                                     # a caller-chosen stream: the updates queue on it back to back
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        a.train_from_pools(env_pool, model_pool, n_updates=K - 2)
    torch.cuda.synchronize()
    logs = a.train_from_pools(env_pool, model_pool, n_updates=2, return_losses=True)     # with the loss read-back, same sequence
    assert all(np.isfinite(v) for v in logs.values())
    # the per-environment-step pattern (_do_training_repeats, mopo.py:780-799): one update per call
    with torch.cuda.stream(side):
        for _ in range(5):
            a.train_from_pools(env_pool, model_pool, n_updates=1)
    torch.cuda.synchronize()
    K += 5
    for it in range(K):
        eb = env_pool.random_batch_device(12, seed=99, offset=2 * it)
        mb = model_pool.random_batch_device(244, seed=99, offset=2 * it + 1)
        batch = {k: torch.cat([eb[k], mb[k]], 0) for k in ('observations', 'actions', 'next_observations', 'rewards', 'terminals')}
        b._do_training(it, batch, return_losses=False)
    pa, pb = a.get_params(), b.get_params()
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
        for k in ('W', 'b'):
            for x, y in zip(pa[name][k], pb[name][k]):
                assert np.array_equal(x, y), (name, k)
    assert pa['log_alpha'] == pb['log_alpha'] and pa['log_alpha'] != 0.0
    pol = GaussianPolicy(od, ad, device=0)
    a.export_policy(pol)
    obs = rng.standard_normal((200, od)).astype(np.float32)
    act, _ = pol.get_action_meta(obs, None, deterministic=True)
    want, _, _ = opol.forward(pa['pi']['W'], pa['pi']['b'], obs, deterministic=True)
    np.testing.assert_allclose(act, want, rtol=1e-3, atol=2e-5)


@pytest.mark.gpu
def test_pool_sampler_matches_oracle_and_draws_are_independent():
    """mopo_pool_sample's indices equal the NumPy restatement bit for bit, and batches of consecutive draw ids (what
    consecutive SAC updates use) share no rows beyond chance."""
    from mopo_b200.replay_pools import SimpleReplayPool
    from oracle import philox
    rng = np.random.default_rng(3)
    pool = SimpleReplayPool(5, 2, 50000, device=0)
    n = 40000
    pool.add_samples({'observations': rng.standard_normal((n, 5)).astype(np.float32), 'actions': rng.standard_normal((n, 2)).astype(np.float32),
                      'next_observations': rng.standard_normal((n, 5)).astype(np.float32), 'rewards': np.arange(n, dtype=np.float32)[:, None],
                      'terminals': np.zeros((n, 1), bool)})
    seen = []
    for off in (0, 1, 2, 3, (1 << 33) + 5):
        b = pool.random_batch_device(244, seed=99, offset=off)
        idx = b['rewards'].cpu().numpy()[:, 0].astype(np.int64)           # rewards hold the row id
        assert np.array_equal(idx, philox.pool_indices(244, n, 99, off))
        seen.append(idx)
    for a in range(len(seen)):
        for c in range(a + 1, len(seen)):
            assert len(np.intersect1d(seen[a], seen[c])) <= 8                # expectation 244^2 / 40000 = 1.5


@pytest.mark.gpu
def test_full_device_loop_and_checkpoint(tmp_path):
    """Rows f1-f4 together, nothing leaving HBM: offline pool -> BNN.train -> elites -> actor-driven rollout into the model
    pool -> SAC updates drawn from both pools -> actor exported -> next rollout uses the new actor; then the SAC variables
    survive a save / load round trip (targets included) and training continues identically from the checkpoint."""
    import torch
    from mopo_b200 import _lib
    import mopo_b200.static as static
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    from mopo_b200.algorithms.sac import SACUpdate
    from mopo_b200.policies import GaussianPolicy
    od, ad, n = 11, 3, 4000
    rng = np.random.RandomState(0)
    data = _batch(rng, n, od, ad)
    data['next_observations'] = (data['observations'] + 0.1 * np.tanh(data['observations'] @ rng.standard_normal((od, od)) * 0.3)).astype(np.float32)
    data['terminals'] = np.zeros((n, 1), bool)
    env_pool = SimpleReplayPool(od, ad, n)
    env_pool.add_samples(data)
    model = construct_model(obs_dim=od, act_dim=ad, num_networks=7, num_elites=5, separate_mean_var=True, precision=_lib.PREC_BF16, seed=1)
    env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=2)
    pol = GaussianPolicy(od, ad, device=0, precision=_lib.PREC_BF16, seed=3)
    sac = SACUpdate(od, ad, device=0, seed=4)
    sac.export_policy(pol)
    ro = ModelRollout(env, env_pool, rollout_batch_size=1000, rollout_schedule=(0, 1, 2, 2), epoch_length=1000, model_train_freq=1000,
                      model_retain_epochs=2, policy=pol, seed=5)
    ro._train_model(batch_size=256, max_epochs=3, holdout_ratio=0.2)
    ro._set_rollout_length()
    ro._reallocate_model_pool()
    obs0 = data['observations'][:1000]
    ro._rollout_model(1000, start_obs=obs0, quiet=True)
    torch.cuda.synchronize()
    a_before = ro._model_pool.fields['actions'][:1000].clone()
    assert ro._model_pool.size == 2000
    logs = sac.train_from_pools(env_pool, ro._model_pool, n_updates=20, return_losses=True)
    assert all(np.isfinite(v) for v in logs.values())
    sac.export_policy(pol)
    ro._rollout_model(1000, start_obs=obs0, quiet=True)
    torch.cuda.synchronize()
    a_after = ro._model_pool.fields['actions'][2000:3000]
    assert ro._model_pool.size == 4000 and not torch.equal(a_before, a_after)          # the updated actor drives the rollout
    # checkpoint round trip
    path = str(tmp_path / 'sac.npz')
    sac.save(path)
    clone = SACUpdate(od, ad, device=0, seed=4)
    clone.load(path)
    pa, pb = sac.get_params(), clone.get_params()
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
        for k in ('W', 'b'):
            for x, y in zip(pa[name][k], pb[name][k]):
                assert np.array_equal(x, y), (name, k)
    assert pa['log_alpha'] == pb['log_alpha']
    assert not np.array_equal(pa['q1']['W'][0], pa['q1_targ']['W'][0])                  # targets lag the main critics


@pytest.mark.gpu
def test_sac_real_ratio_one_uses_only_the_env_pool():
    """real_ratio == 1.0 (mopo.py:816-819): no model pool; the whole batch comes from the env pool."""
    import torch
    from mopo_b200.algorithms.sac import SACUpdate
    from mopo_b200.replay_pools import SimpleReplayPool
    od, ad = 11, 3
    rng = np.random.RandomState(3)
    env_pool = SimpleReplayPool(od, ad, 500)
    env_pool.add_samples(_batch(rng, 500, od, ad))
    a = SACUpdate(od, ad, hidden_sizes=(64, 64), device=0, real_ratio=1.0, batch_size=64, seed=5)
    b = SACUpdate(od, ad, hidden_sizes=(64, 64), device=0, real_ratio=1.0, batch_size=64, seed=5)
    assert a.batch_sizes() == (64, 0)
    a.train_from_pools(env_pool, None, n_updates=3)
    for it in range(3):
        eb = env_pool.random_batch_device(64, seed=5, offset=2 * it)
        b._do_training(it, {k: eb[k] for k in ('observations', 'actions', 'next_observations', 'rewards', 'terminals')}, return_losses=False)
    torch.cuda.synchronize()
    pa, pb = a.get_params(), b.get_params()
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
        for k in ('W', 'b'):
            for x, y in zip(pa[name][k], pb[name][k]):
                assert np.array_equal(x, y), (name, k)
    empty = SimpleReplayPool(od, ad, 10)
    with pytest.raises(RuntimeError):                     # sampling an empty pool (np.random.randint(0, 0) raises in the reference)
        a.train_from_pools(empty, None, n_updates=1)


@pytest.mark.gpu
def test_checkpoint_resume_continues_the_same_sequence(tmp_path):
    """save() / load() carry the Adam slots, the Adam step and the update index behind the random draws: 4 + 3 updates with a
    checkpoint in between equal 7 uninterrupted updates bit for bit (a resumed run that restarted its optimisers or replayed
    its draws from update 0 would not)."""
    from mopo_b200.algorithms.sac import SACUpdate
    from mopo_b200.replay_pools import SimpleReplayPool
    od, ad = 11, 3
    rng = np.random.RandomState(8)
    env_pool, model_pool = SimpleReplayPool(od, ad, 2000), SimpleReplayPool(od, ad, 9000)
    env_pool.add_samples(_batch(rng, 2000, od, ad))
    model_pool.add_samples(_batch(rng, 9000, od, ad))
    a = SACUpdate(od, ad, device=0, seed=5)
    b = SACUpdate(od, ad, device=0, seed=5)
    a.train_from_pools(env_pool, model_pool, n_updates=7)
    b.train_from_pools(env_pool, model_pool, n_updates=4)
    path = str(tmp_path / 'sac.npz')
    b.save(path)
    c = SACUpdate(od, ad, device=0, seed=5)
    c.load(path)
    st = c.get_opt_state()
    assert st['t'] == 4 and st['draw'] == 4 and any(np.abs(m).max() > 0 for m in st['m'])
    c.train_from_pools(env_pool, model_pool, n_updates=3)
    pa, pc = a.get_params(), c.get_params()
    for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
        for k in ('W', 'b'):
            for x, y in zip(pa[name][k], pc[name][k]):
                assert np.array_equal(x, y), (name, k)
    assert pa['log_alpha'] == pc['log_alpha']
    oa, oc = a.get_opt_state(), c.get_opt_state()
    assert oa['t'] == oc['t'] == 7 and oa['draw'] == oc['draw']
    for x, y in zip(oa['m'] + oa['v'], oc['m'] + oc['v']):
        assert np.array_equal(x, y)

