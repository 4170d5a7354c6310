"""Host-side logic of the shims that needs no GPU: the rollout-length schedule and model-pool sizing of
mopo/algorithms/mopo.py:675-711, format_samples_for_training (constructor.py:46-57), the scaler's fit rule
(models/utils.py:69-86) and the layer spec (fc.py:48-52)."""
import numpy as np
import pytest


class _Pool:
    _observation_space, _action_space = 17, 6


def _ro(**kw):
    from mopo_b200.algorithms.mopo import ModelRollout
    return ModelRollout(fake_env=None, env_pool=_Pool(), **kw)


@pytest.mark.parametrize('schedule', [(20, 100, 1, 1), (20, 100, 1, 5), (0, 10, 5, 25), (20, 100, 5, 5)])
def test_rollout_length_schedule(schedule, capsys):
    ro = _ro(rollout_schedule=schedule)
    lo, hi, a, b = schedule
    for epoch in (0, lo, lo + 1, (lo + hi) // 2, hi, hi + 50):
        ro._epoch = epoch
        ro._set_rollout_length()
        if epoch <= lo:                                       # mopo.py:677-678
            want = a
        else:                                                 # mopo.py:680-682
            want = int(min((epoch - lo) / (hi - lo), 1) * (b - a) + a)
        assert ro._rollout_length == want
    assert '[ Model Length ]' in capsys.readouterr().out


def test_model_pool_sizing_and_growth(monkeypatch):
    """mopo.py:689-711: size = retain_epochs * int(rollout_length * batch * epoch_length / train_freq); a longer rollout
    re-allocates and carries the old samples over."""
    import mopo_b200.algorithms.mopo as M
    made = []

    class FakePool:
        def __init__(self, obs_space, act_space, max_size, device=None):
            self._max_size, self.size, self.samples = max_size, 0, None
            made.append(self)

        def return_all_samples(self, as_numpy=True):
            return {'n': self.size}

        def add_samples(self, samples):
            self.size = samples['n']

    monkeypatch.setattr(M, 'SimpleReplayPool', FakePool)

    class Env:
        class model:
            class device:
                index = 0
    ro = M.ModelRollout(fake_env=Env(), env_pool=_Pool(), rollout_batch_size=50000, rollout_schedule=(20, 100, 5, 5),
                        epoch_length=1000, model_train_freq=1000, model_retain_epochs=5)
    ro._set_rollout_length()
    ro._reallocate_model_pool()
    assert made[-1]._max_size == 5 * int(5 * 50000 * 1000 / 1000) == 1250000
    made[-1].size = 777
    ro._reallocate_model_pool()                               # same size: nothing new
    assert len(made) == 1
    ro._rollout_length = 7
    ro._reallocate_model_pool()
    assert len(made) == 2 and made[-1]._max_size == 5 * 7 * 50000 and made[-1].size == 777 and ro._model_pool is made[-1]


def test_format_samples_and_scaler_fit():
    from mopo_b200.models.constructor import format_samples_for_training
    from mopo_b200.models.bnn import TensorStandardScaler
    rng = np.random.RandomState(0)
    s = {'observations': rng.standard_normal((50, 4)), 'actions': rng.standard_normal((50, 2)),
         'next_observations': rng.standard_normal((50, 4)), 'rewards': rng.standard_normal((50, 1))}
    x, y = format_samples_for_training(s)
    assert np.array_equal(x, np.concatenate([s['observations'], s['actions']], -1))
    assert np.array_equal(y, np.concatenate([s['rewards'], s['next_observations'] - s['observations']], -1))   # [reward | delta obs]
    data = rng.standard_normal((200, 6)).astype(np.float32)
    data[:, 3] = 2.5                                          # constant column: sigma < 1e-12 -> 1 (utils.py:81)
    sc = TensorStandardScaler(6)
    assert np.array_equal(sc.transform(data), data)           # mu 0, sigma 1 until fit
    sc.fit(data)
    np.testing.assert_allclose(sc.mu[0], data.mean(0), rtol=1e-6)
    assert sc.sigma[0, 3] == 1.0 and sc.fitted
    np.testing.assert_allclose(sc.inverse_transform(sc.transform(data)), data, atol=1e-5)
    assert np.array_equal(sc.cached_mu, sc.mu)


def test_fc_spec_mirrors_reference_repr_and_init_law():
    from mopo_b200.models.fc import FC
    l = FC(200, input_dim=23, activation='swish', weight_decay=2.5e-5, ensemble_size=7)
    assert repr(l) == "FC(output_dim=200, input_dim=23, activation='swish', weight_decay=2.5e-05, ensemble_size=7)"
    c = l.copy()
    c.set_output_dim(36)
    assert l.get_output_dim() == 200 and c.get_output_dim() == 36
    with pytest.raises(ValueError):
        FC(10, activation='gelu')
    l.construct_vars(np.random.RandomState(0))
    std = 1.0 / (2.0 * np.sqrt(23))
    assert l.weights.shape == (7, 23, 200) and np.abs(l.weights).max() <= 2 * std + 1e-7     # truncated at 2 sigma (fc.py:145-149)
    assert abs(l.weights.std() - 0.88 * std) < 0.05 * std and not l.biases.any()
    with pytest.raises(RuntimeError):
        FC(10).construct_vars(np.random.RandomState(0))


def test_offline_loader_formats(tmp_path, capsys):
    """mopo/off_policy/loader.py:8-126 on the host side (the pool here is the oracle's NumPy ring; the GPU tests run the same
    loader into the device pool): contiguous matrix, BEAR pickle with its three extra keys, softlearning checkpoint
    directory (ascending epochs, stop at max_size, the 250k rule, the saved pool + return statistics), bad widths."""
    import gzip, os, pickle
    from mopo_b200.off_policy import loader
    from oracle.pool import SimpleReplayPool
    od, ad, n = 5, 2, 300
    rng = np.random.RandomState(4)
    fields = {'observations': rng.standard_normal((n, od)), 'actions': rng.standard_normal((n, ad)),
              'next_observations': rng.standard_normal((n, od)), 'rewards': rng.standard_normal((n, 1)),
              'terminals': rng.uniform(size=(n, 1)) < 0.02}
    mat = np.concatenate([fields[k].astype(np.float64) for k in ('observations', 'actions', 'next_observations', 'rewards', 'terminals')], 1)
    want = SimpleReplayPool(od, ad, 1000); want.add_samples(fields)

    def same(pool):
        return pool.size == want.size and all(np.array_equal(pool.fields[k], want.fields[k]) for k in want.fields)

    # contiguous
    p_npy = str(tmp_path / 'pool.npy'); np.save(p_npy, mat)
    pool = SimpleReplayPool(od, ad, 1000); loader.restore_pool(pool, p_npy, max_size=1000)
    assert same(pool)
    parsed = loader.parse_contiguous(mat, od, ad)
    assert parsed['terminals'].dtype == bool and parsed['actions'].shape == (n, ad)
    with pytest.raises(AssertionError, match='Expected 14 dimensions, found 13'):
        loader.parse_contiguous(mat[:, :-1], od, ad)
    # BEAR: a gzip pickle that np.load rejects -> second format; extras dropped
    p_bear = str(tmp_path / 'bear.pkl')
    with gzip.open(p_bear, 'wb') as f:
        pickle.dump(dict(fields, log_pis=np.zeros(n), data_policy_mean=np.zeros((n, ad)), data_policy_logvar=np.zeros((n, ad))), f)
    pool = SimpleReplayPool(od, ad, 1000); loader.restore_pool(pool, p_bear, max_size=1000)
    assert same(pool)
    assert 'BEAR pool' in capsys.readouterr().out
    # softlearning checkpoints: epochs out of lexical order, 100 rows each
    root = tmp_path / 'run'
    for ep, lo in ((2, 0), (10, 100), (30, 200)):
        d = root / ('checkpoint_%d' % ep); d.mkdir(parents=True)
        with gzip.open(str(d / 'replay_pool.pkl'), 'wb') as f:
            pickle.dump({k: v[lo:lo + 100] for k, v in fields.items()}, f)
    assert [os.path.basename(os.path.dirname(p)) for p in loader.checkpoint_files(str(root), 1000)] == ['checkpoint_2', 'checkpoint_10', 'checkpoint_30']
    assert len(loader.checkpoint_files(str(root), 250e3)) == 1
    out = tmp_path / 'out'; out.mkdir()
    pool = SimpleReplayPool(od, ad, 1000); loader.restore_pool(pool, str(root), max_size=1000, save_path=str(out))
    assert same(pool)
    stats = pickle.load(open(str(out / 'pool_stat_300.pkl'), 'rb'))
    ends = [i for i in range(n) if fields['terminals'][i, 0]]
    paths, s = [], 0
    for e in ends: paths.append(fields['rewards'][s:e + 1].astype(np.float32).sum()); s = e + 1
    assert np.isclose(stats['buffer_max'], max(paths), rtol=1e-5) and np.isclose(stats['avg_return'], fields['rewards'].astype(np.float32).sum() / len(paths), rtol=1e-4)
    saved = pickle.load(open(str(out / 'pool_300.pkl'), 'rb'))
    assert np.array_equal(saved['observations'], want.fields['observations'][:n])
    pool = SimpleReplayPool(od, ad, 1000); loader.restore_pool(pool, str(root), max_size=150)      # stops after two checkpoints
    assert pool.size == 200
