"""Offline-data ingest into the device replay pool: the pool MOPO's rollouts draw their start states from
(`sampler.random_batch(B)['observations']`, reference mopo/algorithms/mopo.py:727-728).

`restore_pool` keeps the reference's entry point and its dispatch rule (mopo/off_policy/loader.py:8-20: a name containing
'd4rl', a softlearning checkpoint directory, a contiguous `.npy` matrix, else a gzip-pickled BEAR buffer).  Everything
below it is organised around the on-disk layouts, which are the contract:

  contiguous  float matrix [N, obs | act | next_obs | rew | done]                    (loader.py:102-126)
  bear        gzip pickle of a field dict + three policy-statistic keys to drop      (loader.py:88-99)
  checkpoints <root>/checkpoint_<epoch>/replay_pool.pkl, gzip pickles of field dicts (loader.py:32-49, written by
              FlexibleReplayPool.save_latest_experience, flexible_replay_pool.py:136-152)
  d4rl        d4rl.qlearning_dataset(gym.make(name)), 1-D rewards / terminals         (loader.py:23-29)

Each parser yields field dicts in the pool's schema; every dict goes to HBM with one `add_samples`.
"""
import glob
import gzip
import os
import pickle
import re

import numpy as np

POOL_FIELDS = ('observations', 'actions', 'next_observations', 'rewards', 'terminals')
_BEAR_EXTRAS = ('log_pis', 'data_policy_mean', 'data_policy_logvar')


def _pool_dims(pool):
    return pool.fields['observations'].shape[1], pool.fields['actions'].shape[1]


def _as_columns(a):
    a = np.asarray(a)
    return a[:, None] if a.ndim == 1 else a


def _report(tag, fields):
    done = int(np.asarray(fields['terminals']).sum())
    episodes = done if done else 1000            # the reference's stand-in when a buffer holds no terminal (loader.py:91)
    print('[ mopo/off_policy ] {}: {} rows | {} trajectories | avg return: {}'.format(
        tag, len(fields['rewards']), episodes, float(np.asarray(fields['rewards']).sum()) / episodes))


def parse_contiguous(matrix, obs_dim, act_dim):
    """[N, obs | act | next_obs | rew | done] -> field dict (column widths from the pool)."""
    widths = {'observations': obs_dim, 'actions': act_dim, 'next_observations': obs_dim, 'rewards': 1, 'terminals': 1}
    order = ('observations', 'actions', 'next_observations', 'rewards', 'terminals')
    total = sum(widths[k] for k in order)
    if matrix.ndim != 2 or matrix.shape[1] != total:
        raise AssertionError('Expected {} dimensions, found {}'.format(total, matrix.shape[1] if matrix.ndim == 2 else matrix.shape))
    out, col = {}, 0
    for k in order:
        out[k] = matrix[:, col:col + widths[k]]
        col += widths[k]
    out['terminals'] = out['terminals'].astype(bool)
    return out


def _read_gzip_pickle(path):
    with gzip.open(path, 'rb') as f:
        return pickle.load(f)


def read_file(path, obs_dim, act_dim):
    """A single file: the contiguous matrix if NumPy can read it as one, else the BEAR pickle (the reference tries them in
    this order, loader.py:15-19)."""
    try:
        fields = parse_contiguous(np.load(path), obs_dim, act_dim)
        tag = 'contiguous pool ' + path
    except Exception:
        raw = _read_gzip_pickle(path)
        fields = {k: v for k, v in raw.items() if k not in _BEAR_EXTRAS}
        tag = 'BEAR pool ' + path
    _report(tag, fields)
    return fields


def checkpoint_files(root, max_size):
    """replay_pool.pkl of every checkpoint_<epoch> under root, by ascending epoch; a 250k-row request skips the two
    earliest checkpoints, as the reference does (loader.py:41-43)."""
    found = []
    for d in glob.glob(os.path.join(root, 'checkpoint_*')):
        m = re.search(r'checkpoint_(\d+)$', d)
        if m:
            found.append((int(m.group(1)), os.path.join(d, 'replay_pool.pkl')))
    found.sort()
    if max_size == 250e3:
        found = found[2:]
    return [p for _, p in found]


def restore_pool_softlearning(replay_pool, experiment_root, max_size, save_path=None):
    """loader.py:32-85: append checkpointed experience until the pool holds max_size rows; optionally dump the restored
    rows and their return statistics (1000-step or terminal-delimited paths) next to each other under save_path."""
    for path in checkpoint_files(experiment_root, max_size):
        chunk = _read_gzip_pickle(path)
        rows = {len(v) for v in chunk.values()}
        assert len(rows) == 1, {k: np.shape(v) for k, v in chunk.items()}
        replay_pool.add_samples(chunk)
        print('[ mopo/off_policy ] checkpoint {}: pool size {}'.format(path, replay_pool.size))
        if replay_pool.size >= max_size:
            break
    if save_path is None:
        return
    data = replay_pool.return_all_samples()
    rew, term = np.asarray(data['rewards']).ravel(), np.asarray(data['terminals']).ravel().astype(bool)
    returns, acc, start = [], 0.0, 0
    for i in range(len(rew)):                    # a path ends at a terminal or after 1000 steps (loader.py:60-66)
        acc += float(rew[i])
        if term[i] or i - start + 1 == 1000:
            returns.append(acc)
            acc, start = 0.0, i + 1
    n_paths = max(len(returns), 1)
    stats = dict(avg_return=float(rew.sum()) / n_paths, buffer_max=max(returns, default=-np.inf), buffer_min=min(returns, default=np.inf))
    size = replay_pool.size
    with open(os.path.join(save_path, 'pool_stat_{}.pkl'.format(size)), 'wb') as f:
        pickle.dump(stats, f)
    with open(os.path.join(save_path, 'pool_{}.pkl'.format(size)), 'wb') as f:
        pickle.dump({k: np.asarray(v) for k, v in data.items()}, f)
    print('[ mopo/off_policy ] restored pool saved under {}: {}'.format(save_path, stats))


def restore_pool_d4rl(replay_pool, name, dataset=None):
    """loader.py:23-29.  `dataset` lets a caller hand in an already loaded qlearning_dataset dict (no gym / d4rl import)."""
    if dataset is None:
        import gym
        import d4rl
        dataset = d4rl.qlearning_dataset(gym.make(name))
    replay_pool.add_samples({k: _as_columns(dataset[k]) if k in ('rewards', 'terminals') else np.asarray(dataset[k])
                             for k in POOL_FIELDS})


def restore_pool_contiguous(replay_pool, load_path):
    replay_pool.add_samples(parse_contiguous(np.load(load_path), *_pool_dims(replay_pool)))


def restore_pool_bear(replay_pool, load_path):
    raw = _read_gzip_pickle(load_path)
    replay_pool.add_samples({k: v for k, v in raw.items() if k not in _BEAR_EXTRAS})


def restore_pool(replay_pool, experiment_root, max_size, save_path=None):
    if 'd4rl' in experiment_root:
        restore_pool_d4rl(replay_pool, experiment_root[5:])
    else:
        assert os.path.exists(experiment_root)
        if os.path.isdir(experiment_root):
            restore_pool_softlearning(replay_pool, experiment_root, max_size, save_path)
        else:
            replay_pool.add_samples(read_file(experiment_root, *_pool_dims(replay_pool)))
    print('[ mbpo/off_policy ] Replay pool has size: {}'.format(replay_pool.size))
