"""Offline-data ingest into the device replay pool (reference mopo/off_policy/loader.py): the pool the rollout draws
its start states from (`sampler.random_batch(B)['observations']`, mopo.py:727-728).

restore_pool (loader.py:8-20) dispatches on the path exactly as the reference does; the contiguous `.npy` layout
`[obs | act | next_obs | rew | done]` (loader.py:102-126) and the BEAR pickle (loader.py:88-99) are host-side parsing
followed by ONE `add_samples` into HBM.  D4RL needs the `d4rl` + `gym` packages, as in the reference (loader.py:23-29).
"""
import gzip
import os
import pickle

import numpy as np


def restore_pool(replay_pool, experiment_root, max_size, save_path=None):
    if 'd4rl' in experiment_root:
        restore_pool_d4rl(replay_pool, experiment_root[5:])
    else:
        assert os.path.exists(experiment_root)
        if os.path.isdir(experiment_root):
            raise NotImplementedError('softlearning checkpoint directories (loader.py:32-85) hold pickled TF-era pools: '
                                      'convert them with the reference and load the contiguous .npy')
        try:
            restore_pool_contiguous(replay_pool, experiment_root)
        except Exception:                                   # loader.py:16-19: fall back to the BEAR pickle format
            restore_pool_bear(replay_pool, experiment_root)
    print('[ mbpo/off_policy ] Replay pool has size: {}'.format(replay_pool.size))


def restore_pool_d4rl(replay_pool, name, dataset=None):
    """loader.py:23-29.  `dataset` lets a caller hand in an already loaded qlearning_dataset dict."""
    if dataset is None:
        import gym
        import d4rl
        dataset = d4rl.qlearning_dataset(gym.make(name))
    data = {k: np.asarray(dataset[k]) for k in ('observations', 'actions', 'next_observations', 'rewards', 'terminals')}
    data['rewards'] = np.expand_dims(data['rewards'], axis=1) if data['rewards'].ndim == 1 else data['rewards']
    data['terminals'] = np.expand_dims(data['terminals'], axis=1) if data['terminals'].ndim == 1 else data['terminals']
    replay_pool.add_samples(data)


def restore_pool_bear(replay_pool, load_path):
    print('[ mopo/off_policy ] Loading BEAR replay pool from: {}'.format(load_path))
    data = pickle.load(gzip.open(load_path, 'rb'))
    num_trajectories = data['terminals'].sum() or 1000
    avg_return = data['rewards'].sum() / num_trajectories
    print('[ mopo/off_policy ] {} trajectories | avg return: {}'.format(num_trajectories, avg_return))
    for key in ['log_pis', 'data_policy_mean', 'data_policy_logvar']:
        data.pop(key, None)
    replay_pool.add_samples(data)


def restore_pool_contiguous(replay_pool, load_path):
    print('[ mopo/off_policy ] Loading contiguous replay pool from: {}'.format(load_path))
    data = np.load(load_path)
    state_dim = replay_pool.fields['observations'].shape[1]
    action_dim = replay_pool.fields['actions'].shape[1]
    expected_dim = state_dim + action_dim + state_dim + 1 + 1
    actual_dim = data.shape[1]
    assert expected_dim == actual_dim, 'Expected {} dimensions, found {}'.format(expected_dim, actual_dim)
    ends = np.cumsum([state_dim, action_dim, state_dim, 1, 1])
    states, actions, next_states, rewards, dones = np.split(data, ends, axis=1)[:5]
    replay_pool.add_samples({
        'observations': states,
        'actions': actions,
        'next_observations': next_states,
        'rewards': rewards,
        'terminals': dones.astype(bool),
    })
