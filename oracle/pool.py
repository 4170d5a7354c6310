"""NumPy restatement of the reference model_pool ring buffer.  TEST INFRASTRUCTURE ONLY.

Follows softlearning/replay_pools/flexible_replay_pool.py:39-48 (zero-init fields, _advance),
:57-83 (add_samples, modular index, last-write-wins on wrap), :85-92 (random_batch),
:121-135 (batch_by_indices), :157-161 (return_all_samples) and the 5-field schema of
softlearning/replay_pools/simple_replay_pool.py:48-70.
"""
import numpy as np


class SimpleReplayPool:
    def __init__(self, obs_dim, act_dim, max_size):
        self._max_size = int(max_size)
        self._pointer = 0
        self._size = 0
        n = self._max_size
        self.fields = {                                              # simple_replay_pool.py:48-70 (dict order)
            'actions': np.zeros((n, act_dim), np.float32),
            'rewards': np.zeros((n, 1), np.float32),
            'terminals': np.zeros((n, 1), bool),
            'observations': np.zeros((n, obs_dim), np.float32),
            'next_observations': np.zeros((n, obs_dim), np.float32),
        }

    @property
    def size(self):
        return self._size

    def add_samples(self, samples):                                  # flexible_replay_pool.py:57-83
        num = samples['observations'].shape[0]
        index = np.arange(self._pointer, self._pointer + num) % self._max_size
        for name in self.fields:
            self.fields[name][index] = samples[name]
        self._pointer = (self._pointer + num) % self._max_size       # :45-48
        self._size = min(self._size + num, self._max_size)

    def random_indices(self, batch_size):                            # :85-87
        if self._size == 0:
            return np.arange(0, 0)
        return np.random.randint(0, self._size, batch_size)

    def batch_by_indices(self, indices):                             # :121-135
        if np.any(indices % self._max_size > self.size):
            raise ValueError("Tried to retrieve batch with indices greater than current size")
        return {k: v[indices] for k, v in self.fields.items()}

    def random_batch(self, batch_size):
        return self.batch_by_indices(self.random_indices(batch_size))

    def return_all_samples(self):                                    # :157-161
        return {k: v[:self.size] for k, v in self.fields.items()}
