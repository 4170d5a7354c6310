"""Device-resident SimpleReplayPool (reference softlearning/replay_pools/simple_replay_pool.py:37-106
over flexible_replay_pool.py:9-182): the model_pool ring buffer lives in HBM, 5 fields in the
reference's dict order, zero-initialised, modular write index, last write wins."""
import ctypes as C
import numpy as np
import torch

from . import _lib

FIELD_ORDER = ('actions', 'rewards', 'terminals', 'observations', 'next_observations')   # simple_replay_pool.py:48-70
_FIELD_ID = {'actions': _lib.F_ACTIONS, 'rewards': _lib.F_REWARDS, 'terminals': _lib.F_TERMINALS,
             'observations': _lib.F_OBS, 'next_observations': _lib.F_NEXT_OBS}


def _space_dim(space):
    if isinstance(space, (int, np.integer)):
        return int(space)
    if hasattr(space, 'spaces'):                                          # gym Dict: simple_replay_pool.py:42-46
        return int(sum(np.prod(s.shape) for s in space.spaces.values()))
    return int(np.prod(space.shape))


class _RawDevArray:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {'shape': shape, 'typestr': typestr, 'data': (ptr, False), 'version': 2}


class SimpleReplayPool:
    def __init__(self, observation_space, action_space, max_size, device=None):
        self._observation_space, self._action_space = observation_space, action_space
        self.obs_dim, self.act_dim = _space_dim(observation_space), _space_dim(action_space)
        self._max_size = int(max_size)
        self.device = torch.device('cuda', torch.cuda.current_device() if device is None else device)
        lib = _lib.load()
        h = C.c_void_p()
        _lib.check(lib.mopo_pool_create(self.obs_dim, self.act_dim, self._max_size, self.device.index or 0, C.byref(h)))
        self._handle = h
        widths = {'actions': self.act_dim, 'rewards': 1, 'terminals': 1, 'observations': self.obs_dim,
                  'next_observations': self.obs_dim}
        self.fields = {}
        with torch.cuda.device(self.device):
            for name in FIELD_ORDER:
                p = lib.mopo_pool_field(h, _FIELD_ID[name])
                ts = '|u1' if name == 'terminals' else '<f4'
                t = torch.as_tensor(_RawDevArray(p, (self._max_size, widths[name]), ts), device=self.device)
                self.fields[name] = t.view(torch.bool) if name == 'terminals' else t

    @property
    def size(self): return int(_lib.load().mopo_pool_size(self._handle))

    @property
    def _size(self): return self.size

    @property
    def _pointer(self): return int(_lib.load().mopo_pool_pointer(self._handle))

    @property
    def field_names(self): return list(FIELD_ORDER)

    def _dev(self, x, dtype):
        return torch.as_tensor(x).to(device=self.device, dtype=dtype).contiguous()

    def add_samples(self, samples):                                       # flexible_replay_pool.py:57-83
        n = samples['observations'].shape[0]
        obs = self._dev(samples['observations'], torch.float32)          # f64 -> f32 cast as fields are float32
        act = self._dev(samples['actions'], torch.float32)
        nobs = self._dev(samples['next_observations'], torch.float32)
        rew = self._dev(samples['rewards'], torch.float32).reshape(n, 1)
        term = self._dev(samples['terminals'], torch.bool).reshape(n, 1).view(torch.uint8)
        for t, w in ((obs, self.obs_dim), (act, self.act_dim), (nobs, self.obs_dim)):
            assert t.shape == (n, w), 'value shape: {}, expected: {}'.format(tuple(t.shape), (n, w))
        _lib.check(_lib.load().mopo_pool_add(self._handle, obs.data_ptr(), act.data_ptr(), nobs.data_ptr(),
                                             rew.data_ptr(), term.data_ptr(), n, _lib.stream_ptr(self.device)))

    def advance(self, n):
        """Rows [pointer, pointer + n) of `fields` were filled in place (dist.allgather_new_rows): move the ring forward."""
        _lib.check(_lib.load().mopo_pool_advance(self._handle, int(n)))

    def add_sample(self, sample):
        self.add_samples({k: (v[None, ...]) for k, v in sample.items()})

    def random_indices(self, batch_size):                                 # :85-87 (host RNG, as the reference)
        if self.size == 0:
            return np.arange(0, 0)
        return np.random.randint(0, self.size, batch_size)

    def batch_by_indices(self, indices, field_name_filter=None, as_numpy=True):   # :121-135
        idx_np = indices.cpu().numpy() if torch.is_tensor(indices) else np.asarray(indices)
        if np.any(idx_np % self._max_size > self.size):
            raise ValueError("Tried to retrieve batch with indices greater than current size")
        idx = torch.as_tensor(idx_np % self._max_size).to(device=self.device, dtype=torch.int64).contiguous()
        return self._gather(idx, as_numpy)

    def _gather(self, idx, as_numpy):
        n = idx.shape[0]
        f32 = dict(dtype=torch.float32, device=self.device)
        out = {'actions': torch.empty((n, self.act_dim), **f32), 'rewards': torch.empty((n, 1), **f32),
               'terminals': torch.empty((n, 1), dtype=torch.uint8, device=self.device),
               'observations': torch.empty((n, self.obs_dim), **f32), 'next_observations': torch.empty((n, self.obs_dim), **f32)}
        _lib.check(_lib.load().mopo_pool_gather(self._handle, idx.data_ptr(), n, out['observations'].data_ptr(),
                                                out['actions'].data_ptr(), out['next_observations'].data_ptr(),
                                                out['rewards'].data_ptr(), out['terminals'].data_ptr(),
                                                _lib.stream_ptr(self.device)))
        out['terminals'] = out['terminals'].bool()
        return {k: v.cpu().numpy() for k, v in out.items()} if as_numpy else out

    def random_batch(self, batch_size, field_name_filter=None, as_numpy=True, **kwargs):   # :89-92
        return self.batch_by_indices(self.random_indices(batch_size), as_numpy=as_numpy)

    def random_batch_device(self, batch_size, seed=0, offset=0):
        """Philox-indexed batch sampled and gathered on the device (no host RNG round trip)."""
        n = int(batch_size)
        f32 = dict(dtype=torch.float32, device=self.device)
        idx = torch.empty((n,), dtype=torch.int64, device=self.device)
        out = {'actions': torch.empty((n, self.act_dim), **f32), 'rewards': torch.empty((n, 1), **f32),
               'terminals': torch.empty((n, 1), dtype=torch.uint8, device=self.device),
               'observations': torch.empty((n, self.obs_dim), **f32), 'next_observations': torch.empty((n, self.obs_dim), **f32)}
        _lib.check(_lib.load().mopo_pool_sample(self._handle, n, seed, offset, idx.data_ptr(), out['observations'].data_ptr(),
                                                out['actions'].data_ptr(), out['next_observations'].data_ptr(),
                                                out['rewards'].data_ptr(), out['terminals'].data_ptr(),
                                                _lib.stream_ptr(self.device)))
        out['terminals'] = out['terminals'].bool()
        out['indices'] = idx
        return out

    def last_n_batch(self, last_n, field_name_filter=None, **kwargs):     # :94-99
        idx = np.arange(self._pointer - min(self.size, last_n), self._pointer) % self._max_size
        return self.batch_by_indices(idx, **kwargs)

    def return_all_samples(self, as_numpy=True):                          # :157-161
        n = self.size
        out = {k: v[:n] for k, v in self.fields.items()}
        return {k: v.cpu().numpy() for k, v in out.items()} if as_numpy else out

    def __del__(self):
        try:
            if self._handle is not None:
                self.fields = {}
                _lib.load().mopo_pool_destroy(self._handle)
                self._handle = None
        except Exception:
            pass
