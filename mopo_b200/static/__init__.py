"""Per-environment static functions, device-side (reference: mopo/static/__init__.py:7-25).

``mopo_b200.static[domain]`` mirrors ``mopo.static[domain]``: an object with
``termination_fn(obs, act, next_obs) -> bool[B,1]``.  The predicate itself is compiled into the
step kernel (``term_kind``); ``termination_fn`` here runs the same device code on caller arrays.
"""
import sys
import types
import numpy as np

from .. import _lib

_KINDS = {
    'halfcheetah': _lib.TERM_NEVER, 'halfcheetahjump': _lib.TERM_NEVER, 'halfcheetahvel': _lib.TERM_NEVER,
    'halfcheetahveljump': _lib.TERM_NEVER, 'point2denv': _lib.TERM_NEVER, 'point2dwallenv': _lib.TERM_NEVER,
    'pendulum': _lib.TERM_NEVER,       # pendulum.py:9 returns float zeros; bool here so `~term` works
    'hopper': _lib.TERM_HOPPER, 'walker2d': _lib.TERM_WALKER2D,
    'ant': _lib.TERM_ANT, 'antangle': _lib.TERM_ANT, 'humanoid': _lib.TERM_HUMANOID,
}


class StaticFns:
    def __init__(self, domain):
        self.domain = domain
        self.term_kind = _KINDS[domain]

    def termination_fn(self, obs, act, next_obs):
        import torch
        assert len(obs.shape) == len(next_obs.shape) == len(act.shape) == 2      # mopo/static/*.py:7
        is_np = isinstance(next_obs, np.ndarray)
        t = torch.as_tensor(next_obs)
        dev = t.device if t.is_cuda else torch.device('cuda', torch.cuda.current_device())
        t = t.to(device=dev, dtype=torch.float32).contiguous()
        out = torch.empty((t.shape[0], 1), dtype=torch.uint8, device=dev)
        lib = _lib.load()
        _lib.check(lib.mopo_termination_fn(self.term_kind, t.data_ptr(), t.shape[0], t.shape[1], out.data_ptr(),
                                           _lib.stream_ptr(dev)))
        out = out.bool()
        return out.cpu().numpy() if is_np else out

    def __repr__(self):
        return 'StaticFns(%r)' % self.domain


class _Registry(dict):
    """dict like the reference's ``static_fns`` (mopo/static/__init__.py:23-25)."""


static_fns = _Registry({d: StaticFns(d) for d in _KINDS})


class _Module(types.ModuleType):
    def __getitem__(self, k):
        return static_fns[k]

    def __contains__(self, k):
        return k in static_fns

    def keys(self):
        return static_fns.keys()


_m = sys.modules[__name__]
_m.__class__ = _Module
