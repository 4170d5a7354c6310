"""NumPy restatement of the reference's per-environment static termination functions.
TEST INFRASTRUCTURE ONLY.  Follows mopo/static/<env>.py (line refs per function);
pinned against the imported reference functions by tests/golden/make_golden.py.
"""
import numpy as np


def _never(obs, act, next_obs):
    # halfcheetah.py:9-11 (also halfcheetahjump/vel/veljump, point2denv, point2dwallenv)
    return np.zeros((len(obs), 1), dtype=bool)


def _hopper(obs, act, next_obs):
    # hopper.py:9-18.  NB hopper.py:12 applies np.abs to the BOOLEAN (next_obs[:,1:] < 100):
    # there is no lower bound on next_obs[:,1:].
    height, angle = next_obs[:, 0], next_obs[:, 1]
    not_done = np.isfinite(next_obs).all(axis=-1) \
        * np.abs(next_obs[:, 1:] < 100).all(axis=-1) \
        * (height > .7) \
        * (np.abs(angle) < .2)
    return (~not_done)[:, None]


def _walker2d(obs, act, next_obs):
    # walker2d.py:9-17
    height, angle = next_obs[:, 0], next_obs[:, 1]
    not_done = (height > 0.8) * (height < 2.0) * (angle > -1.0) * (angle < 1.0)
    return (~not_done)[:, None]


def _ant(obs, act, next_obs):
    # ant.py:9-16, antangle.py (identical)
    x = next_obs[:, 0]
    not_done = np.isfinite(next_obs).all(axis=-1) * (x >= 0.2) * (x <= 1.0)
    return (~not_done)[:, None]


def _humanoid(obs, act, next_obs):
    # humanoid.py:10-14
    z = next_obs[:, 0]
    return ((z < 1.0) + (z > 2.0))[:, None]


TERMINATION_FNS = {
    'halfcheetah': _never, 'halfcheetahjump': _never, 'halfcheetahvel': _never,
    'halfcheetahveljump': _never, 'point2denv': _never, 'point2dwallenv': _never,
    'hopper': _hopper, 'walker2d': _walker2d, 'ant': _ant, 'antangle': _ant, 'humanoid': _humanoid,
    # pendulum.py:9 returns FLOAT zeros (breaks `~term` at mopo.py:753); treated as never-terminate.
    'pendulum': _never,
}


class StaticFns:
    def __init__(self, domain):
        self.domain = domain.lower()
        self._fn = TERMINATION_FNS[self.domain]

    def termination_fn(self, obs, act, next_obs):
        assert len(obs.shape) == len(next_obs.shape) == len(act.shape) == 2
        return self._fn(obs, act, next_obs)
