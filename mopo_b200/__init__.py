"""mopo_b200 -- B200-native implementation of MOPO's model-rollout hot path.

Drop-in surface (same names / argument meaning as xionghuichen/mopo):
  mopo_b200.models.constructor.construct_model   <- mopo/models/constructor.py:7-43
  mopo_b200.models.bnn.BNN                       <- mopo/models/bnn.py (predict / random_inds / save / load)
  mopo_b200.models.fake_env.FakeEnv              <- mopo/models/fake_env.py:37-131
  mopo_b200.static[domain]                       <- mopo/static/__init__.py
  mopo_b200.replay_pools.SimpleReplayPool        <- softlearning/replay_pools/simple_replay_pool.py
  mopo_b200.algorithms.mopo.ModelRollout         <- MOPO._rollout_model & co, mopo/algorithms/mopo.py:675-765
All arithmetic runs in hand-written sm_100a CUDA behind the C ABI in include/mopo_b200.h.
"""
from . import _lib  # noqa: F401  (does not load the .so until first use)

__version__ = '0.1.0'
