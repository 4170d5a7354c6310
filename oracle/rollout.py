"""Literal restatement of MOPO._rollout_model (mopo/algorithms/mopo.py:723-765).
TEST INFRASTRUCTURE ONLY.  The policy call (mopo.py:734) is replaced by a caller-supplied
``action_fn(step, obs) -> act`` so actions can be precomputed / injected on both sides.
"""
import numpy as np


def rollout_model(fake_env, pool, start_obs, rollout_length, action_fn,
                  deterministic=False, noise_fn=None, inds_fn=None):
    """Returns ({'mean_rollout_length'}, steps_added).  noise_fn(step, B) -> [E,B,D1] and
    inds_fn(step, B) -> [B] inject randomness (None -> global np.random as the reference)."""
    obs = start_obs                                                   # mopo.py:727-728
    rollout_batch_size = len(start_obs)
    steps_added = []
    for i in range(rollout_length):                                   # :730
        act = action_fn(i, obs)                                       # :734
        kw = {}
        if noise_fn is not None:
            kw['noise'] = noise_fn(i, len(obs))
        if inds_fn is not None:
            kw['model_inds'] = inds_fn(i, len(obs))
        next_obs, rew, term, info = fake_env.step(obs, act, deterministic=deterministic, **kw)   # :747
        steps_added.append(len(obs))                                  # :748
        pool.add_samples({'observations': obs, 'actions': act, 'next_observations': next_obs,
                          'rewards': rew, 'terminals': term})         # :750-751
        nonterm_mask = ~term.squeeze(-1)                              # :753
        if nonterm_mask.sum() == 0:                                   # :754-756
            break
        obs = next_obs[nonterm_mask]                                  # :758
    return {'mean_rollout_length': sum(steps_added) / rollout_batch_size}, steps_added
