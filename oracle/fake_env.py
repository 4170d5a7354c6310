"""NumPy restatement of the reference's FakeEnv.step.  TEST INFRASTRUCTURE ONLY.

Follows mopo/models/fake_env.py:37-131 (step) and :20-35 (_get_logprob) statement by
statement; the only additions are the injectable ``noise`` ([E,B,D1], replaces the
np.random.normal draw at :72) and ``model_inds`` ([B], replaces model.random_inds at :77)
so that a Philox-based device implementation can be compared on identical randomness.
Pinned against the imported reference FakeEnv by tests/golden/make_golden.py.
"""
import numpy as np


class FakeEnv:
    def __init__(self, model, config, penalty_coeff=0., penalty_learned_var=False,
                 penalty_learned_var_random=False):
        self.model = model
        self.config = config
        self.penalty_coeff = penalty_coeff
        self.penalty_learned_var = penalty_learned_var
        self.penalty_learned_var_random = penalty_learned_var_random

    def _get_logprob(self, x, means, variances):                      # fake_env.py:20-35
        k = x.shape[-1]
        log_prob = -1 / 2 * (k * np.log(2 * np.pi) + np.log(variances).sum(-1)
                             + (np.power(x - means, 2) / variances).sum(-1))
        prob = np.exp(log_prob).sum(0)
        log_prob = np.log(prob)
        stds = np.std(means, 0).mean(-1)
        return log_prob, stds

    def step(self, obs, act, deterministic=False, noise=None, model_inds=None):
        assert len(obs.shape) == len(act.shape)                       # :38
        if len(obs.shape) == 1:                                       # :39-44
            obs, act, return_single = obs[None], act[None], True
        else:
            return_single = False
        inputs = np.concatenate((obs, act), axis=-1)                  # :46
        batch_length = inputs.shape[0]
        all_means, all_vars = [], []
        if self.model.num_nets * batch_length > 500000:               # :50-56 chunked predict
            group = int(500000 / self.model.num_nets)
            for i in range(int(np.ceil(batch_length / group))):
                m, v = self.model.predict(inputs[i * group:(i + 1) * group], factored=True)
                all_means.append(m); all_vars.append(v)
        else:
            m, v = self.model.predict(inputs[:], factored=True)
            all_means.append(m); all_vars.append(v)
        means = np.concatenate(all_means, axis=1)
        vars_ = np.concatenate(all_vars, axis=1)
        means[:, :, 1:] += obs                                        # :66  model predicts [rew, delta_obs]
        stds = np.sqrt(vars_)                                         # :67
        if deterministic:
            ensemble_samples = means                                  # :70
        else:
            if noise is None:
                noise = np.random.normal(size=means.shape)            # :72
            ensemble_samples = means + noise * stds
        if not deterministic:
            num_models, batch_size, _ = means.shape                   # :76-81
            if model_inds is None:
                model_inds = self.model.random_inds(batch_size)
            batch_inds = np.arange(0, batch_size)
            samples = ensemble_samples[model_inds, batch_inds]
            model_means = means[model_inds, batch_inds]
            model_stds = stds[model_inds, batch_inds]
        else:
            samples = np.mean(ensemble_samples, axis=0)               # :84-86
            model_means = np.mean(means, axis=0)
            model_stds = np.mean(stds, axis=0)
        log_prob, dev = self._get_logprob(samples, means, vars_)      # :88
        rewards, next_obs = samples[:, :1], samples[:, 1:]            # :90
        terminals = self.config.termination_fn(obs, act, next_obs)    # :91
        batch_size = model_means.shape[0]
        return_means = np.concatenate((model_means[:, :1], terminals, model_means[:, 1:]), axis=-1)   # :94
        return_stds = np.concatenate((model_stds[:, :1], np.zeros((batch_size, 1)), model_stds[:, 1:]), axis=-1)
        if self.penalty_coeff != 0:                                   # :97-115
            if not self.penalty_learned_var:
                ens_obs = means[:, :, 1:]
                diffs = ens_obs - np.mean(ens_obs, axis=0)
                dists = np.linalg.norm(diffs, axis=2)
                penalty = np.max(dists, axis=0)
            else:
                penalty = np.amax(np.linalg.norm(stds, axis=2), axis=0)
            penalty = np.expand_dims(penalty, 1)
            assert penalty.shape == rewards.shape
            unpenalized_rewards = rewards
            penalized_rewards = rewards - self.penalty_coeff * penalty
        else:
            penalty = None
            unpenalized_rewards = rewards
            penalized_rewards = rewards
        if return_single:                                             # :121-127
            next_obs, return_means, return_stds = next_obs[0], return_means[0], return_stds[0]
            unpenalized_rewards, penalized_rewards, terminals = unpenalized_rewards[0], penalized_rewards[0], terminals[0]
        info = {'mean': return_means, 'std': return_stds, 'log_prob': log_prob, 'dev': dev,
                'unpenalized_rewards': unpenalized_rewards, 'penalty': penalty,
                'penalized_rewards': penalized_rewards}
        return next_obs, penalized_rewards, terminals, info
