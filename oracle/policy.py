"""TEST INFRASTRUCTURE ONLY (imported by tests/ and __graft_entry__.smoke()).

NumPy restatement of the SAC actor the rollout queries: mlp_gaussian_policy + apply_squashing_func
(reference mopo/algorithms/mopo.py:275-312, hidden_sizes (256, 256) with ReLU, simple_run/base.py:62-66) as read out by
MOPO.get_action_meta (mopo.py:468-485).  The reference graph is TensorFlow 1.14 (absent here): parity unpinned at the TF
boundary, the ops are dense + relu + clip + exp + tanh restated in float64 ("truth") or float32 ("TF-like").
"""
import numpy as np

LOG_STD_MAX, LOG_STD_MIN = 2.0, -20.0          # mopo.py:271-272


def init_params(obs_dim, act_dim, hidden=256, seed=0, dtype=np.float32):
    """tf.layers.dense defaults: glorot_uniform kernels, zero biases (mopo.py:275-280 passes kernel_initializer=None)."""
    rng = np.random.RandomState(seed)
    shapes = [(obs_dim, hidden), (hidden, hidden), (hidden, act_dim), (hidden, act_dim)]
    weights, biases = [], []
    for fan_in, fan_out in shapes:
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        weights.append(rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(dtype))
        biases.append((0.1 * rng.standard_normal(fan_out)).astype(dtype))     # non-zero so that the bias path is exercised
    return weights, biases


def forward(weights, biases, obs, noise=None, deterministic=False, dtype=np.float64):
    """-> (action, mu, std): action = tanh(mu + noise * std), or tanh(mu) when deterministic (mopo.py:296-303, 287-289)."""
    W = [np.asarray(w, dtype) for w in weights]
    b = [np.asarray(x, dtype) for x in biases]
    x = np.asarray(obs, dtype)
    h = np.maximum(x @ W[0] + b[0], 0)
    h = np.maximum(h @ W[1] + b[1], 0)
    mu = h @ W[2] + b[2]
    log_std = np.clip(h @ W[3] + b[3], LOG_STD_MIN, LOG_STD_MAX)
    std = np.exp(log_std)
    if deterministic:
        return np.tanh(mu), mu, std
    pi = mu + np.asarray(noise, dtype) * std
    return np.tanh(pi), mu, std
