"""NumPy restatement of the reference ensemble forward.  TEST INFRASTRUCTURE ONLY.

Follows (paths relative to the reference repo):
  mopo/models/constructor.py:30-36   architecture (4 x hidden swish + linear head, optional var head)
  mopo/models/fc.py:15-22,96-106     activations, einsum/matmul + bias
  mopo/models/fc.py:145-154          init law: trunc-normal(sigma=1/(2 sqrt(in))), bias 0
  mopo/models/utils.py:58-96         TensorStandardScaler (mu=0, sigma=1 init; fit; transform)
  mopo/models/bnn.py:185-199,210-213 head doubling in combined mode; max/min logvar init (+0.5 / -10)
  mopo/models/bnn.py:631-675         _compile_outputs (both separate_mean_var modes)
  mopo/models/bnn.py:337-344         elites (_model_inds) and random_inds
"""
import numpy as np


def _trunc_normal(rng, shape, std):
    """tf.truncated_normal_initializer: N(0, std) redrawn until |x| <= 2 std."""
    out = rng.standard_normal(size=shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * std).astype(np.float32)


def init_params(obs_dim, act_dim, rew_dim=1, hidden_dim=200, num_networks=7, n_hidden=4,
                separate_mean_var=True, seed=0, bias_std=0.0):
    """Random-init parameter dict with the reference's layout and init law.

    weights[l]: [E, in_l, out_l] float32, biases[l]: [E, 1, out_l] float32 (fc.py:145-154).
    In combined mode the last layer has 2*D1 outputs (bnn.py:185); in separate mode
    there is an extra ``var_weights/var_biases`` head fed by the last hidden activation.
    ``bias_std`` > 0 draws non-zero biases (not the init law; used to make tests sensitive to bias handling).
    """
    rng = np.random.RandomState(seed)
    E, D1, din = num_networks, obs_dim + rew_dim, obs_dim + act_dim
    dims = [din] + [hidden_dim] * n_hidden
    weights, biases = [], []
    for i in range(n_hidden):
        weights.append(_trunc_normal(rng, (E, dims[i], dims[i + 1]), 1.0 / (2.0 * np.sqrt(dims[i]))))
        biases.append((rng.standard_normal((E, 1, dims[i + 1])) * bias_std).astype(np.float32))
    out_last = D1 if separate_mean_var else 2 * D1
    weights.append(_trunc_normal(rng, (E, dims[-1], out_last), 1.0 / (2.0 * np.sqrt(dims[-1]))))
    biases.append((rng.standard_normal((E, 1, out_last)) * bias_std).astype(np.float32))
    p = {
        'obs_dim': obs_dim, 'act_dim': act_dim, 'rew_dim': rew_dim, 'hidden_dim': hidden_dim,
        'num_networks': E, 'separate_mean_var': bool(separate_mean_var),
        'weights': weights, 'biases': biases,
        'mu': np.zeros((1, din), np.float32), 'sigma': np.ones((1, din), np.float32),
        'max_logvar': np.full((1, D1), 0.5, np.float32), 'min_logvar': np.full((1, D1), -10.0, np.float32),
    }
    if separate_mean_var:
        p['var_weights'] = _trunc_normal(rng, (E, hidden_dim, D1), 1.0 / (2.0 * np.sqrt(hidden_dim)))
        p['var_biases'] = (rng.standard_normal((E, 1, D1)) * bias_std).astype(np.float32)
    return p


def fit_scaler(params, data):
    """TensorStandardScaler.fit (mopo/models/utils.py:69-86)."""
    mu = np.mean(data, axis=0, keepdims=True)
    sigma = np.std(data, axis=0, keepdims=True)
    sigma[sigma < 1e-12] = 1.0
    params['mu'] = mu.astype(np.float32)
    params['sigma'] = sigma.astype(np.float32)
    return params


def _swish(x):
    # fc.py:21  x * sigmoid(x)
    with np.errstate(over="ignore"):
        return x * (1.0 / (1.0 + np.exp(-x)))


def _softplus(x):
    return np.logaddexp(np.zeros((), x.dtype), x)


def forward(params, inputs, dtype=np.float32, ret_log_var=False):
    """BNN._compile_outputs (bnn.py:631-675) -> (mean[E,B,D1], var[E,B,D1]).

    inputs 2-D [B,in]: shared by all nets (fc.py:98-99); 3-D [E,B,in]: per-net inputs (fc.py:100-101).
    """
    dt = np.dtype(dtype)
    c = lambda a: np.asarray(a, dtype=dt)
    x = (c(inputs) - c(params['mu'])) / c(params['sigma'])          # utils.py:96
    W, b = params['weights'], params['biases']
    n_layers = len(W)
    cur = x
    hidden_last = None
    for l in range(n_layers):
        if cur.ndim == 2:
            raw = np.matmul(cur[None], c(W[l])) + c(b[l])              # fc.py:99 einsum('ij,ajk->aik') == broadcast matmul (BLAS)
        else:
            raw = np.matmul(cur, c(W[l])) + c(b[l])                  # fc.py:101
        if l < n_layers - 1:
            cur = _swish(raw)
            hidden_last = cur
        else:
            cur = raw                                                # end_act None (bnn.py:58)
    if params['separate_mean_var']:
        mean = cur                                                   # bnn.py:660-663
        raw_lv = np.matmul(hidden_last, c(params['var_weights'])) + c(params['var_biases'])  # bnn.py:665-667
    else:
        D1 = cur.shape[-1] // 2
        mean, raw_lv = cur[..., :D1], cur[..., D1:]                  # bnn.py:650-654
    max_lv, min_lv = c(params['max_logvar']), c(params['min_logvar'])
    logvar = max_lv - _softplus(max_lv - raw_lv)                     # bnn.py:654 / 669
    logvar = min_lv + _softplus(logvar - min_lv)                     # bnn.py:655 / 670
    if ret_log_var:
        return mean, logvar
    return mean, np.exp(logvar)                                      # bnn.py:675


class _Scaler:
    def __init__(self, params):
        self.cached_mu, self.cached_sigma = params['mu'], params['sigma']


class OracleModel:
    """Duck-typed stand-in for BNN as FakeEnv uses it (fake_env.py:50-61,77):
    predict / num_nets / random_inds / _model_inds / scaler."""

    def __init__(self, params, num_elites=5, model_inds=None, dtype=np.float32):
        self.params = params
        self.num_nets = params['num_networks']
        self.num_elites = num_elites
        self._model_inds = list(range(num_elites)) if model_inds is None else list(model_inds)
        self.scaler = _Scaler(params)
        self.dtype = dtype
        self.forced_inds = None   # test hook: inject elite picks

    def predict(self, inputs, factored=False):
        mean, var = forward(self.params, inputs, dtype=self.dtype)
        if inputs.ndim == 2 and not factored:                        # bnn.py:261-263
            m = mean.mean(0)
            return m, var.mean(0) + np.square(mean - m).mean(0)
        return mean, var

    def random_inds(self, batch_size):                               # bnn.py:342-344
        if self.forced_inds is not None:
            assert len(self.forced_inds) == batch_size
            return np.asarray(self.forced_inds)
        return np.random.choice(self._model_inds, size=batch_size)


def scale_weights(params, gain):
    """Multiply every weight tensor by ``gain`` (test helper: the init law makes activations
    shrink layer by layer; a gain of 2-4 gives O(1) activations so GEMM errors are visible)."""
    params['weights'] = [w * np.float32(gain) for w in params['weights']]
    if 'var_weights' in params:
        params['var_weights'] = params['var_weights'] * np.float32(gain)
    return params


def synthetic_case(domain, B, seed=0, gain=3.0, separate_mean_var=True, bias_std=0.1):
    """Seeded synthetic D4RL-shaped case shared by the golden generator, tests and bench
    (SURVEY.md 8(d)): params (reference init law x gain), start obs, per-step actions."""
    shapes = {'halfcheetah': (17, 6), 'hopper': (11, 3), 'walker2d': (17, 6), 'ant': (27, 8),
              'humanoid': (45, 17)}
    obs_dim, act_dim = shapes[domain]
    p = init_params(obs_dim, act_dim, seed=seed, separate_mean_var=separate_mean_var, bias_std=bias_std)
    scale_weights(p, gain)
    rng = np.random.RandomState(seed + 1)
    obs = rng.standard_normal((B, obs_dim)).astype(np.float32)
    if domain == 'hopper':
        obs *= 0.05; obs[:, 0] += 1.25
    elif domain == 'walker2d':
        obs *= 0.05; obs[:, 0] += 1.25
    elif domain == 'ant':
        obs *= 0.05; obs[:, 0] += 0.6
    elif domain == 'humanoid':
        obs *= 0.05; obs[:, 0] += 1.4
    act = np.tanh(rng.standard_normal((B, act_dim))).astype(np.float32)
    fit_scaler(p, np.concatenate([obs, act], -1))
    return p, obs, act
