"""NumPy float64 restatement of the reference's ensemble TRAINING step.  TEST INFRASTRUCTURE ONLY.

Follows (paths relative to the reference repo):
  mopo/models/bnn.py:631-675   _compile_outputs: scaler, swish stack, head, the two softplus log-variance clamps
  mopo/models/bnn.py:677-700   _compile_losses: per net  mean_b mean_d[(mean - t)^2 exp(-logvar)] + mean_b mean_d[logvar];
                               mse_loss (inc_var_loss=False) = mean_b mean_d[(mean - t)^2]
  mopo/models/bnn.py:242-249   train_loss = sum_nets(losses) + sum(decays) + 0.01 sum(max_logvar) - 0.01 sum(min_logvar);
                               train_op = AdamOptimizer(lr 1e-3).minimize(train_loss, var_list = all W, b, max/min logvar)
  mopo/models/fc.py:155-156    decay_l = weight_decay_l * tf.nn.l2_loss(W_l) = weight_decay_l * sum(W_l^2) / 2
  mopo/models/constructor.py:30-36   weight decays 2.5e-5, 5e-5, 7.5e-5, 7.5e-5, 1e-4 (+ 1e-4 on the separate var head)
  mopo/models/bnn.py:369-503   train(): holdout split, bootstrap idxs [E, N], batches idxs[:, k*bs:(k+1)*bs], shuffle_rows,
                               _save_best (:309-329: a net's snapshot is renewed when its holdout mse improves by > 1 %)

The optimiser is TF1's tf.train.AdamOptimizer (a dependency, not vendored in the reference; TF 1.x is not installed
here), restated from its published update rule:
  lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t);  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;  p -= lr_t * m / (sqrt(v) + eps)
with b1 = 0.9, b2 = 0.999, eps = 1e-8.  PARITY UNPINNED against a TF run (TF cannot be run here): the analytic gradient
below is pinned instead by central finite differences of `loss` (tests/test_train.py), and `loss` restates bnn.py line
by line.  Parameters use the dict of oracle/ensemble.py (reference layout).
"""
import numpy as np

DECAYS_MLP = (0.000025, 0.00005, 0.000075, 0.000075, 0.0001)     # constructor.py:31-35
DECAY_VAR = 0.0001                                               # constructor.py:36
B1, B2, EPS = 0.9, 0.999, 1e-8


def _sig(x):
    return 1.0 / (1.0 + np.exp(-x))


def _softplus(x):
    return np.logaddexp(0.0, x)


def trainable(params):
    """The optimiser's var_list (bnn.py:212-224): per layer [W, b] (+ var head) + [max_logvar, min_logvar], as float64 copies."""
    v = {'weights': [np.array(w, np.float64) for w in params['weights']],
         'biases': [np.array(b, np.float64).reshape(w.shape[0], 1, -1) for w, b in zip(params['weights'], params['biases'])],
         'max_logvar': np.array(params['max_logvar'], np.float64).reshape(1, -1),
         'min_logvar': np.array(params['min_logvar'], np.float64).reshape(1, -1)}
    if params.get('separate_mean_var'):
        v['var_weights'] = np.array(params['var_weights'], np.float64)
        v['var_biases'] = np.array(params['var_biases'], np.float64).reshape(v['var_weights'].shape[0], 1, -1)
    return v


def _forward(v, xn):
    """xn [E, n, in] already scaled.  Returns mean, logvar and the caches the backward pass needs."""
    L = len(v['weights']) - 1
    hs, zs = [xn], []
    for l in range(L):
        z = hs[-1] @ v['weights'][l] + v['biases'][l]
        zs.append(z)
        hs.append(z * _sig(z))                                   # fc.py:21
    out = hs[-1] @ v['weights'][L] + v['biases'][L]
    if 'var_weights' in v:                                       # bnn.py:657-667
        mean = out
        raw = hs[-1] @ v['var_weights'] + v['var_biases']
    else:                                                        # bnn.py:646-649
        D1 = out.shape[-1] // 2
        mean, raw = out[..., :D1], out[..., D1:]
    l1 = v['max_logvar'] - _softplus(v['max_logvar'] - raw)      # bnn.py:669
    lv = v['min_logvar'] + _softplus(l1 - v['min_logvar'])       # bnn.py:670
    return mean, lv, (hs, zs, raw, l1)


def scale(params, x):
    """TensorStandardScaler.transform with the fitted (float32) mu / sigma (utils.py:88-96)."""
    return (np.asarray(x, np.float64) - np.asarray(params['mu'], np.float64)) / np.asarray(params['sigma'], np.float64)


def mse(params, v, x, t):
    """bnn.py:248 mse_loss: per-net mean squared error of the mean head.  x [E, n, in] raw inputs, t [E, n, D1]."""
    mean, _, _ = _forward(v, scale(params, x))
    return ((mean - t) ** 2).mean(-1).mean(-1)


def loss(params, v, x, t, decays=DECAYS_MLP, decay_var=DECAY_VAR, deterministic=False):
    """bnn.py:242-245 train_loss (scalar); deterministic (bnn.py:246-247): sum over nets of the plain MSE + the decays (TF
    minimises the [E] loss vector, i.e. its sum), no log-variance bound terms."""
    mean, lv, _ = _forward(v, scale(params, x))
    if deterministic:
        per_net = ((mean - t) ** 2).mean(-1).mean(-1)
    else:
        per_net = ((mean - t) ** 2 * np.exp(-lv)).mean(-1).mean(-1) + lv.mean(-1).mean(-1)
    tot = per_net.sum()
    for w, d in zip(v['weights'], decays):
        tot += d * 0.5 * (w ** 2).sum()
    if 'var_weights' in v:
        tot += decay_var * 0.5 * (v['var_weights'] ** 2).sum()
    if deterministic:
        return tot
    return tot + 0.01 * v['max_logvar'].sum() - 0.01 * v['min_logvar'].sum()


def grads(params, v, x, t, decays=DECAYS_MLP, decay_var=DECAY_VAR, deterministic=False):
    """Analytic gradient of `loss` w.r.t. every entry of `v` (same dict structure)."""
    xn = scale(params, x)
    mean, lv, (hs, zs, raw, l1) = _forward(v, xn)
    E, n, D1 = mean.shape
    L = len(v['weights']) - 1
    inv = np.exp(-lv)
    c = 1.0 / (n * D1)
    d_mean = 2.0 * (mean - t) * inv * c
    d_lv = (1.0 - (mean - t) ** 2 * inv) * c
    s_min = _sig(l1 - v['min_logvar'])                           # d lv / d l1
    s_max = _sig(v['max_logvar'] - raw)                          # d l1 / d raw
    d_raw = d_lv * s_min * s_max
    g = {'weights': [None] * (L + 1), 'biases': [None] * (L + 1)}
    g['max_logvar'] = (d_lv * s_min * (1.0 - s_max)).sum((0, 1)).reshape(1, -1) + 0.01
    g['min_logvar'] = (d_lv * (1.0 - s_min)).sum((0, 1)).reshape(1, -1) - 0.01
    if deterministic:                                            # the loss does not see the log-variance at all
        d_mean = 2.0 * (mean - t) * c
        d_raw = np.zeros_like(d_raw)
        g['max_logvar'], g['min_logvar'] = np.zeros_like(g['max_logvar']), np.zeros_like(g['min_logvar'])
    hT = np.swapaxes(hs[-1], 1, 2)
    if 'var_weights' in v:
        g['weights'][L] = hT @ d_mean + decays[L] * v['weights'][L]
        g['biases'][L] = d_mean.sum(1, keepdims=True)
        g['var_weights'] = hT @ d_raw + decay_var * v['var_weights']
        g['var_biases'] = d_raw.sum(1, keepdims=True)
        d_h = d_mean @ np.swapaxes(v['weights'][L], 1, 2) + d_raw @ np.swapaxes(v['var_weights'], 1, 2)
    else:
        d_out = np.concatenate([d_mean, d_raw], -1)
        g['weights'][L] = hT @ d_out + decays[L] * v['weights'][L]
        g['biases'][L] = d_out.sum(1, keepdims=True)
        d_h = d_out @ np.swapaxes(v['weights'][L], 1, 2)
    for l in range(L - 1, -1, -1):
        s = _sig(zs[l])
        d_z = d_h * (s + zs[l] * s * (1.0 - s))
        g['weights'][l] = np.swapaxes(hs[l], 1, 2) @ d_z + decays[l] * v['weights'][l]
        g['biases'][l] = d_z.sum(1, keepdims=True)
        if l > 0:
            d_h = d_z @ np.swapaxes(v['weights'][l], 1, 2)
    return g


def _leaves(d):
    keys = ['max_logvar', 'min_logvar'] + [k for k in ('var_weights', 'var_biases') if k in d]
    out = [(d, k) for k in keys]
    for k in ('weights', 'biases'):
        out += [(d[k], i) for i in range(len(d[k]))]
    return out


def adam_init(v):
    z = {k: ([np.zeros_like(a) for a in val] if isinstance(val, list) else np.zeros_like(val)) for k, val in v.items()}
    return {'m': z, 'v': {k: ([a.copy() for a in val] if isinstance(val, list) else val.copy()) for k, val in z.items()}, 't': 0}


def adam_step(v, g, st, lr=1e-3):
    """tf.train.AdamOptimizer._apply_dense, in place on `v`."""
    st['t'] += 1
    t = st['t']
    lr_t = lr * np.sqrt(1.0 - B2 ** t) / (1.0 - B1 ** t)
    for (pc, pk), (gc, gk), (mc, mk), (vc, vk) in zip(_leaves(v), _leaves(g), _leaves(st['m']), _leaves(st['v'])):
        mc[mk] = B1 * mc[mk] + (1.0 - B1) * gc[gk]
        vc[vk] = B2 * vc[vk] + (1.0 - B2) * gc[gk] ** 2
        pc[pk] = pc[pk] - lr_t * mc[mk] / (np.sqrt(vc[vk]) + EPS)


def train_batches(params, v, st, inputs, targets, idxs, batch_size, lr=1e-3, decays=DECAYS_MLP, decay_var=DECAY_VAR,
                  deterministic=False):
    """One epoch's inner loop (bnn.py:421-427): batch k trains net e on rows idxs[e, k*bs:(k+1)*bs].  Returns the
    train_loss of every batch evaluated BEFORE its update."""
    losses = []
    for k in range(int(np.ceil(idxs.shape[-1] / batch_size))):
        bi = idxs[:, k * batch_size:(k + 1) * batch_size]
        x, t = np.asarray(inputs, np.float64)[bi], np.asarray(targets, np.float64)[bi]
        losses.append(loss(params, v, x, t, decays, decay_var, deterministic))
        adam_step(v, grads(params, v, x, t, decays, decay_var, deterministic), st, lr)
    return np.array(losses)


def save_best(snapshots, epoch, holdout_losses):
    """bnn.py:309-329 without the state copy: returns (nets whose snapshot is renewed, updated flag)."""
    renewed = []
    for i, current in enumerate(holdout_losses):
        _, best = snapshots[i]
        if (best - current) / best > 0.01:
            snapshots[i] = (epoch, current)
            renewed.append(i)
    return renewed
