"""NumPy float64 restatement of MOPO's SAC update step.  TEST INFRASTRUCTURE ONLY.

Follows (paths relative to the reference repo):
  mopo/algorithms/mopo.py:271-283   mlp (tf.layers.dense stack), gaussian_likelihood (EPS = 1e-8)
  mopo/algorithms/mopo.py:286-309   apply_squashing_func, mlp_gaussian_policy (log_std clipped to [-20, 2])
  mopo/algorithms/mopo.py:311-325   mlp_actor_critic: pi, q1, q2 = mlp(concat(obs, act), [256, 256, 1]) (ReLU)
  mopo/algorithms/mopo.py:339-352   the four graphs: main(s, a), main(s, pi(s)), pi(s'), target(s', pi(s'))
  mopo/algorithms/mopo.py:359-394   alpha = exp(log_alpha); policy loss mean(alpha logp_pi - min(q1_pi, q2_pi));
                                    q_target = reward_scale r + discount (1 - done)(min(q1', q2') - alpha logp_pi')
  mopo/algorithms/mopo.py:411-445   q_i loss = 0.5 mean (q_i - stop_grad(q_target))^2; Adam(Q_lr) per critic, Adam(policy_lr)
                                    for pi, alpha loss -mean(log_alpha (logp_pi + target_entropy)) with Adam(policy_lr)
  mopo/algorithms/mopo.py:448-452   Polyak: target <- (1 - tau) target + tau main; target_init copies
  mopo/algorithms/mopo.py:801-821   _training_batch: int(batch_size * real_ratio) env rows then the model rows

All four train ops run in ONE session.run on the same feed (mopo.py:455-457, 838); the control dependencies there only
order the variable writes, so every gradient is restated from the PRE-update variables.  The optimiser is TF1 Adam
(oracle/train.py).  PARITY UNPINNED against a TF run (TF is not installed here); the analytic gradients below are pinned
by central finite differences of `losses` (tests/test_sac.py).
"""
import numpy as np

LOG_STD_MAX, LOG_STD_MIN, EPS = 2.0, -20.0, 1e-8
B1, B2, ADAM_EPS = 0.9, 0.999, 1e-8


def _glorot(rng, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))                      # tf.layers.dense default kernel_initializer
    return rng.uniform(-lim, lim, size=(fan_in, fan_out))


def init_state(obs_dim, act_dim, hidden=256, seed=0, bias_std=0.0):
    """Variables of scopes main/{pi,q1,q2}, target/{q1,q2} and log_alpha (initializer 0.0, mopo.py:359-363)."""
    rng = np.random.RandomState(seed)

    def mlp(din, dout):
        dims = [din, hidden, hidden]
        W = [_glorot(rng, dims[i], dims[i + 1]) for i in range(2)] + [_glorot(rng, hidden, dout)]
        b = [rng.standard_normal(w.shape[1]) * bias_std for w in W]
        return {'W': W, 'b': b}

    pi = mlp(obs_dim, act_dim)
    pi['W'].append(_glorot(rng, hidden, act_dim)); pi['b'].append(rng.standard_normal(act_dim) * bias_std)   # mu, log_std heads
    st = {'pi': pi, 'q1': mlp(obs_dim + act_dim, 1), 'q2': mlp(obs_dim + act_dim, 1), 'log_alpha': np.zeros(())}
    for k in ('q1', 'q2'):                                          # target_init (mopo.py:451-452)
        st[k + '_targ'] = {'W': [w.copy() for w in st[k]['W']], 'b': [b.copy() for b in st[k]['b']]}
    return st


def pi_forward(P, obs, eps):
    h0 = np.maximum(obs @ P['W'][0] + P['b'][0], 0.0)
    h1 = np.maximum(h0 @ P['W'][1] + P['b'][1], 0.0)
    mu = h1 @ P['W'][2] + P['b'][2]
    ls_raw = h1 @ P['W'][3] + P['b'][3]
    ls = np.clip(ls_raw, LOG_STD_MIN, LOG_STD_MAX)
    std = np.exp(ls)
    u = mu + eps * std
    logp = (-0.5 * (((u - mu) / (std + EPS)) ** 2 + 2 * ls + np.log(2 * np.pi))).sum(-1)
    logp = logp - (2 * (np.log(2.0) - u - np.logaddexp(0.0, -2 * u))).sum(-1)
    return np.tanh(u), logp, (h0, h1, mu, ls_raw, ls, std, u)


def q_forward(Q, obs, act):
    x = np.concatenate([obs, act], -1)
    h0 = np.maximum(x @ Q['W'][0] + Q['b'][0], 0.0)
    h1 = np.maximum(h0 @ Q['W'][1] + Q['b'][1], 0.0)
    return (h1 @ Q['W'][2] + Q['b'][2])[:, 0], (x, h0, h1)


def losses(st, batch, eps_pi, eps_next, hp):
    """(q1_loss, q2_loss, policy_loss, alpha_loss) and the intermediates the gradients need."""
    alpha = np.exp(st['log_alpha'])
    a_pi, logp_pi, c_pi = pi_forward(st['pi'], batch['observations'], eps_pi)
    a_nx, logp_nx, _ = pi_forward(st['pi'], batch['next_observations'], eps_next)
    q1, c_q1 = q_forward(st['q1'], batch['observations'], batch['actions'])
    q2, c_q2 = q_forward(st['q2'], batch['observations'], batch['actions'])
    q1_pi, c_q1pi = q_forward(st['q1'], batch['observations'], a_pi)
    q2_pi, c_q2pi = q_forward(st['q2'], batch['observations'], a_pi)
    q1_t, _ = q_forward(st['q1_targ'], batch['next_observations'], a_nx)
    q2_t, _ = q_forward(st['q2_targ'], batch['next_observations'], a_nx)
    next_value = np.minimum(q1_t, q2_t) - alpha * logp_nx
    y = hp['reward_scale'] * batch['rewards'][:, 0] + hp['discount'] * (1.0 - batch['terminals'][:, 0].astype(np.float64)) * next_value
    out = {'q1_loss': 0.5 * ((q1 - y) ** 2).mean(), 'q2_loss': 0.5 * ((q2 - y) ** 2).mean(),
           'policy_loss': (alpha * logp_pi - np.minimum(q1_pi, q2_pi)).mean(),
           'alpha_loss': -(st['log_alpha'] * (logp_pi + hp['target_entropy'])).mean()}
    cache = dict(alpha=alpha, a_pi=a_pi, logp_pi=logp_pi, c_pi=c_pi, q1=q1, q2=q2, c_q1=c_q1, c_q2=c_q2, q1_pi=q1_pi, q2_pi=q2_pi,
                 c_q1pi=c_q1pi, c_q2pi=c_q2pi, y=y)
    return out, cache


def _q_backward(Q, cache, dq):
    """dq [B] -> (grads of W / b, d input [B, obs + act])."""
    x, h0, h1 = cache
    dz2 = dq[:, None]
    gW2, gb2 = h1.T @ dz2, dz2.sum(0)
    dz1 = (dz2 @ Q['W'][2].T) * (h1 > 0)
    gW1, gb1 = h0.T @ dz1, dz1.sum(0)
    dz0 = (dz1 @ Q['W'][1].T) * (h0 > 0)
    gW0, gb0 = x.T @ dz0, dz0.sum(0)
    return {'W': [gW0, gW1, gW2], 'b': [gb0, gb1, gb2]}, dz0 @ Q['W'][0].T


def grads(st, batch, eps_pi, hp, cache):
    """d q1_loss / d q1 vars, d q2_loss / d q2 vars, d policy_loss / d pi vars, d alpha_loss / d log_alpha."""
    obs = batch['observations']
    n = obs.shape[0]
    od = obs.shape[1]
    g = {}
    g['q1'], _ = _q_backward(st['q1'], cache['c_q1'], (cache['q1'] - cache['y']) / n)
    g['q2'], _ = _q_backward(st['q2'], cache['c_q2'], (cache['q2'] - cache['y']) / n)
    first = cache['q1_pi'] <= cache['q2_pi']                       # tf.minimum: ties send the gradient to the first argument
    _, dx1 = _q_backward(st['q1'], cache['c_q1pi'], np.where(first, -1.0 / n, 0.0))
    _, dx2 = _q_backward(st['q2'], cache['c_q2pi'], np.where(first, 0.0, -1.0 / n))
    d_a = (dx1 + dx2)[:, od:]                                       # d policy_loss / d a_pi through -min q
    h0, h1, mu, ls_raw, ls, std, u = cache['c_pi']
    a = cache['a_pi']
    alpha = cache['alpha']
    r2 = (u - mu) / (std + EPS) ** 2
    g_u = alpha / n * (-r2 + 2.0 * a) + d_a * (1.0 - a * a)         # total d L / d u
    d_mu = g_u + alpha / n * r2
    d_ls = g_u * (eps_pi * std) + alpha / n * ((u - mu) ** 2 * std / (std + EPS) ** 3 - 1.0)
    d_ls = d_ls * ((ls_raw >= LOG_STD_MIN) & (ls_raw <= LOG_STD_MAX))
    P = st['pi']
    gW2, gb2, gW3, gb3 = h1.T @ d_mu, d_mu.sum(0), h1.T @ d_ls, d_ls.sum(0)
    dz1 = (d_mu @ P['W'][2].T + d_ls @ P['W'][3].T) * (h1 > 0)
    gW1, gb1 = h0.T @ dz1, dz1.sum(0)
    dz0 = (dz1 @ P['W'][1].T) * (h0 > 0)
    g['pi'] = {'W': [obs.T @ dz0, gW1, gW2, gW3], 'b': [dz0.sum(0), gb1, gb2, gb3]}
    g['log_alpha'] = -(cache['logp_pi'] + hp['target_entropy']).mean()
    return g


def adam_init(st):
    z = lambda net: {'W': [np.zeros_like(w) for w in net['W']], 'b': [np.zeros_like(b) for b in net['b']]}
    return {'m': {'pi': z(st['pi']), 'q1': z(st['q1']), 'q2': z(st['q2']), 'log_alpha': np.zeros(())},
            'v': {'pi': z(st['pi']), 'q1': z(st['q1']), 'q2': z(st['q2']), 'log_alpha': np.zeros(())}, 't': 0}


def update(st, opt, batch, eps_pi, eps_next, hp):
    """One _do_training (mopo.py:829-843): the four Adam steps from pre-update gradients, then the Polyak update."""
    out, cache = losses(st, batch, eps_pi, eps_next, hp)
    g = grads(st, batch, eps_pi, hp, cache)
    opt['t'] += 1
    t = opt['t']

    def step(p, gr, m, v, lr):
        lr_t = lr * np.sqrt(1.0 - B2 ** t) / (1.0 - B1 ** t)
        m[...] = B1 * m + (1.0 - B1) * gr
        v[...] = B2 * v + (1.0 - B2) * gr ** 2
        p[...] = p - lr_t * m / (np.sqrt(v) + ADAM_EPS)

    for net, lr in (('pi', hp['policy_lr']), ('q1', hp['Q_lr']), ('q2', hp['Q_lr'])):
        for k in ('W', 'b'):
            for i in range(len(st[net][k])):
                step(st[net][k][i], g[net][k][i], opt['m'][net][k][i], opt['v'][net][k][i], lr)
    la = np.array(st['log_alpha'], np.float64)
    step(la, g['log_alpha'], opt['m']['log_alpha'], opt['v']['log_alpha'], hp['policy_lr'])
    st['log_alpha'] = la
    for k in ('q1', 'q2'):                                          # mopo.py:448-449
        for kk in ('W', 'b'):
            for i in range(len(st[k][kk])):
                st[k + '_targ'][kk][i] = (1.0 - hp['tau']) * st[k + '_targ'][kk][i] + hp['tau'] * st[k][kk][i]
    return out


DEFAULT_HP = {'reward_scale': 1.0, 'discount': 0.99, 'tau': 5e-3, 'policy_lr': 3e-4, 'Q_lr': 3e-4, 'target_entropy': -6.0}
