"""Ensemble training (reference mopo/models/bnn.py:369-503).

CPU: the oracle's analytic gradient against central finite differences of the loss restated from bnn.py:677-700,
242-244; Adam known answers; the _save_best rule.  GPU: the device optimiser loop (mopo_ens_train_*) against the
float64 oracle batch by batch, mse_loss, snapshot / restore, and BNN.train end to end."""
import ctypes as C
import numpy as np
import pytest

from oracle import ensemble as ens
from oracle import train as otr


def _problem(obs_dim, act_dim, H, E, sep, n_rows, seed=0, bias_std=0.1):
    p = ens.init_params(obs_dim, act_dim, 1, H, E, 4, separate_mean_var=sep, seed=seed, bias_std=bias_std)
    rng = np.random.RandomState(seed + 1)
    inputs = (rng.standard_normal((n_rows, obs_dim + act_dim)) * 2 + 0.5).astype(np.float32)
    A = rng.standard_normal((obs_dim + act_dim, obs_dim + 1)) * 0.3
    targets = (np.tanh(inputs @ A) + 0.05 * rng.standard_normal((n_rows, obs_dim + 1))).astype(np.float32)
    ens.fit_scaler(p, inputs)
    return p, inputs, targets, rng


@pytest.mark.parametrize('sep', [True, False])
def test_oracle_gradient_matches_finite_differences(sep):
    p, inputs, targets, rng = _problem(5, 2, 16, 3, sep, 50)
    bi = rng.randint(50, size=(3, 12))
    x, t = inputs[bi].astype(np.float64), targets[bi].astype(np.float64)
    v = otr.trainable(p)
    v['max_logvar'] += rng.standard_normal(v['max_logvar'].shape) * 0.3
    v['min_logvar'] += rng.standard_normal(v['min_logvar'].shape) * 0.3
    g = otr.grads(p, v, x, t)
    for (vc, vk), (gc, gk) in zip(otr._leaves(v), otr._leaves(g)):
        a = vc[vk]
        assert gc[gk].shape == a.shape
        for _ in range(6):
            i = tuple(rng.randint(s) for s in a.shape)
            old, h = a[i], 1e-5
            a[i] = old + h; lp = otr.loss(p, v, x, t)
            a[i] = old - h; lm = otr.loss(p, v, x, t)
            a[i] = old
            fd = (lp - lm) / (2 * h)
            assert abs(fd - gc[gk][i]) <= 1e-4 * abs(fd) + 1e-8, (vk, i, fd, gc[gk][i])


def test_oracle_loss_known_answer():
    """Zero weights: mean = 0, raw = 0 -> logvar = min + softplus(max - softplus(max) - min); loss in closed form."""
    p = ens.init_params(3, 2, 1, 8, 2, 4, separate_mean_var=True, seed=0)
    for k in ('weights', 'biases'):
        p[k] = [np.zeros_like(a) for a in p[k]]
    p['var_weights'] = np.zeros_like(p['var_weights']); p['var_biases'] = np.zeros_like(p['var_biases'])
    v = otr.trainable(p)
    x = np.random.RandomState(0).standard_normal((2, 7, 5)); t = np.full((2, 7, 4), 2.0)
    l1 = 0.5 - np.log1p(np.exp(0.5))
    lv = -10.0 + np.log1p(np.exp(l1 + 10.0))
    want = 2 * (4.0 * np.exp(-lv) + lv) + 0.01 * 4 * 0.5 + 0.01 * 4 * 10.0
    np.testing.assert_allclose(otr.loss(p, v, x, t), want, rtol=1e-12)
    np.testing.assert_allclose(otr.mse(p, v, x, t), [4.0, 4.0])


def test_oracle_adam_first_step_is_lr_sign():
    p, inputs, targets, rng = _problem(4, 2, 8, 2, True, 30)
    v = otr.trainable(p)
    before = [vc[vk].copy() for vc, vk in otr._leaves(v)]
    bi = rng.randint(30, size=(2, 10))
    g = otr.grads(p, v, inputs[bi].astype(np.float64), targets[bi].astype(np.float64))
    st = otr.adam_init(v)
    otr.adam_step(v, g, st, lr=1e-3)
    assert st['t'] == 1
    for b0, (vc, vk), (gc, gk) in zip(before, otr._leaves(v), otr._leaves(g)):
        big = np.abs(gc[gk]) > 1e-3          # eps / (sqrt(1 - b2) |g|) < 4e-4
        np.testing.assert_allclose((vc[vk] - b0)[big], -1e-3 * np.sign(gc[gk][big]), rtol=2e-3)


def test_oracle_save_best_rule():
    snaps = {i: (None, 1e10) for i in range(3)}
    assert otr.save_best(snaps, 0, [1.0, 2.0, 3.0]) == [0, 1, 2]
    assert otr.save_best(snaps, 1, [0.995, 1.97, 3.5]) == [1]          # 0.5 % no, 1.5 % yes, worse no
    assert snaps[0] == (0, 1.0) and snaps[1] == (1, 1.97) and snaps[2] == (0, 3.0)


# ------------------------------------------------------------------------------------------------------------------
def _handle(p, device=0):
    from mopo_b200 import _lib
    lib = _lib.load()
    n_hidden = len(p['weights']) - 1
    cfg = _lib.EnsCfg(p['obs_dim'], p['act_dim'], p['rew_dim'], p['hidden_dim'], n_hidden, p['num_networks'],
                      min(5, p['num_networks']), int(p['separate_mean_var']), device)
    h = C.c_void_p()
    _lib.check(lib.mopo_ens_create(C.byref(cfg), C.byref(h)))
    keep = [np.ascontiguousarray(a, np.float32) for a in p['weights'] + p['biases']]
    n = n_hidden + 1
    W = (C.c_void_p * n)(*[a.ctypes.data for a in keep[:n]])
    b = (C.c_void_p * n)(*[a.ctypes.data for a in keep[n:]])
    vw = np.ascontiguousarray(p['var_weights'], np.float32) if p['separate_mean_var'] else None
    vb = np.ascontiguousarray(p['var_biases'], np.float32) if p['separate_mean_var'] else None
    _lib.check(lib.mopo_ens_set_params(h, W, b, _lib.ptr(vw), _lib.ptr(vb), p['mu'].ctypes.data, p['sigma'].ctypes.data,
                                       p['max_logvar'].ctypes.data, p['min_logvar'].ctypes.data, 0, None))
    return lib, h


def _get_params(lib, h, p):
    from mopo_b200 import _lib
    n = len(p['weights'])
    ws = [np.empty_like(np.asarray(w, np.float32)) for w in p['weights']]
    bs = [np.empty((w.shape[0], 1, w.shape[2]), np.float32) for w in p['weights']]
    W = (C.c_void_p * n)(*[a.ctypes.data for a in ws])
    b = (C.c_void_p * n)(*[a.ctypes.data for a in bs])
    vw = np.empty_like(p['var_weights']) if p['separate_mean_var'] else None
    vb = np.empty((vw.shape[0], 1, vw.shape[2]), np.float32) if p['separate_mean_var'] else None
    mx, mn = np.empty_like(p['max_logvar']), np.empty_like(p['min_logvar'])
    _lib.check(lib.mopo_ens_get_params(h, W, b, _lib.ptr(vw), _lib.ptr(vb), mx.ctypes.data, mn.ctypes.data, None))
    out = {'weights': ws, 'biases': bs, 'max_logvar': mx, 'min_logvar': mn}
    if vw is not None:
        out['var_weights'], out['var_biases'] = vw, vb
    return out


def _decays(p):
    n_hidden = len(p['weights']) - 1
    d = list(otr.DECAYS_MLP[:n_hidden]) + [otr.DECAYS_MLP[-1], otr.DECAY_VAR]
    return (C.c_float * len(d))(*d)


def _compare_params(got, v, atol):
    worst = 0.0
    for (gc, gk), (vc, vk) in zip(otr._leaves(got), otr._leaves(v)):
        worst = max(worst, float(np.abs(gc[gk].reshape(vc[vk].shape) - vc[vk]).max()))
    assert worst <= atol, worst
    return worst


@pytest.mark.gpu
@pytest.mark.parametrize('sep', [True, False])
@pytest.mark.parametrize('shape', [(5, 2, 32, 3, 16, 40), (17, 6, 200, 7, 256, 600), (11, 3, 200, 7, 100, 250), (6, 3, 30, 2, 10, 25)])   # last: odd pitches -> scalar-load fallbacks
def test_train_batches_match_oracle(shape, sep):
    """mopo_ens_train_batches == the float64 oracle: the train_loss of every batch and the parameters after the batches
    (ragged last batch included)."""
    import torch
    from mopo_b200 import _lib
    od, ad, H, E, bs, n_cols = shape
    p, inputs, targets, rng = _problem(od, ad, H, E, sep, 500, seed=od)
    idx = rng.randint(500, size=(E, n_cols)).astype(np.int32)
    lib, h = _handle(p)
    d_in, d_tg, d_idx = (torch.as_tensor(a).cuda() for a in (inputs, targets, idx))
    nb = int(np.ceil(n_cols / bs))
    d_loss = torch.zeros(nb, device='cuda')
    _lib.check(lib.mopo_ens_train_begin(h, 1e-3, _decays(p)))
    _lib.check(lib.mopo_ens_train_batches(h, d_in.data_ptr(), d_tg.data_ptr(), 500, d_idx.data_ptr(), n_cols, n_cols, bs,
                                          d_loss.data_ptr(), None))
    torch.cuda.synchronize()
    got = _get_params(lib, h, p)
    v = otr.trainable(p)
    st = otr.adam_init(v)
    want_loss = otr.train_batches(p, v, st, inputs, targets, idx, bs)
    np.testing.assert_allclose(d_loss.cpu().numpy(), want_loss, rtol=2e-4)
    # every Adam step moves an entry by ~lr = 1e-3; fp32 vs float64 gradients agree to ~1e-6 relative, so the parameters
    # agree far below one step except where a gradient is within fp32 noise of zero (none at these seeds)
    _compare_params(got, v, 2e-5 * nb)
    _lib.check(lib.mopo_ens_train_end(h, None))
    lib.mopo_ens_destroy(h)


@pytest.mark.gpu
@pytest.mark.parametrize('sep', [True, False])
def test_train_deterministic_loss_and_persistent_adam_state(sep):
    """(1) deterministic=True (bnn.py:246-247): the loss is MSE + decays, the log-variance half and max/min_logvar get no
    gradient.  (2) The optimiser belongs to the model, not to a train() call (one AdamOptimizer built in finalize, bnn.py:180,
    252): a second begin/batches/end run continues with the first run's moments and step count -- the oracle is run as ONE
    sequence over both runs' batches."""
    import torch
    from mopo_b200 import _lib
    od, ad, H, E, bs, n_cols = 11, 3, 200, 7, 64, 192
    p, inputs, targets, rng = _problem(od, ad, H, E, sep, 500, seed=3)
    idx = rng.randint(500, size=(E, 2 * n_cols)).astype(np.int32)
    lib, h = _handle(p)
    d_in, d_tg = torch.as_tensor(inputs).cuda(), torch.as_tensor(targets).cuda()
    nb = n_cols // bs
    got_loss = []
    for run, det in ((0, 1), (1, 0)):                       # a deterministic run, then a probabilistic one on the same optimiser
        d_idx = torch.as_tensor(np.ascontiguousarray(idx[:, run * n_cols:(run + 1) * n_cols])).cuda()
        d_loss = torch.zeros(nb, device='cuda')
        _lib.check(lib.mopo_ens_train_begin(h, 1e-3, _decays(p)))
        _lib.check(lib.mopo_ens_train_set_deterministic(h, det))
        _lib.check(lib.mopo_ens_train_batches(h, d_in.data_ptr(), d_tg.data_ptr(), 500, d_idx.data_ptr(), n_cols, n_cols, bs,
                                              d_loss.data_ptr(), None))
        _lib.check(lib.mopo_ens_train_end(h, None))
        got_loss.append(d_loss.cpu().numpy())
    got = _get_params(lib, h, p)
    v = otr.trainable(p)
    mx0, mn0 = v['max_logvar'].copy(), v['min_logvar'].copy()
    st = otr.adam_init(v)
    want0 = otr.train_batches(p, v, st, inputs, targets, idx[:, :n_cols], bs, deterministic=True)
    assert np.array_equal(v['max_logvar'], mx0) and np.array_equal(v['min_logvar'], mn0)     # untouched by the deterministic loss
    want1 = otr.train_batches(p, v, st, inputs, targets, idx[:, n_cols:], bs)                  # same Adam state: t continues at nb + 1
    np.testing.assert_allclose(got_loss[0], want0, rtol=2e-4)
    np.testing.assert_allclose(got_loss[1], want1, rtol=2e-4)
    _compare_params(got, v, 2e-5 * 2 * nb)
    # a fresh optimiser gives different parameters: the state did carry over
    _lib.check(lib.mopo_ens_train_reset(h))
    lib.mopo_ens_destroy(h)


@pytest.mark.gpu
def test_train_mse_snapshot_restore():
    import torch
    from mopo_b200 import _lib
    p, inputs, targets, rng = _problem(11, 3, 64, 4, True, 400, seed=3)
    lib, h = _handle(p)
    d_in, d_tg = torch.as_tensor(inputs).cuda(), torch.as_tensor(targets).cuda()
    idx = rng.randint(400, size=(4, 128)).astype(np.int32)
    d_idx = torch.as_tensor(idx).cuda()
    mse = torch.zeros(4, device='cuda')
    with pytest.raises(RuntimeError):
        _lib.check(lib.mopo_ens_train_mse(h, d_in.data_ptr(), d_tg.data_ptr(), None, 0, 100, mse.data_ptr(), None))   # before begin
    _lib.check(lib.mopo_ens_train_begin(h, 1e-3, _decays(p)))
    v = otr.trainable(p)
    # shared rows (the tiled holdout set) and per-net rows
    _lib.check(lib.mopo_ens_train_mse(h, d_in.data_ptr(), d_tg.data_ptr(), None, 0, 100, mse.data_ptr(), None))
    x = np.tile(inputs[None, :100], [4, 1, 1]).astype(np.float64); t = np.tile(targets[None, :100], [4, 1, 1]).astype(np.float64)
    np.testing.assert_allclose(mse.cpu().numpy(), otr.mse(p, v, x, t), rtol=1e-4)
    _lib.check(lib.mopo_ens_train_mse(h, d_in.data_ptr(), d_tg.data_ptr(), d_idx.data_ptr(), 128, 77, mse.data_ptr(), None))
    np.testing.assert_allclose(mse.cpu().numpy(), otr.mse(p, v, inputs[idx[:, :77]].astype(np.float64),
                                                          targets[idx[:, :77]].astype(np.float64)), rtol=1e-4)
    # restore needs a saved state for every net (bnn.py:298 KeyError)
    _lib.check(lib.mopo_ens_train_snapshot(h, 1, None))
    with pytest.raises(RuntimeError):
        _lib.check(lib.mopo_ens_train_restore(h, None))
    with pytest.raises(ValueError):
        _lib.check(lib.mopo_ens_train_snapshot(h, 4, None))
    for e in (0, 2, 3):
        _lib.check(lib.mopo_ens_train_snapshot(h, e, None))
    snap0 = _get_params(lib, h, p)
    _lib.check(lib.mopo_ens_train_batches(h, d_in.data_ptr(), d_tg.data_ptr(), 400, d_idx.data_ptr(), 128, 128, 32, None, None))
    _lib.check(lib.mopo_ens_train_snapshot(h, 2, None))                # net 2 keeps its newer weights
    after = _get_params(lib, h, p)
    assert not np.array_equal(after['weights'][0], snap0['weights'][0])
    _lib.check(lib.mopo_ens_train_restore(h, None))
    got = _get_params(lib, h, p)
    for k in ('weights', 'biases'):
        for a, s0, s1 in zip(got[k], snap0[k], after[k]):
            for e in range(4):
                assert np.array_equal(a[e], (s1 if e == 2 else s0)[e])
    assert np.array_equal(got['var_weights'][2], after['var_weights'][2]) and np.array_equal(got['var_weights'][0], snap0['var_weights'][0])
    # max / min logvar are not part of the saved state (bnn.py:287-290): they keep the trained values
    assert np.array_equal(got['max_logvar'], after['max_logvar']) and not np.array_equal(got['max_logvar'], snap0['max_logvar'])
    _lib.check(lib.mopo_ens_train_end(h, None))
    lib.mopo_ens_destroy(h)


def _dynamics(n, obs_dim=11, act_dim=3, seed=0):
    rng = np.random.RandomState(seed)
    obs = rng.standard_normal((n, obs_dim)).astype(np.float32)
    act = np.tanh(rng.standard_normal((n, act_dim))).astype(np.float32)
    A = rng.standard_normal((obs_dim + act_dim, obs_dim)) * 0.2
    x = np.concatenate([obs, act], -1)
    next_obs = (obs + np.sin(x @ A) * 0.5 + 0.02 * rng.standard_normal((n, obs_dim))).astype(np.float32)
    rew = (x[:, :1] * 0.5 + act[:, :1] + 0.02 * rng.standard_normal((n, 1))).astype(np.float32)
    return {'observations': obs, 'actions': act, 'next_observations': next_obs, 'rewards': rew}


@pytest.mark.gpu
def test_bnn_train_one_epoch_matches_oracle_loop():
    """BNN.train with the reference's random draws pinned (permutation, bootstrap idxs): one epoch of the loop equals
    the oracle's epoch -- holdout losses, elites, the snapshot taken at epoch 0 and the host copies of the weights."""
    from mopo_b200.models.constructor import construct_model, format_samples_for_training
    samples = _dynamics(700, seed=5)
    inputs, targets = format_samples_for_training(samples)
    model = construct_model(obs_dim=11, act_dim=3, hidden_dim=200, num_networks=7, num_elites=5, separate_mean_var=True, seed=11)
    rng = np.random.RandomState(9)
    perm = rng.permutation(700)
    n_hold = min(int(700 * 0.2), 1000)
    idxs = rng.randint(700 - n_hold, size=(7, 700 - n_hold))
    p = {'obs_dim': 11, 'act_dim': 3, 'rew_dim': 1, 'hidden_dim': 200, 'num_networks': 7, 'separate_mean_var': True,
         'weights': [l.weights.copy() for l in model.layers], 'biases': [l.biases.copy() for l in model.layers],
         'var_weights': model.var_layers[0].weights.copy(), 'var_biases': model.var_layers[0].biases.copy(),
         'max_logvar': model.max_logvar.copy(), 'min_logvar': model.min_logvar.copy()}
    metrics = model.train(inputs, targets, batch_size=256, max_epochs=1, holdout_ratio=0.2, permutation=perm, idxs=idxs)
    tr_in, ho_in = inputs[perm[n_hold:]], inputs[perm[:n_hold]]
    tr_tg, ho_tg = targets[perm[n_hold:]], targets[perm[:n_hold]]
    ens.fit_scaler(p, tr_in)
    np.testing.assert_allclose(model.scaler.mu, p['mu'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(model.scaler.sigma, p['sigma'], rtol=1e-5)
    v = otr.trainable(p)
    st = otr.adam_init(v)
    otr.train_batches(p, v, st, tr_in, tr_tg, idxs, 256)
    want = otr.mse(p, v, np.tile(ho_in[None], [7, 1, 1]).astype(np.float64), np.tile(ho_tg[None], [7, 1, 1]).astype(np.float64))
    assert model._grad_updates == 3
    np.testing.assert_allclose(metrics['val_loss'], np.sort(want)[:5].mean(), rtol=2e-3)
    assert model._model_inds == np.argsort(want)[:5].tolist()
    assert all(model._snapshots[i][0] == 0 for i in range(7))
    np.testing.assert_allclose([model._snapshots[i][1] for i in range(7)], want, rtol=2e-3)
    got = {'weights': [l.weights for l in model.layers], 'biases': [l.biases for l in model.layers],
           'var_weights': model.var_layers[0].weights, 'var_biases': model.var_layers[0].biases,
           'max_logvar': model.max_logvar, 'min_logvar': model.min_logvar}
    _compare_params(got, v, 1e-4)


@pytest.mark.gpu
def test_bnn_train_learns_and_feeds_the_rollout(tmp_path):
    """End to end (mopo.py:_train_model -> _rollout_model): train on synthetic dynamics until the early-stop rule fires
    or max_epochs, check the holdout loss fell, elites are the best holdout nets, the bf16 images were rebuilt from the
    trained weights and save() / load round-trips them."""
    import torch
    from mopo_b200 import _lib
    from mopo_b200.models.constructor import construct_model, format_samples_for_training
    samples = _dynamics(6000, seed=1)
    inputs, targets = format_samples_for_training(samples)
    model = construct_model(obs_dim=11, act_dim=3, hidden_dim=200, num_networks=7, num_elites=5, separate_mean_var=True, seed=3)
    with pytest.raises(AttributeError):
        model._model_inds
    before_mean, _ = model.predict(inputs[:500], factored=True)
    mse0 = ((before_mean - targets[None, :500]) ** 2).mean((1, 2))
    metrics = model.train(inputs, targets, batch_size=256, max_epochs=30, holdout_ratio=0.2, max_epochs_since_update=5)
    mean, var = model.predict(inputs[:500], factored=True)
    mse1 = ((mean - targets[None, :500]) ** 2).mean((1, 2))
    assert np.all(mse1 < 0.2 * mse0), (mse0, mse1)
    assert metrics['val_loss'] < 0.2 * float(mse0.mean())
    assert len(model._model_inds) == 5 and len(set(model._model_inds)) == 5
    # tensor-core images follow the trained weights: bf16 prediction tracks the fp32 one
    mean_bf, _ = model.predict(inputs[:500], factored=True, precision=_lib.PREC_BF16)
    assert np.abs(mean_bf - mean).max() < 0.1 and ((mean_bf - targets[None, :500]) ** 2).mean() < 1.5 * mse1.mean() + 1e-3
    # host copies == device weights: save -> load into a fresh model -> same predictions
    model.save(str(tmp_path), 0)
    import os
    os.rename(os.path.join(str(tmp_path), 'BNN_0.mat'), os.path.join(str(tmp_path), 'BNN.mat'))
    clone = construct_model(obs_dim=11, act_dim=3, hidden_dim=200, num_networks=7, num_elites=5, separate_mean_var=True,
                            load_dir=str(tmp_path))
    mean2, var2 = clone.predict(inputs[:500], factored=True)
    np.testing.assert_array_equal(mean2, mean)
    np.testing.assert_array_equal(var2, var)
    # no holdout -> no saved state -> the reference's KeyError in _set_state (bnn.py:298), RuntimeError here
    with pytest.raises(RuntimeError):
        model.train(inputs[:500], targets[:500], batch_size=256, max_epochs=1, holdout_ratio=0.0)


@pytest.mark.gpu
def test_train_graph_replay_equals_direct_launches():
    """On a non-default stream the full batches of an epoch replay one captured CUDA graph (column offset and Adam
    step count live on the device); the result is bit-identical to launching every kernel directly, over two epochs
    with a ragged last batch, and the launch count claims the same number of kernels."""
    import torch
    from mopo_b200 import _lib
    p, inputs, targets, rng = _problem(17, 6, 200, 7, True, 1000, seed=21)
    n_cols, bs = 6 * 64 + 17, 64
    idx = rng.randint(1000, size=(7, n_cols)).astype(np.int32)
    d_in, d_tg, d_idx = (torch.as_tensor(a).cuda() for a in (inputs, targets, idx))
    results, counts = [], []
    side = torch.cuda.Stream()
    for stream in (None, side):
        lib, h = _handle(p)
        _lib.check(lib.mopo_ens_train_begin(h, 1e-3, _decays(p)))
        torch.cuda.synchronize()
        c0 = lib.mopo_launch_count()
        for _ in range(2):
            _lib.check(lib.mopo_ens_train_batches(h, d_in.data_ptr(), d_tg.data_ptr(), 1000, d_idx.data_ptr(), n_cols, n_cols, bs,
                                                  None, stream.cuda_stream if stream is not None else None))
        torch.cuda.synchronize()
        counts.append(lib.mopo_launch_count() - c0)
        results.append(_get_params(lib, h, p))
        _lib.check(lib.mopo_ens_train_end(h, None))
        lib.mopo_ens_destroy(h)
    assert counts[0] == counts[1] and counts[0] % 14 == 0
    for (ac, ak), (bc, bk) in zip(otr._leaves(results[0]), otr._leaves(results[1])):
        assert np.array_equal(ac[ak], bc[bk]), ak


@pytest.mark.gpu
def test_train_model_then_rollout_stays_on_device():
    """MOPO's model phase end to end (mopo.py:528-529, 554-560): offline env pool -> _train_model (BNN.train on the pool's
    device fields) -> elites -> _rollout_model with the on-device actor on the tensor-core path into the model pool."""
    import torch
    from mopo_b200 import _lib
    import mopo_b200.static as static
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    from mopo_b200.policies import GaussianPolicy
    samples = _dynamics(5000, obs_dim=17, act_dim=6, seed=2)
    samples['terminals'] = np.zeros((5000, 1), bool)
    env_pool = SimpleReplayPool(17, 6, 5000)
    env_pool.add_samples(samples)
    model = construct_model(obs_dim=17, act_dim=6, hidden_dim=200, num_networks=7, num_elites=5, separate_mean_var=True,
                            precision=_lib.PREC_BF16, seed=5)
    env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=7)
    from oracle import policy as opol
    pol = GaussianPolicy(17, 6, device=0, precision=_lib.PREC_BF16, seed=9)
    pol.set_params(*opol.init_params(17, 6, seed=4))
    ro = ModelRollout(env, env_pool, rollout_batch_size=2000, rollout_schedule=(0, 1, 3, 3), epoch_length=1000,
                      model_train_freq=1000, model_retain_epochs=1, policy=pol, seed=3)
    metrics = ro._train_model(batch_size=256, max_epochs=8, holdout_ratio=0.2)
    assert np.isfinite(metrics['val_loss']) and len(model._model_inds) == 5
    ro._set_rollout_length()
    ro._reallocate_model_pool()
    stats = ro._rollout_model(2000, quiet=True)
    torch.cuda.synchronize()
    assert stats['mean_rollout_length'] == 3.0 and ro._model_pool.size == 6000
    f = {k: v.cpu().numpy() for k, v in ro._model_pool.fields.items()}
    assert np.isfinite(f['next_observations']).all() and np.isfinite(f['rewards']).all()
    # the trained model is a model of these dynamics: one-step predictions from pool states track the generator
    x = np.concatenate([samples['observations'][:300], samples['actions'][:300]], -1)
    mean, _ = model.predict(x, factored=False, precision=_lib.PREC_FP32)
    want = np.concatenate([samples['rewards'][:300], samples['next_observations'][:300] - samples['observations'][:300]], -1)
    base = (want ** 2).mean()
    assert ((mean - want) ** 2).mean() < 0.5 * base


@pytest.mark.gpu
@pytest.mark.parametrize('cfg', [dict(num_networks=7, num_elites=5, separate_mean_var=False),
                                 dict(num_networks=1, num_elites=1, separate_mean_var=True)])
def test_bnn_train_other_layouts(cfg, tmp_path):
    """BNN.train with the combined [mean | logvar] head (bnn.py:183-189) and with a single-net 'ensemble': the loop runs,
    the holdout loss falls, and the trained host copies round-trip through save / load."""
    import os
    from mopo_b200.models.constructor import construct_model, format_samples_for_training
    samples = _dynamics(3000, seed=4)
    inputs, targets = format_samples_for_training(samples)
    model = construct_model(obs_dim=11, act_dim=3, hidden_dim=200, seed=2, **cfg)
    m0, _ = model.predict(inputs[:300], factored=True)
    mse0 = ((m0 - targets[None, :300]) ** 2).mean()
    metrics = model.train(inputs, targets, batch_size=256, max_epochs=15, holdout_ratio=0.2)
    m1, v1 = model.predict(inputs[:300], factored=True)
    assert ((m1 - targets[None, :300]) ** 2).mean() < 0.3 * mse0 and np.isfinite(metrics['val_loss'])
    assert len(model._model_inds) == cfg['num_elites']
    model.save(str(tmp_path), 0)
    os.rename(os.path.join(str(tmp_path), 'BNN_0.mat'), os.path.join(str(tmp_path), 'BNN.mat'))
    clone = construct_model(obs_dim=11, act_dim=3, hidden_dim=200, load_dir=str(tmp_path), **cfg)
    m2, v2 = clone.predict(inputs[:300], factored=True)
    np.testing.assert_array_equal(m2, m1)
    np.testing.assert_array_equal(v2, v1)
