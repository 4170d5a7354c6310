"""Generate tests/golden/*.npz from the REFERENCE's own code (run in the build container only).

    python tests/golden/make_golden.py

Imports /root/reference/mopo/models/fake_env.py (FakeEnv), mopo/static/*.py, the reference's
FlexibleReplayPool AND its model code -- mopo/models/constructor.py, bnn.py, fc.py, utils.py --
all unmodified.  TensorFlow 1.14 is absent from this image, so ``tensorflow`` is
tests/golden/tf_numpy_stub.py: a graph-mode stand-in that evaluates the dozen ops those files call
in float32 NumPy.  FakeEnv.step is therefore driven by the reference's own BNN (construct_model ->
finalize -> predict via sess.run), loaded with the seeded synthetic parameters, with
np.random.normal / model.random_inds patched so the injected noise and elite indices are recorded.
fwd_*.npz hold BNN.predict itself (2-D factored / aggregated, 3-D per-net inputs, both head modes).
What stays third-party is the arithmetic inside TensorFlow's kernels.  The committed fixtures travel
to the GPU box; /root/reference does not.
"""
import os, sys, types, importlib.util
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference'
sys.path.insert(0, ROOT)
from oracle import ensemble as ens  # noqa: E402


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    sys.path.insert(0, HERE)
    import tf_numpy_stub
    tf_numpy_stub.install()
    sys.path.insert(0, REF)
    fe = _load('ref_fake_env', os.path.join(REF, 'mopo/models/fake_env.py'))
    static = {}
    for f in sorted(os.listdir(os.path.join(REF, 'mopo/static'))):
        if f.startswith('__') or not f.endswith('.py'):
            continue
        static[f[:-3]] = _load('ref_static_' + f[:-3], os.path.join(REF, 'mopo/static', f)).StaticFns
    # flexible_replay_pool.py does `from .replay_pool import ReplayPool`: mount the directory as a bare
    # package (without running its __init__, which imports gym via simple_replay_pool.py:4).
    pkg = types.ModuleType('ref_replay_pools')
    pkg.__path__ = [os.path.join(REF, 'softlearning/replay_pools')]
    sys.modules['ref_replay_pools'] = pkg
    import importlib
    pool = importlib.import_module('ref_replay_pools.flexible_replay_pool')
    return fe.FakeEnv, static, pool.FlexibleReplayPool


def reference_model(p, name):
    """The reference's own BNN (constructor.py:7-43 -> bnn.py finalize), loaded with the parameter dict ``p``."""
    from mopo.models.constructor import construct_model       # /root/reference, through the TF stand-in
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        m = construct_model(obs_dim=p['obs_dim'], act_dim=p['act_dim'], rew_dim=p['rew_dim'], hidden_dim=p['hidden_dim'],
                            num_networks=p['num_networks'], num_elites=5, separate_mean_var=p['separate_mean_var'], name=name)
    for layer, w, b in zip(m.layers, p['weights'], p['biases']):
        layer.weights.load(w); layer.biases.load(b)
    if p['separate_mean_var']:
        m.var_layers[0].weights.load(p['var_weights']); m.var_layers[0].biases.load(p['var_biases'])
    m.scaler.mu.load(p['mu']); m.scaler.sigma.load(p['sigma']); m.scaler.cache()
    m.max_logvar.load(p['max_logvar']); m.min_logvar.load(p['min_logvar'])
    return m


def forward_case(domain, B, seed, separate):
    """BNN.predict of the reference (bnn.py:508-546 over _compile_outputs, bnn.py:631-675): every input mode."""
    p, obs, act = ens.synthetic_case(domain, B, seed=seed, separate_mean_var=separate)
    m = reference_model(p, 'fwd_%s_%d' % (domain, seed))
    x = np.concatenate([obs, act], -1)
    rng = np.random.RandomState(seed + 3)
    x3 = (x[None] + 0.1 * rng.standard_normal((p['num_networks'],) + x.shape)).astype(np.float32)
    mean_fac, var_fac = m.predict(x, factored=True)
    mean, var = m.predict(x, factored=False)
    mean3, var3 = m.predict(x3, factored=True)
    assert mean_fac.dtype == np.float32 and mean_fac.shape == (p['num_networks'], B, p['obs_dim'] + 1)
    return dict(domain=domain, B=B, seed=seed, separate=separate, x3=x3, mean_fac=mean_fac, var_fac=var_fac,
                mean=mean, var=var, mean3=mean3, var3=var3)


def step_case(FakeEnv, static, domain, B, seed, learned_var, deterministic, coeff=1.0, separate=True, model_inds_set=None):
    p, obs, act = ens.synthetic_case(domain, B, seed=seed, separate_mean_var=separate)
    model = reference_model(p, 'step_%s_%d' % (domain, seed))
    model._model_inds = list(range(5)) if model_inds_set is None else list(model_inds_set)
    rng = np.random.RandomState(seed + 7)
    D1 = obs.shape[1] + 1
    noise = rng.standard_normal((model.num_nets, B, D1))
    inds = rng.choice(model._model_inds, size=B)
    model.random_inds = lambda batch_size: inds           # bnn.py:342-344 draws from global np.random: inject instead
    env = FakeEnv(model, static[domain], penalty_coeff=coeff, penalty_learned_var=learned_var)
    orig = np.random.normal
    np.random.normal = lambda size=None: noise.reshape(size)
    try:
        next_obs, rew, term, info = env.step(obs, act, deterministic=deterministic)
    finally:
        np.random.normal = orig
    out = dict(domain=domain, B=B, seed=seed, learned_var=learned_var, deterministic=deterministic,
               coeff=coeff, separate=separate, elites=np.asarray(model._model_inds),
               noise=noise, model_inds=inds, next_obs=next_obs, rew=rew, term=term,
               info_mean=info['mean'], info_std=info['std'], info_log_prob=info['log_prob'], info_dev=info['dev'],
               unpenalized=info['unpenalized_rewards'],
               penalty=np.zeros((0,)) if info['penalty'] is None else info['penalty'])
    return out


def termination_case(static, domain, obs_dim, seed):
    """Adversarial next_obs rows: thresholds +- ulp, NaN, +-Inf, large magnitudes."""
    rng = np.random.RandomState(seed)
    B = 256
    nobs = rng.standard_normal((B, obs_dim)).astype(np.float32)
    nobs[:, 0] = rng.uniform(0.0, 2.5, B)
    nobs[:, 1] = rng.uniform(-1.3, 1.3, B) * (0.25 if domain == 'hopper' else 1.0)
    thr = np.float32([0.7, 0.8, 2.0, 0.2, 1.0, -1.0, -0.2, 100.0, 0.0])
    r = 0
    for t in thr:
        for v in (t, np.nextafter(t, np.float32(np.inf)), np.nextafter(t, np.float32(-np.inf))):
            for col in (0, 1, 2):
                if col < obs_dim:
                    nobs[r % B, col] = v; r += 1
    for v in (np.nan, np.inf, -np.inf, 500.0, -500.0, 99.99999, 100.00001):
        for col in (0, 1, 2, obs_dim - 1):
            nobs[r % B, col] = v; r += 1
    # healthy rows so both outcomes occur
    nobs[-32:, 0] = 1.25; nobs[-32:, 1] = 0.01
    if domain in ('ant', 'antangle'):
        nobs[-32:, 0] = 0.6
    obs = np.zeros_like(nobs); act = np.zeros((B, 1), np.float32)
    # In the reference pipeline next_obs is float64 (samples = f32 mean + f64 noise * f32 std, fake_env.py:72), so the
    # thresholds are compared in double: feed the f32-representable values as float64.
    term = static[domain].termination_fn(obs, act, nobs.astype(np.float64))
    return dict(domain=domain, next_obs=nobs, term=np.asarray(term))


def pool_case(FlexibleReplayPool, seed):
    rng = np.random.RandomState(seed)
    obs_dim, act_dim, max_size = 5, 2, 37
    fields = {'actions': {'shape': (act_dim,), 'dtype': 'float32'}, 'rewards': {'shape': (1,), 'dtype': 'float32'},
              'terminals': {'shape': (1,), 'dtype': 'bool'}, 'observations': {'shape': (obs_dim,), 'dtype': 'float32'},
              'next_observations': {'shape': (obs_dim,), 'dtype': 'float32'}}
    pool = FlexibleReplayPool(max_size, fields)
    chunks, ptrs, sizes = [], [], []
    for n in (10, 20, 15, 1, 50, 37, 3):     # wraps, n > max_size (last-write-wins), n == max_size
        s = {'observations': rng.standard_normal((n, obs_dim)), 'actions': rng.standard_normal((n, act_dim)),
             'next_observations': rng.standard_normal((n, obs_dim)), 'rewards': rng.standard_normal((n, 1)),
             'terminals': rng.rand(n, 1) < 0.3}
        pool.add_samples(s)
        chunks.append(s); ptrs.append(pool._pointer); sizes.append(pool.size)
    out = dict(obs_dim=obs_dim, act_dim=act_dim, max_size=max_size, n_chunks=len(chunks), ptrs=np.asarray(ptrs), sizes=np.asarray(sizes))
    for i, s in enumerate(chunks):
        for k, v in s.items():
            out['chunk%d_%s' % (i, k)] = v
    for k, v in pool.fields.items():
        out['final_' + k] = v
    return out


def main():
    FakeEnv, static, FlexibleReplayPool = load_reference()
    cases = []
    B = 96
    for domain in ('halfcheetah', 'hopper', 'walker2d'):
        cases.append(step_case(FakeEnv, static, domain, B, seed=11, learned_var=True, deterministic=False))
    cases.append(step_case(FakeEnv, static, 'hopper', B, seed=12, learned_var=False, deterministic=False))
    cases.append(step_case(FakeEnv, static, 'walker2d', B, seed=13, learned_var=True, deterministic=True))
    cases.append(step_case(FakeEnv, static, 'halfcheetah', B, seed=14, learned_var=True, deterministic=False, coeff=0.0))
    cases.append(step_case(FakeEnv, static, 'halfcheetah', B, seed=15, learned_var=True, deterministic=False, separate=False,
                           model_inds_set=[6, 2, 3, 0, 5]))
    cases.append(step_case(FakeEnv, static, 'ant', B, seed=16, learned_var=True, deterministic=False))
    cases.append(step_case(FakeEnv, static, 'humanoid', B, seed=17, learned_var=False, deterministic=False, coeff=2.5))
    for i, c in enumerate(cases):
        np.savez_compressed(os.path.join(HERE, 'step_%02d_%s.npz' % (i, c['domain'])), **c)
    fwd = [forward_case('halfcheetah', 64, 31, True), forward_case('hopper', 50, 32, False),
           forward_case('humanoid', 40, 33, True), forward_case('walker2d', 33, 34, True)]
    for i, c in enumerate(fwd):
        np.savez_compressed(os.path.join(HERE, 'fwd_%02d_%s.npz' % (i, c['domain'])), **c)
    dims = {'halfcheetah': 17, 'hopper': 11, 'walker2d': 17, 'ant': 27, 'antangle': 27, 'humanoid': 45,
            'pendulum': 3, 'point2denv': 4}
    for i, (d, od) in enumerate(dims.items()):
        if d == 'pendulum':
            continue    # returns float zeros (pendulum.py:9); not a bool mask
        np.savez_compressed(os.path.join(HERE, 'term_%s.npz' % d), **termination_case(static, d, od, 100 + i))
    np.savez_compressed(os.path.join(HERE, 'pool_ring.npz'), **pool_case(FlexibleReplayPool, 5))
    print('wrote', sorted(f for f in os.listdir(HERE) if f.endswith('.npz')))


if __name__ == '__main__':
    main()
