"""The SAC update of MOPO on the device (reference mopo/algorithms/mopo.py): the actor / twin-critic / alpha graph of
:271-457 and the training step `_do_training` (:829-843) with `_training_batch`'s env / model pool mix (:801-821).

`SACUpdate` owns the networks as a C handle (mopo_sac_*): every gradient, the four Adam optimisers and the Polyak target
update run in libmopo_b200.so; batches are drawn from the device replay pools without touching the host.  Variables use
tf.layers.dense's layout (kernel [in, out], bias [out]) and the reference's scopes: main/pi (hidden0, hidden1, mu,
log_std), main/q1, main/q2 (hidden0, hidden1, out), target/q1, target/q2, log_alpha.
"""
import ctypes as C
import numpy as np
import torch

from .. import _lib

LOSS_NAMES = ('Q/q1_loss', 'sac_Q/q2_loss', 'sac_pi/policy_loss', 'sac_pi/alpha_loss')


def _glorot(rng, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))                       # tf.layers.dense default kernel_initializer
    return rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32)


class SACUpdate:
    def __init__(self, obs_dim, act_dim, hidden_sizes=(256, 256), device=0, lr=3e-4, Q_lr=None, reward_scale=1.0, discount=0.99,
                 tau=5e-3, target_entropy='auto', target_update_interval=1, real_ratio=0.05, batch_size=256, seed=0):
        if len(hidden_sizes) != 2 or hidden_sizes[0] != hidden_sizes[1]:
            raise ValueError('SACUpdate: two equal hidden layers (network_kwargs hidden_sizes=(256, 256)), got %r' % (hidden_sizes,))
        self.obs_dim, self.act_dim, self.hidden = int(obs_dim), int(act_dim), int(hidden_sizes[0])
        self.device = torch.device('cuda', device)
        if target_entropy == 'auto':                               # mopo.py:165-168
            target_entropy = -float(act_dim)
        self.hp = _lib.SacHParams(reward_scale, discount, tau, lr, lr if Q_lr is None else Q_lr, float(target_entropy),
                                  int(target_update_interval), 0)
        self._real_ratio, self._batch_size = float(real_ratio), int(batch_size)
        self.seed = int(seed)
        self._num_train_steps = 0
        h = C.c_void_p()
        _lib.check(_lib.load().mopo_sac_create(self.obs_dim, self.act_dim, self.hidden, device, C.byref(self.hp), C.byref(h)))
        self._handle = h
        rng = np.random.RandomState(self.seed)
        od, ad, H = self.obs_dim, self.act_dim, self.hidden
        pi_W = [_glorot(rng, od, H), _glorot(rng, H, H), _glorot(rng, H, ad), _glorot(rng, H, ad)]
        q = lambda: [_glorot(rng, od + ad, H), _glorot(rng, H, H), _glorot(rng, H, 1)]
        zeros = lambda Ws: [np.zeros(w.shape[1], np.float32) for w in Ws]
        q1_W, q2_W = q(), q()
        self.set_params(pi_W, zeros(pi_W), q1_W, zeros(q1_W), q2_W, zeros(q2_W), 0.0)

    def __del__(self):
        try:
            if getattr(self, '_handle', None):
                _lib.load().mopo_sac_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def _shapes(self):
        od, ad, H = self.obs_dim, self.act_dim, self.hidden
        return [(od, H), (H, H), (H, ad), (H, ad)], [(od + ad, H), (H, H), (H, 1)]

    # ---- variables ----------------------------------------------------------------------------------
    def set_params(self, pi_W, pi_b, q1_W, q1_b, q2_W, q2_b, log_alpha=0.0):
        """Install the main networks; copies them into the targets (target_init, mopo.py:451-452) and resets Adam."""
        ps, qs = self._shapes()
        keep = []

        def pack(Ws, bs, shapes, who):
            if len(Ws) != len(shapes) or len(bs) != len(shapes):
                raise ValueError('SACUpdate.set_params: %s needs %d dense layers' % (who, len(shapes)))
            wa, ba = [], []
            for i, (w, b) in enumerate(zip(Ws, bs)):
                w = np.ascontiguousarray(np.asarray(w, np.float32)); b = np.ascontiguousarray(np.asarray(b, np.float32).reshape(-1))
                if w.shape != shapes[i] or b.shape != (shapes[i][1],):
                    raise ValueError('SACUpdate.set_params: %s layer %d has shape %r / %r, expected %r' % (who, i, w.shape, b.shape, shapes[i]))
                wa.append(w); ba.append(b)
            keep.extend(wa + ba)
            return (C.c_void_p * len(wa))(*[a.ctypes.data for a in wa]), (C.c_void_p * len(ba))(*[a.ctypes.data for a in ba])

        pw, pb = pack(pi_W, pi_b, ps, 'pi')
        w1, b1 = pack(q1_W, q1_b, qs, 'q1')
        w2, b2 = pack(q2_W, q2_b, qs, 'q2')
        _lib.check(_lib.load().mopo_sac_set_params(self._handle, pw, pb, w1, b1, w2, b2, float(log_alpha), _lib.stream_ptr(self.device)))

    def get_params(self):
        """dict of host arrays: pi / q1 / q2 / q1_targ / q2_targ -> {'W': [...], 'b': [...]}, log_alpha."""
        ps, qs = self._shapes()
        out, ptrs = {}, []
        for name, shapes in (('pi', ps), ('q1', qs), ('q2', qs), ('q1_targ', qs), ('q2_targ', qs)):
            W = [np.empty(s, np.float32) for s in shapes]
            b = [np.empty(s[1], np.float32) for s in shapes]
            out[name] = {'W': W, 'b': b}
            ptrs += [(C.c_void_p * len(W))(*[a.ctypes.data for a in W]), (C.c_void_p * len(b))(*[a.ctypes.data for a in b])]
        la = C.c_float()
        _lib.check(_lib.load().mopo_sac_get_params(self._handle, *ptrs, C.byref(la), _lib.stream_ptr(self.device)))
        out['log_alpha'] = float(la.value)
        return out

    def _opt_sizes(self):
        sizes = (C.c_int64 * 13)()
        _lib.check(_lib.load().mopo_sac_opt_sizes(self._handle, sizes))
        return [int(x) for x in sizes]

    def get_opt_state(self):
        """Adam slots of the 13 variables (C-ABI order), the Adam step and the update index that keys the random draws."""
        sizes = self._opt_sizes()
        m = [np.empty(n, np.float32) for n in sizes]
        v = [np.empty(n, np.float32) for n in sizes]
        t, draw = C.c_int64(), C.c_uint64()
        _lib.check(_lib.load().mopo_sac_get_opt_state(self._handle, (C.c_void_p * 13)(*[a.ctypes.data for a in m]),
                                                      (C.c_void_p * 13)(*[a.ctypes.data for a in v]), C.byref(t), C.byref(draw),
                                                      _lib.stream_ptr(self.device)))
        return {'m': m, 'v': v, 't': int(t.value), 'draw': int(draw.value)}

    def set_opt_state(self, state):
        sizes = self._opt_sizes()
        m = [np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1)) for a in state['m']]
        v = [np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1)) for a in state['v']]
        if [a.size for a in m] != sizes or [a.size for a in v] != sizes:
            raise ValueError('SACUpdate.set_opt_state: slot sizes %r, expected %r' % ([a.size for a in m], sizes))
        _lib.check(_lib.load().mopo_sac_set_opt_state(self._handle, (C.c_void_p * 13)(*[a.ctypes.data for a in m]),
                                                      (C.c_void_p * 13)(*[a.ctypes.data for a in v]), int(state['t']), int(state['draw']),
                                                      _lib.stream_ptr(self.device)))
        self._num_train_steps = int(state['t'])

    def save(self, path):
        """Checkpoint as one .npz: every variable (main + target networks, log_alpha) and the optimiser state (Adam slots, Adam
        step, the update index behind the minibatch draws and the policy noise) -- what the reference's TF-session checkpoint
        holds, so that a resumed run continues the same sequence instead of replaying it from update 0."""
        p = self.get_params()
        flat = {'log_alpha': np.float32(p['log_alpha'])}
        for name in ('pi', 'q1', 'q2', 'q1_targ', 'q2_targ'):
            for k in ('W', 'b'):
                for i, a in enumerate(p[name][k]):
                    flat['%s/%s%d' % (name, k, i)] = a
        o = self.get_opt_state()
        for i in range(13):
            flat['opt/m%d' % i] = o['m'][i]
            flat['opt/v%d' % i] = o['v'][i]
        flat['opt/t'] = np.int64(o['t'])
        flat['opt/draw'] = np.uint64(o['draw'])
        np.savez(path, **flat)

    def load(self, path):
        d = np.load(path)
        get = lambda name, k, n: [d['%s/%s%d' % (name, k, i)] for i in range(n)]
        self.set_params(get('pi', 'W', 4), get('pi', 'b', 4), get('q1', 'W', 3), get('q1', 'b', 3), get('q2', 'W', 3), get('q2', 'b', 3),
                        float(d['log_alpha']))
        # set_params runs target_init and resets the optimisers; a checkpoint taken mid-training carries its own targets and slots
        self._set_targets(get('q1_targ', 'W', 3), get('q1_targ', 'b', 3), get('q2_targ', 'W', 3), get('q2_targ', 'b', 3))
        if 'opt/t' in d.files:                      # (checkpoints written before the optimiser state was saved: variables only)
            self.set_opt_state({'m': [d['opt/m%d' % i] for i in range(13)], 'v': [d['opt/v%d' % i] for i in range(13)],
                                't': int(d['opt/t']), 'draw': int(d['opt/draw'])})

    def _set_targets(self, q1t_W, q1t_b, q2t_W, q2t_b):
        keep = []

        def pack(arrs):
            a = [np.ascontiguousarray(np.asarray(x, np.float32)) for x in arrs]
            keep.extend(a)
            return (C.c_void_p * len(a))(*[x.ctypes.data for x in a])

        _lib.check(_lib.load().mopo_sac_set_targets(self._handle, pack(q1t_W), pack(q1t_b), pack(q2t_W), pack(q2t_b),
                                                    _lib.stream_ptr(self.device)))

    def export_policy(self, policy):
        """Push the actor (main/pi) into the rollout's GaussianPolicy handle, device to device."""
        _lib.check(_lib.load().mopo_sac_export_policy(self._handle, policy._handle, _lib.stream_ptr(self.device)))

    # ---- MOPO._do_training (mopo.py:829-843) ----------------------------------------------------------
    def _do_training(self, iteration, batch, eps=None, return_losses=True):
        """One update on `batch` (dict with the reference's field names; NumPy or torch).  `eps` [2n, act] pins the
        graph's two tf.random_normal draws (pi(s) rows first, then pi(s')) for parity tests."""
        dev = self.device
        f = lambda k, dt: torch.as_tensor(batch[k]).to(device=dev, dtype=dt).contiguous()
        obs, act, nobs = f('observations', torch.float32), f('actions', torch.float32), f('next_observations', torch.float32)
        rew = f('rewards', torch.float32).reshape(-1)
        term = f('terminals', torch.uint8).reshape(-1)
        n = obs.shape[0]
        if obs.shape != (n, self.obs_dim) or act.shape != (n, self.act_dim) or nobs.shape != obs.shape or rew.shape != (n,) or term.shape != (n,):
            raise ValueError('SACUpdate._do_training: batch shapes do not match obs %d / act %d' % (self.obs_dim, self.act_dim))
        e = None if eps is None else torch.as_tensor(eps).to(device=dev, dtype=torch.float32).contiguous()
        if e is not None and e.shape != (2 * n, self.act_dim):
            raise ValueError('eps must be [2n, act_dim]')
        losses = torch.empty(4, dtype=torch.float32, device=dev) if return_losses else None
        b = _lib.SacBatch(obs.data_ptr(), act.data_ptr(), rew.data_ptr(), term.data_ptr(), nobs.data_ptr(), n, _lib.ptr(e), self.seed,
                          _lib.ptr(losses))
        _lib.check(_lib.load().mopo_sac_update(self._handle, C.byref(b), _lib.stream_ptr(dev)))
        self._num_train_steps += 1
        if not return_losses:
            return {}
        return dict(zip(LOSS_NAMES, losses.cpu().numpy().tolist()))

    # ---- _training_batch + _do_training_repeats (mopo.py:780-821) on the device pools -------------------
    def batch_sizes(self, batch_size=None):
        batch_size = batch_size or self._batch_size
        env_batch_size = int(batch_size * self._real_ratio)                # mopo.py:803
        return env_batch_size, batch_size - env_batch_size

    def train_from_pools(self, env_pool, model_pool, n_updates=1, batch_size=None, return_losses=False):
        """n_updates x (_training_batch + _do_training) in ONE C call: rows are drawn from the device pools (env rows first,
        then model rows, mopo.py:808-815) and never leave HBM."""
        n_env, n_model = self.batch_sizes(batch_size)
        if model_pool is None:                                             # real_ratio == 1.0 (mopo.py:816-819)
            n_env, n_model = n_env + n_model, 0
        losses = torch.empty((n_updates, 4), dtype=torch.float32, device=self.device) if return_losses else None
        # on a side stream, so that the n_updates one-kernel updates queue behind each other beside the caller's stream
        cur = torch.cuda.current_stream(self.device)
        if getattr(self, '_stream', None) is None:
            self._stream = torch.cuda.Stream(self.device)
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            _lib.check(_lib.load().mopo_sac_update_from_pools(self._handle, env_pool._handle, model_pool._handle if model_pool is not None else None,
                                                              n_env, n_model, self.seed, int(n_updates), _lib.ptr(losses),
                                                              _lib.stream_ptr(self.device)))
        cur.wait_stream(self._stream)
        self._num_train_steps += int(n_updates)
        if return_losses:
            m = losses.mean(0).cpu().numpy().tolist()
            return dict(zip(LOSS_NAMES, m))
        return {}
