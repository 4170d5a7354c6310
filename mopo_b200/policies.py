"""Device mirror of the SAC actor the rollout queries (reference mopo/algorithms/mopo.py): the graph of
mlp_gaussian_policy + apply_squashing_func (:281-312, hidden_sizes (256, 256), ReLU) behind `get_action_meta`
(:468-485).  Held as a C handle (mopo_policy_*) so that `MOPO._rollout_model`'s `act = policy(obs)` (:734) runs on
the device between two model steps: pass the object as `ModelRollout(..., policy=...)`.

Parameters use tf.layers.dense's layout: kernel [in, out], bias [out]; names follow the reference's variable scopes
(`pi/dense`, `pi/dense_1`, `pi/dense_2` = mu, `pi/dense_3` = log_std).
"""
import ctypes as C
import numpy as np
import torch

from . import _lib

LOG_STD_MAX, LOG_STD_MIN = 2, -20          # mopo.py:271-272


class GaussianPolicy:
    def __init__(self, obs_dim, act_dim, hidden_sizes=(256, 256), device=0, precision=None, seed=0):
        if len(hidden_sizes) != 2 or hidden_sizes[0] != hidden_sizes[1]:
            raise ValueError('GaussianPolicy: two equal hidden layers (network_kwargs hidden_sizes=(256, 256)), got %r'
                             % (hidden_sizes,))
        self.obs_dim, self.act_dim, self.hidden = int(obs_dim), int(act_dim), int(hidden_sizes[0])
        self.device = torch.device('cuda', device)
        self.precision = _lib.PREC_FP32 if precision is None else precision
        self.seed = int(seed)
        self._calls = 0
        h = C.c_void_p()
        _lib.check(_lib.load().mopo_policy_create(self.obs_dim, self.act_dim, self.hidden, device, C.byref(h)))
        self._handle = h

    def __del__(self):
        try:
            if getattr(self, '_handle', None):
                _lib.load().mopo_policy_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ---- parameters ---------------------------------------------------------------------------------
    def set_params(self, weights, biases):
        """weights = [W0 [obs,H], W1 [H,H], Wmu [H,act], Wls [H,act]], biases likewise (NumPy or torch, any device)."""
        shapes = [(self.obs_dim, self.hidden), (self.hidden, self.hidden), (self.hidden, self.act_dim), (self.hidden, self.act_dim)]
        ws, bs = [], []
        for i, (w, b) in enumerate(zip(weights, biases)):
            w = np.ascontiguousarray(torch.as_tensor(w).detach().cpu().numpy(), dtype=np.float32)
            b = np.ascontiguousarray(torch.as_tensor(b).detach().cpu().numpy(), dtype=np.float32).reshape(-1)
            if w.shape != shapes[i] or b.shape != (shapes[i][1],):
                raise ValueError('GaussianPolicy.set_params: layer %d has shape %r / %r, expected %r' % (i, w.shape, b.shape, shapes[i]))
            ws.append(w); bs.append(b)
        if len(ws) != 4:
            raise ValueError('GaussianPolicy.set_params: 4 dense layers expected')
        wp = (C.c_void_p * 4)(*[w.ctypes.data for w in ws])
        bp = (C.c_void_p * 4)(*[b.ctypes.data for b in bs])
        _lib.check(_lib.load().mopo_policy_set_params(self._handle, wp, bp, 0, _lib.stream_ptr(self.device)))
        self._params = (ws, bs)

    # ---- MOPO.get_action_meta (mopo.py:468-485) ---------------------------------------------------------
    def get_action_meta(self, state, hidden=None, deterministic=False, noise=None, precision=None, return_stats=False,
                        step_id=None):
        np_in = not torch.is_tensor(state)
        obs = torch.as_tensor(state).to(device=self.device, dtype=torch.float32)
        single = obs.dim() == 1
        if single:
            obs = obs[None]
        obs = obs.contiguous()
        B = obs.shape[0]
        act = torch.empty(B, self.act_dim, device=self.device)
        mu = torch.empty_like(act) if return_stats else None
        std = torch.empty_like(act) if return_stats else None
        nz = None if noise is None else torch.as_tensor(noise).to(device=self.device, dtype=torch.float32).contiguous()
        a = _lib.PolicyArgs(obs=obs.data_ptr(), B=B, deterministic=int(bool(deterministic)), random_uniform=0,
                            precision=self.precision if precision is None else precision, noise=_lib.ptr(nz),
                            seed=self.seed, step_id=self._calls if step_id is None else int(step_id), origin=None, origin_base=0, actions=act.data_ptr(),
                            mu_out=_lib.ptr(mu), std_out=_lib.ptr(std))
        self._calls += 1
        _lib.check(_lib.load().mopo_policy_act(self._handle, C.byref(a), _lib.stream_ptr(self.device)))
        out = act[0] if single else act
        if np_in:
            out = out.cpu().numpy()
        if return_stats:
            return out, hidden, (mu.cpu().numpy(), std.cpu().numpy())
        return out, hidden

    def actions_np(self, obs):
        return self.get_action_meta(np.asarray(obs))[0]
