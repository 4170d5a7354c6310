"""FakeEnv -- same constructor and step() contract as the reference (mopo/models/fake_env.py:5-131).

step() runs ONE fused device pass (mopo_env_step: scaler, ensemble forward, mean += obs, std,
Gaussian draw, elite gather, penalty, termination, diagnostics); NumPy in -> NumPy out,
torch CUDA in -> torch CUDA out.  Randomness: by default Philox4x32-10 on the device keyed by
(seed, step counter, row); `noise=` / `model_inds=` inject the reference's draws for parity runs
(noise may be the full [E,B,D1] array drawn at fake_env.py:72 or the already gathered [B,D1]).
"""
import ctypes as C
import numpy as np
import torch

from .. import _lib


class FakeEnv:
    def __init__(self, model, config, penalty_coeff=0., penalty_learned_var=False,
                 penalty_learned_var_random=False, seed=0):
        self.model = model
        self.config = config
        self.penalty_coeff = penalty_coeff
        self.penalty_learned_var = penalty_learned_var
        self.penalty_learned_var_random = penalty_learned_var_random
        if not hasattr(config, 'term_kind'):
            raise TypeError("FakeEnv needs device static fns: pass mopo_b200.static[domain] "
                            "(the termination predicate is compiled into the step kernel)")
        self.seed = int(seed)
        self._step_counter = 0

    @property
    def penalty_kind(self):
        if self.penalty_coeff == 0:
            return _lib.PEN_NONE                                          # fake_env.py:116-119
        return _lib.PEN_LEARNED_VAR if self.penalty_learned_var else _lib.PEN_DISAGREEMENT

    def step(self, obs, act, deterministic=False, noise=None, model_inds=None, precision=None, diagnostics=True):
        assert len(obs.shape) == len(act.shape)                           # fake_env.py:38
        is_np = not torch.is_tensor(obs)
        dev = self.model.device
        obs_t = torch.as_tensor(obs).to(device=dev, dtype=torch.float32)
        act_t = torch.as_tensor(act).to(device=dev, dtype=torch.float32)
        if obs_t.dim() == 1:                                              # fake_env.py:39-44
            obs_t, act_t, return_single = obs_t[None], act_t[None], True
        else:
            return_single = False
        obs_t, act_t = obs_t.contiguous(), act_t.contiguous()
        B, obs_dim = obs_t.shape
        D1 = self.model.out_dim
        if obs_t.shape[1] + act_t.shape[1] != self.model.in_dim:
            raise ValueError("Invalid input dimension.")
        f32 = dict(dtype=torch.float32, device=dev)
        next_obs = torch.empty((B, obs_dim), **f32)
        rew = torch.empty((B, 1), **f32)
        term = torch.empty((B, 1), dtype=torch.uint8, device=dev)
        pen = torch.empty((B, 1), **f32)
        unpen = torch.empty((B, 1), **f32)
        mm = torch.empty((B, D1), **f32)
        ms = torch.empty((B, D1), **f32)
        lp = torch.empty((B,), **f32) if diagnostics else None
        dv = torch.empty((B,), **f32) if diagnostics else None
        inds_out = torch.empty((B,), dtype=torch.int32, device=dev)
        noise_t = inds_t = None
        if model_inds is not None:
            inds_t = torch.as_tensor(model_inds).to(device=dev, dtype=torch.int32).contiguous()
        if noise is not None:
            noise_t = torch.as_tensor(noise).to(device=dev, dtype=torch.float32)
            if noise_t.dim() == 3:                                        # full [E,B,D1] draw: gather the chosen net
                if inds_t is None:
                    raise ValueError("full [E,B,D1] noise needs model_inds")
                noise_t = noise_t[inds_t.long(), torch.arange(B, device=dev)]
            noise_t = noise_t.contiguous()
        a = _lib.StepArgs(obs=obs_t.data_ptr(), act=act_t.data_ptr(), B=B,
                          term_kind=self.config.term_kind, penalty_kind=self.penalty_kind,
                          penalty_coeff=float(self.penalty_coeff), deterministic=int(bool(deterministic)),
                          precision=self.model.precision if precision is None else precision,
                          noise=_lib.ptr(noise_t), model_inds=_lib.ptr(inds_t), seed=self.seed,
                          step_id=self._step_counter, origin=None,
                          next_obs=next_obs.data_ptr(), rewards=rew.data_ptr(), terminals=term.data_ptr(),
                          penalty=pen.data_ptr(), unpenalized=unpen.data_ptr(), model_means=mm.data_ptr(),
                          model_stds=ms.data_ptr(), log_prob=_lib.ptr(lp), dev=_lib.ptr(dv),
                          model_inds_out=inds_out.data_ptr())
        self._step_counter += 1
        self.model._ensure_handle()
        _lib.check(_lib.load().mopo_env_step(self.model._handle, C.byref(a), _lib.stream_ptr(dev)))
        terminals = term.bool()
        # fake_env.py:94-95: [rew, terminal, obs...] layout for plots
        return_means = torch.cat((mm[:, :1], terminals.to(torch.float32), mm[:, 1:]), dim=-1)
        return_stds = torch.cat((ms[:, :1], torch.zeros((B, 1), **f32), ms[:, 1:]), dim=-1)
        penalty = pen if self.penalty_coeff != 0 else None
        unpenalized, penalized = unpen, rew
        if return_single:                                                 # fake_env.py:121-127
            next_obs, return_means, return_stds = next_obs[0], return_means[0], return_stds[0]
            unpenalized, penalized, terminals = unpenalized[0], penalized[0], terminals[0]
        info = {'mean': return_means, 'std': return_stds, 'log_prob': lp, 'dev': dv,
                'unpenalized_rewards': unpenalized, 'penalty': penalty, 'penalized_rewards': penalized,
                'model_inds': inds_out}
        if is_np:
            cv = lambda t: None if t is None else t.cpu().numpy()
            info = {k: cv(v) for k, v in info.items()}
            return cv(next_obs), cv(penalized), cv(terminals), info
        return next_obs, penalized, terminals, info

    def close(self):
        pass
