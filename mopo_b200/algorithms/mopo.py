"""The slice of the MOPO algorithm class that is the hot path (reference mopo/algorithms/mopo.py):
_set_rollout_length (:675-687), _reallocate_model_pool (:689-711), _train_model (:712-720), _rollout_model (:723-765).

`ModelRollout` carries exactly the attributes those methods read on `self` in the reference
(_rollout_schedule, _epoch, _rollout_batch_size, _epoch_length, _model_train_freq,
_model_retain_epochs, _pool / sampler, fake_env, _model_pool, _rollout_random, _model_type), so a
maintainer can mix it into MOPO or call it standalone.  Two ways to supply actions:
  * `actions=[a_0 .. a_{h-1}]` (device tensors [>=n_i, act]) or a device policy -> ONE C call
    (mopo_rollout): h fused steps, compaction and pool writes with no host round trip;
  * `get_action_meta(obs, hidden)` host callback as in the reference (mopo.py:734): per-step loop.
"""
import ctypes as C
import os
import numpy as np
import torch

from .. import _lib
from ..replay_pools import SimpleReplayPool


class ModelRollout:
    def __init__(self, fake_env, env_pool, rollout_batch_size=100000, rollout_schedule=(20, 100, 1, 1),
                 epoch_length=1000, model_train_freq=250, model_retain_epochs=20, deterministic=False,
                 rollout_random=False, model_type='mlp', get_action_meta=None, policy=None, seed=0, identity_terminal=0):
        self.fake_env = fake_env
        self._pool = env_pool                       # env (offline data) pool; `sampler.random_batch` delegates to it
        self.sampler = env_pool                     # simple_sampler.py:103-108
        self._rollout_batch_size = int(rollout_batch_size)
        self._rollout_schedule = list(rollout_schedule)
        self._epoch_length = epoch_length
        self._model_train_freq = model_train_freq
        self._model_retain_epochs = model_retain_epochs
        self._deterministic = deterministic
        self._rollout_random = rollout_random
        self._model_type = model_type
        self._identity_terminal = identity_terminal          # mopo.py:743
        self._epoch = 0
        self._rollout_length = 1
        self._n_train_repeat = 1
        self._rollout_count = 0
        self.seed = int(seed)
        self._device_policy = policy                # mopo_b200.policies.GaussianPolicy: act = policy(obs) runs on the device
        if get_action_meta is not None:
            self.get_action_meta = get_action_meta
        elif policy is not None:
            self.get_action_meta = policy.get_action_meta

    def make_init_hidden(self, n):
        return None

    # ---- mopo.py:675-687 -----------------------------------------------------------------------
    def _set_rollout_length(self):
        min_epoch, max_epoch, min_length, max_length = self._rollout_schedule
        if self._epoch <= min_epoch:
            y = min_length
        else:
            dx = (self._epoch - min_epoch) / (max_epoch - min_epoch)
            dx = min(dx, 1)
            y = dx * (max_length - min_length) + min_length
        self._rollout_length = int(y)
        print('[ Model Length ] Epoch: {} (min: {}, max: {}) | Length: {} (min: {} , max: {})'.format(
            self._epoch, min_epoch, max_epoch, self._rollout_length, min_length, max_length))

    # ---- mopo.py:689-711 -----------------------------------------------------------------------
    def _reallocate_model_pool(self):
        obs_space = self._pool._observation_space
        act_space = self._pool._action_space
        rollouts_per_epoch = self._rollout_batch_size * self._epoch_length / self._model_train_freq
        model_steps_per_epoch = int(self._rollout_length * rollouts_per_epoch)
        new_pool_size = self._model_retain_epochs * model_steps_per_epoch
        if not hasattr(self, '_model_pool'):
            print('[ MOPO ] Initializing new model pool with size {:.2e}'.format(new_pool_size))
            self._model_pool = SimpleReplayPool(obs_space, act_space, new_pool_size, device=self.fake_env.model.device.index)
        elif self._model_pool._max_size != new_pool_size:
            print('[ MOPO ] Updating model pool | {:.2e} --> {:.2e}'.format(self._model_pool._max_size, new_pool_size))
            samples = self._model_pool.return_all_samples(as_numpy=False)
            new_pool = SimpleReplayPool(obs_space, act_space, new_pool_size, device=self.fake_env.model.device.index)
            new_pool.add_samples(samples)
            assert self._model_pool.size == new_pool.size
            self._model_pool = new_pool

    # ---- mopo.py:712-720 -----------------------------------------------------------------------
    def _train_model(self, **kwargs):
        """env pool -> (obs|act) / (rew|delta obs) -> BNN.train, all in HBM.  Under torch.distributed the optimiser loop is
        latency-bound, so rank 0 trains and the parameters + elites are broadcast (one NCCL call)."""
        from ..models.constructor import format_samples_for_training
        from .. import dist as mdist
        import torch.distributed as tdist
        if self._model_type == 'identity':
            print('[ MOPO ] Identity model, skipping model')
            return {}
        model = self.fake_env.model
        multi = tdist.is_available() and tdist.is_initialized() and tdist.get_world_size() > 1
        model_metrics = {}
        if not multi or tdist.get_rank() == 0:
            env_samples = self._pool.return_all_samples(as_numpy=False)
            train_inputs, train_outputs = format_samples_for_training(env_samples)
            model_metrics = model.train(train_inputs, train_outputs, **kwargs)
        if multi:
            mdist.broadcast_params(model, src=0)
        return model_metrics

    # ---- mopo.py:723-765 -----------------------------------------------------------------------
    def _rollout_model(self, rollout_batch_size, start_obs=None, actions=None, noise=None, model_inds=None,
                       precision=None, origin_base=0, quiet=False, **kwargs):
        if not quiet:
            print('[ Model Rollout ] Starting | Epoch: {} | Rollout length: {} | Batch size: {} | Type: {}'.format(
                self._epoch, self._rollout_length, rollout_batch_size, self._model_type))
        deterministic = kwargs.get('deterministic', self._deterministic)
        if start_obs is None:
            batch = self.sampler.random_batch(rollout_batch_size)           # mopo.py:727
            start_obs = batch['observations']                               # mopo.py:728
        if self._model_type == 'identity':
            steps_added = self._rollout_identity(start_obs, actions)
        elif actions is not None or self._device_policy is not None:
            steps_added = self._rollout_fused(start_obs, actions, noise, model_inds, deterministic, precision, origin_base)
        else:
            steps_added = self._rollout_callback(start_obs, deterministic, precision)
        mean_rollout_length = sum(steps_added) / rollout_batch_size
        rollout_stats = {'mean_rollout_length': mean_rollout_length}
        if not quiet:
            print('[ Model Rollout ] Added: {:.1e} | Model pool: {:.1e} (max {:.1e}) | Length: {} | Train rep: {}'.format(
                sum(steps_added), self._model_pool.size, self._model_pool._max_size, mean_rollout_length, self._n_train_repeat))
        self._last_steps_added = steps_added
        return rollout_stats

    def _rollout_fused(self, start_obs, actions, noise, model_inds, deterministic, precision, origin_base):
        env, model = self.fake_env, self.fake_env.model
        dev = model.device
        h = self._rollout_length
        # Host inputs (NumPy arrays / CPU tensors, as the reference's sampler and policy produce them) go to the C call as
        # they are: mopo_rollout_host stages them on the device itself (start states + step 0's actions on the compute
        # stream, later steps' actions on a copy stream behind per-step events), without per-array Python work.
        def host_f32(x):
            t = torch.as_tensor(x)
            if t.is_cuda:
                return None
            return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.to(torch.float32).contiguous()

        obs_h = host_f32(start_obs)
        acts_h = None
        if obs_h is not None and actions is not None:
            acts_h = [host_f32(a) for a in actions[:h]]
            if len(acts_h) < h or any(a is None or a.shape[0] < obs_h.shape[0] for a in acts_h):
                acts_h = None
        use_host = obs_h is not None and obs_h.dim() == 2 and (acts_h is not None or (actions is None and self._device_policy is not None))
        if os.environ.get('MOPO_NO_HOST_ROLLOUT') == '1':        # A/B switch: stage through torch copies instead
            use_host = False
        if use_host:
            obs, B = obs_h, obs_h.shape[0]
        else:
            obs = torch.as_tensor(start_obs).to(device=dev, dtype=torch.float32, non_blocking=True).contiguous()
            B = obs.shape[0]
        keep = [obs]

        def ptr_array(seq, dtype):
            if seq is None:
                return None
            ts = [torch.as_tensor(s).to(device=dev, dtype=dtype).contiguous() for s in seq]
            assert len(ts) >= h
            keep.extend(ts)
            return (C.c_void_p * h)(*[t.data_ptr() for t in ts[:h]])

        # Host-resident action arrays (NumPy / CPU tensors; pinned memory makes the copies asynchronous): step 0's go up
        # on the compute stream, the later steps' on a side stream with one event each, and the rollout waits for
        # event i only just before the first kernel that reads actions[i] -- the uploads overlap the earlier steps.
        act_events = None
        if not use_host and actions is not None and h > 1 and not all(torch.is_tensor(s) and s.is_cuda for s in actions[:h]):
            cur = torch.cuda.current_stream(dev)
            if getattr(self, '_copy_stream', None) is None:
                self._copy_stream = torch.cuda.Stream(dev)
            up, evs = [], []
            for i in range(h):
                src = torch.as_tensor(actions[i])
                if src.is_cuda or i == 0:
                    up.append(src.to(device=dev, dtype=torch.float32, non_blocking=True).contiguous())
                    evs.append(None)
                    continue
                with torch.cuda.stream(self._copy_stream):
                    t = src.to(device=dev, dtype=torch.float32, non_blocking=True).contiguous()
                    e = torch.cuda.Event()
                    e.record(self._copy_stream)
                t.record_stream(cur)
                up.append(t)
                evs.append(e)
            actions = up
            keep.extend(evs)
            act_events = (C.c_void_p * h)(*[e.cuda_event if e is not None else None for e in evs])

        steps = (C.c_int64 * max(h, 1))()
        act_ptrs = (C.c_void_p * h)(*[t.data_ptr() for t in acts_h]) if (use_host and acts_h is not None) else ptr_array(actions, torch.float32)
        a = _lib.RolloutArgs(start_obs=obs.data_ptr(), B=B, rollout_length=h, actions=act_ptrs,
                             term_kind=env.config.term_kind, penalty_kind=env.penalty_kind,
                             penalty_coeff=float(env.penalty_coeff), deterministic=int(bool(deterministic)),
                             precision=model.precision if precision is None else precision,
                             noise=ptr_array(noise, torch.float32), model_inds=ptr_array(model_inds, torch.int32),
                             seed=self.seed, rollout_id=self._rollout_count, origin_base=origin_base, steps_added=steps,
                             action_events=act_events,
                             policy=self._device_policy._handle if (actions is None and self._device_policy is not None) else None,
                             rollout_random=int(bool(self._rollout_random)), policy_deterministic=0)
        self._rollout_count += 1
        model._ensure_handle()
        call = _lib.load().mopo_rollout_host if use_host else _lib.load().mopo_rollout
        _lib.check(call(model._handle, self._model_pool._handle, C.byref(a), _lib.stream_ptr(dev)))
        if use_host:
            # the host sources of the asynchronous uploads stay alive until the rollout that consumes them has run (the
            # caller may drop or reuse its arrays as soon as this returns, and a freed pinned block is recycled at once)
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(dev))
            pending = [(e, t) for e, t in getattr(self, '_host_keep', []) if not e.query()]
            pending.append((done, (obs_h, acts_h)))
            self._host_keep = pending
        out = [int(steps[i]) for i in range(h)]
        while out and out[-1] == 0:          # the reference breaks out of the loop when nothing survives (mopo.py:754-756)
            out.pop()
        return out

    def _rollout_identity(self, obs, actions):
        """model_type == 'identity' (mopo.py:740-744): the model is bypassed -- next_obs = obs, zero rewards, terminals =
        identity_terminal -- but the loop, the pool appends and the early break are the reference's."""
        steps_added = []
        obs = np.asarray(obs.cpu() if torch.is_tensor(obs) else obs)
        for i in range(self._rollout_length):
            if actions is not None:
                act = np.asarray(actions[i].cpu() if torch.is_tensor(actions[i]) else actions[i])[:len(obs)]
            else:
                act, _ = self.get_action_meta(obs, self.make_init_hidden(1))
            if self._rollout_random:
                act = np.random.uniform(low=-1, high=1, size=np.shape(act))
            next_obs = obs
            rew = np.zeros((len(obs), 1))
            term = (np.ones((len(obs), 1)) * self._identity_terminal).astype(bool)
            steps_added.append(len(obs))
            self._model_pool.add_samples({'observations': obs, 'actions': act, 'next_observations': next_obs,
                                          'rewards': rew, 'terminals': term})
            nonterm_mask = ~term.squeeze(-1)
            if nonterm_mask.sum() == 0:
                break
            obs = next_obs[nonterm_mask]
        return steps_added

    def _rollout_callback(self, obs, deterministic, precision):
        """Literal per-step loop with a host policy callback (mopo.py:730-758)."""
        steps_added = []
        for i in range(self._rollout_length):
            hidden = self.make_init_hidden(1)
            act, hidden = self.get_action_meta(obs, hidden)
            if self._rollout_random:
                act = np.random.uniform(low=-1, high=1, size=act.shape)
            next_obs, rew, term, info = self.fake_env.step(obs, act, deterministic=deterministic, precision=precision,
                                                           diagnostics=False)
            steps_added.append(len(obs))
            self._model_pool.add_samples({'observations': obs, 'actions': act, 'next_observations': next_obs,
                                          'rewards': rew, 'terminals': term})
            nonterm_mask = ~term.squeeze(-1)
            if nonterm_mask.sum() == 0:
                break
            obs = next_obs[nonterm_mask]
        return steps_added
