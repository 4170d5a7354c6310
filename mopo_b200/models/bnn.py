"""BNN: the probabilistic ensemble behind construct_model (reference mopo/models/bnn.py).

Host-side mirror of the reference class for the hot path: layer stack (add/finalize), scaler,
max/min-logvar, elite state (_model_inds / random_inds, bnn.py:337-344), predict (bnn.py:508-546)
and the .mat weight I/O (bnn.py:276-281, 559-592).  All arithmetic is the CUDA library's
(mopo_ens_predict / mopo_env_step / mopo_rollout); there is no TF session and no CPU path.
`train` (bnn.py:369-503) runs the reference's loop with every Adam step on the device (csrc/train.cu).
"""
import ctypes as C
import os
import numpy as np
import torch

from .. import _lib
from .fc import FC


class TensorStandardScaler:
    """mopo/models/utils.py:45-128 (mu=0, sigma=1 until fit; sigma < 1e-12 -> 1)."""

    def __init__(self, x_dim):
        self.fitted = False
        self.mu = np.zeros((1, x_dim), np.float32)
        self.sigma = np.ones((1, x_dim), np.float32)
        self.cached_mu, self.cached_sigma = np.zeros([0, x_dim]), np.ones([1, x_dim])
        self._on_change = None

    def fit(self, data):
        if torch.is_tensor(data) and data.is_cuda:     # training set resident in HBM: reduce there (float64 accumulation)
            d64 = data.double()
            mu = d64.mean(0, keepdim=True).float().cpu().numpy()
            sigma = d64.std(0, unbiased=False, keepdim=True).float().cpu().numpy()
        else:
            data = data.detach().cpu().numpy() if torch.is_tensor(data) else np.asarray(data)
            mu = np.mean(data, axis=0, keepdims=True)
            sigma = np.std(data, axis=0, keepdims=True)
        sigma[sigma < 1e-12] = 1.0
        self.load(mu, sigma)
        self.fitted = True

    def load(self, mu, sigma):
        self.mu = np.asarray(mu, np.float32).reshape(1, -1)
        self.sigma = np.asarray(sigma, np.float32).reshape(1, -1)
        self.cache()
        if self._on_change:
            self._on_change()

    def transform(self, data):
        return (data - self.mu) / self.sigma

    def inverse_transform(self, data):
        return self.sigma * data + self.mu

    def get_vars(self):
        return [self.mu, self.sigma]

    def cache(self):
        self.cached_mu, self.cached_sigma = self.mu.copy(), self.sigma.copy()

    def load_cache(self):
        self.load(self.cached_mu, self.cached_sigma)


class BNN:
    def __init__(self, params):
        self.name = params.get('name', 'model')                        # bnn.py:42
        self.model_dir = params.get('model_dir', None)
        self.layers, self.var_layers = [], []
        self.finalized = False
        self.scaler = None
        self.max_logvar = self.min_logvar = None
        self.num_nets = params.get('num_networks', 1)
        self.num_elites = params['num_elites']
        self.separate_mean_var = params.get('separate_mean_var', False)
        self.deterministic = params.get('deterministic', False)
        self.model_loaded = False
        self.device = torch.device('cuda', params.get('device', torch.cuda.current_device() if torch.cuda.is_available() else 0))
        self.precision = params.get('precision', _lib.PREC_FP32)
        self.seed = int(params.get('seed', 0))
        self._handle = None
        self._rng_calls = 0
        if self.num_nets == 1:
            print("Created a neural network with variance predictions.")
        else:
            print("Created an ensemble of {} neural networks with variance predictions | Elites: {}".format(
                self.num_nets, self.num_elites))

    # ---- structure (bnn.py:100-165) ------------------------------------------------------
    @property
    def is_probabilistic(self): return True

    @property
    def is_tf_model(self): return False

    def add(self, layer, var_layer=False):
        if self.finalized:
            raise RuntimeError("Cannot modify network structure after finalizing.")
        tgt = self.var_layers if var_layer else self.layers
        if len(tgt) == 0 and layer.get_input_dim() is None:
            raise ValueError("Must set input dimension for the first layer.")
        layer = layer.copy()
        layer.set_ensemble_size(self.num_nets)
        if len(tgt) > 0 and not var_layer:
            layer.set_input_dim(tgt[-1].get_output_dim())
        tgt.append(layer)

    def pop(self):
        if len(self.layers) == 0:
            raise RuntimeError("Network is empty.")
        if self.finalized:
            raise RuntimeError("Cannot modify network structure after finalizing.")
        self.layers.pop()

    def __repr__(self):
        return 'BNN(name=%r, nets=%d, layers=%r, var_layers=%r)' % (self.name, self.num_nets, self.layers, self.var_layers)

    def finalize(self, optimizer=None, optimizer_args=None, *args, **kwargs):
        """bnn.py:166-274 minus the TF training graph: fix the head, construct variables with the
        reference init law, create the device ensemble, load `{model_dir}/{name}.mat` if requested."""
        if len(self.layers) == 0 or (self.separate_mean_var and len(self.var_layers) == 0):
            raise RuntimeError("Cannot finalize an empty network.")
        if self.finalized:
            raise RuntimeError("Can only finalize a network once.")
        hidden = self.layers[:-1]
        if any(l.activation != 'swish' for l in hidden) or len({l.output_dim for l in hidden}) > 1 or self.layers[-1].activation is not None:
            raise NotImplementedError("mopo_b200 builds the construct_model architectures only: 'mlp' (equal-width swish "
                                      "hidden layers + linear head) and 'linear' (the head alone)")
        if not self.separate_mean_var:                                  # bnn.py:183-189
            self.layers[-1].set_output_dim(2 * self.layers[-1].get_output_dim())
            self.end_act_name = self.layers[-1].get_activation(as_func=False)
            self.layers[-1].unset_activation()
            d1 = self.layers[-1].get_output_dim() // 2
        else:
            d1 = self.var_layers[-1].get_output_dim()
            if len(self.var_layers) != 1 or self.var_layers[0].get_input_dim() != hidden[-1].output_dim:
                raise NotImplementedError("separate_mean_var expects one linear var head on the last hidden activation")
        self.in_dim = self.layers[0].get_input_dim()
        self.out_dim = d1
        self.hidden_dim = hidden[0].output_dim if hidden else 1          # 'linear': no hidden layer (constructor.py:25-27)
        rng = np.random.RandomState(self.seed)
        self.scaler = TensorStandardScaler(self.in_dim)
        self.max_logvar = np.ones([1, d1], np.float32) / 2.                # bnn.py:196-199
        self.min_logvar = -np.ones([1, d1], np.float32) * 10.
        for layer in self.layers + self.var_layers:
            layer.construct_vars(rng)
        self._obs_dim = kwargs.get('obs_dim')          # set by construct_model; needed to split in_dim
        self.finalized = True
        if self.model_loaded:
            self.load_params()
        else:
            self._sync_device()
        self.scaler._on_change = self._sync_device

    # ---- device handle --------------------------------------------------------------------
    def _ensure_handle(self):
        if self._handle is not None:
            return
        lib = _lib.load()
        obs_dim = self._obs_dim if self._obs_dim is not None else self.out_dim - 1
        cfg = _lib.EnsCfg(obs_dim=obs_dim, act_dim=self.in_dim - obs_dim, rew_dim=self.out_dim - obs_dim,
                          hidden_dim=self.hidden_dim, n_hidden=len(self.layers) - 1, num_networks=self.num_nets,
                          num_elites=self.num_elites, separate_mean_var=int(self.separate_mean_var),
                          device=self.device.index or 0)
        h = C.c_void_p()
        _lib.check(lib.mopo_ens_create(C.byref(cfg), C.byref(h)))
        self._handle = h

    def _sync_device(self):
        """Push host parameters (reference layout) to the device images."""
        if not self.finalized:
            return
        self._ensure_handle()
        lib = _lib.load()
        keep = []

        def arr(a):
            a = np.ascontiguousarray(a, np.float32); keep.append(a); return a.ctypes.data

        n = len(self.layers)
        W = (C.c_void_p * n)(*[arr(l.weights) for l in self.layers])
        b = (C.c_void_p * n)(*[arr(l.biases) for l in self.layers])
        vw = arr(self.var_layers[0].weights) if self.separate_mean_var else None
        vb = arr(self.var_layers[0].biases) if self.separate_mean_var else None
        _lib.check(lib.mopo_ens_set_params(self._handle, W, b, vw, vb, arr(self.scaler.mu), arr(self.scaler.sigma),
                                           arr(self.max_logvar), arr(self.min_logvar), 0, _lib.stream_ptr(self.device)))
        if getattr(self, '_model_inds', None) is not None:
            self._push_elites()

    def _push_elites(self):
        inds = np.ascontiguousarray(self._model_inds_list, np.int32)
        _lib.check(_lib.load().mopo_ens_set_elites(self._handle, inds.ctypes.data_as(C.POINTER(C.c_int32)), len(inds)))

    def set_params(self, weights, biases, var_weights=None, var_biases=None, mu=None, sigma=None,
                   max_logvar=None, min_logvar=None):
        """Install parameters in the reference layout (what TF `var.load` does at bnn.py:276-281)."""
        for l, w, b in zip(self.layers, weights, biases):
            assert w.shape == (self.num_nets, l.input_dim, l.output_dim), (w.shape, l)
            l.weights = np.array(w, np.float32); l.biases = np.array(b, np.float32).reshape(self.num_nets, 1, -1)
        if self.separate_mean_var:
            self.var_layers[0].weights = np.array(var_weights, np.float32)
            self.var_layers[0].biases = np.array(var_biases, np.float32).reshape(self.num_nets, 1, -1)
        if max_logvar is not None: self.max_logvar = np.array(max_logvar, np.float32).reshape(1, -1)
        if min_logvar is not None: self.min_logvar = np.array(min_logvar, np.float32).reshape(1, -1)
        if mu is not None:
            self.scaler._on_change = None
            self.scaler.load(mu, sigma)
            self.scaler._on_change = self._sync_device
        self._sync_device()

    # ---- elites (bnn.py:337-344) ----------------------------------------------------------
    @property
    def _model_inds(self):
        if not hasattr(self, '_model_inds_list'):
            raise AttributeError("'BNN' object has no attribute '_model_inds' (set by _end_train, bnn.py:337-340)")
        return self._model_inds_list

    @_model_inds.setter
    def _model_inds(self, inds):
        self._model_inds_list = [int(i) for i in inds]
        if self._handle is not None:
            self._push_elites()

    def _end_train(self, holdout_losses):
        sorted_inds = np.argsort(holdout_losses)
        self._model_inds = sorted_inds[:self.num_elites].tolist()
        print('Using {} / {} models: {}'.format(self.num_elites, self.num_nets, self._model_inds))

    def random_inds(self, batch_size):
        inds = np.random.choice(self._model_inds, size=batch_size)
        return inds

    def reset(self):
        rng = np.random.RandomState(self.seed + 1 + self._rng_calls); self._rng_calls += 1
        for layer in self.layers + self.var_layers:
            layer.construct_vars(rng)
        self._sync_device()

    # ---- predict (bnn.py:508-546) -----------------------------------------------------------
    def predict(self, inputs, factored=False, *args, **kwargs):
        is_np = not torch.is_tensor(inputs)
        x = torch.as_tensor(inputs)
        if x.dim() not in (2, 3) or (x.dim() == 3 and x.shape[0] != self.num_nets) or x.shape[-1] != self.in_dim:
            raise ValueError("Invalid input dimension.")                   # fc.py:103
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        B = x.shape[-2]
        mean = torch.empty((self.num_nets, B, self.out_dim), dtype=torch.float32, device=self.device)
        var = torch.empty_like(mean)
        self._ensure_handle()
        lib, prec = _lib.load(), kwargs.get('precision', self.precision)
        if x.dim() == 3:                                                   # bnn.py:541-545: net e sees inputs[e]; always factored
            _lib.check(lib.mopo_ens_predict3d(self._handle, x.data_ptr(), B, mean.data_ptr(), var.data_ptr(), prec, _lib.stream_ptr(self.device)))
            return (mean.cpu().numpy(), var.cpu().numpy()) if is_np else (mean, var)
        _lib.check(lib.mopo_ens_predict(self._handle, x.data_ptr(), B, mean.data_ptr(), var.data_ptr(), prec, _lib.stream_ptr(self.device)))
        if not factored:                                                   # bnn.py:261-263
            m = mean.mean(0)
            mean, var = m, var.mean(0) + (mean - m).square().mean(0)
        if is_np:
            return mean.cpu().numpy(), var.cpu().numpy()
        return mean, var

    # ---- train (bnn.py:369-503) ---------------------------------------------------------------
    def _start_train(self):                                                # bnn.py:331-334
        self._snapshots = {i: (None, 1e10) for i in range(self.num_nets)}
        self._epochs_since_update = 0

    def _save_best(self, epoch, holdout_losses):                           # bnn.py:309-329
        updated = False
        lib, st = _lib.load(), _lib.stream_ptr(self.device)
        for i in range(len(holdout_losses)):
            current = holdout_losses[i]
            _, best = self._snapshots[i]
            improvement = (best - current) / best
            if improvement > 0.01:
                self._snapshots[i] = (epoch, current)
                _lib.check(lib.mopo_ens_train_snapshot(self._handle, i, st))    # _save_state(i), bnn.py:287-290
                updated = True
        if updated:
            self._epochs_since_update = 0
        else:
            self._epochs_since_update += 1
        return self._epochs_since_update > self._max_epochs_since_update

    # ---- device master images (in-place collectives: mopo_b200/dist.py) -----------------------------
    def _device_images(self):
        """torch views of the fp32 master images in HBM, fixed order: mu, sigma, (W_l, b_l) for l = 0..L (the head image is
        the fused [mean | logvar] head), max_logvar, min_logvar."""
        from ..replay_pools import _RawDevArray
        self._ensure_handle()
        lib = _lib.load()
        specs = [(_lib.IMG_MU, 0), (_lib.IMG_SIGMA, 0)]
        for l in range(len(self.layers)):
            specs += [(_lib.IMG_W, l), (_lib.IMG_B, l)]
        specs += [(_lib.IMG_MAX_LOGVAR, 0), (_lib.IMG_MIN_LOGVAR, 0)]
        views = []
        with torch.cuda.device(self.device):
            for which, layer in specs:
                ptr, cnt = C.c_void_p(), C.c_int64()
                _lib.check(lib.mopo_ens_param_image(self._handle, which, layer, C.byref(ptr), C.byref(cnt)))
                views.append(torch.as_tensor(_RawDevArray(ptr.value, (cnt.value,), '<f4'), device=self.device))
        return views

    def param_bytes(self):
        return 4 * sum(int(np.prod(np.shape(a))) for a in self._all_vars())

    def _after_device_update(self, sync_host=True):
        """The master images were overwritten on the device: rebuild the tensor-core operand images and (optionally) refresh
        the host mirrors that save() / the scaler read."""
        _lib.check(_lib.load().mopo_ens_refresh_images(self._handle, _lib.stream_ptr(self.device)))
        if sync_host:
            self._pull_params()
            v = self._device_images()
            self.scaler._on_change = None
            self.scaler.load(v[0].cpu().numpy().reshape(1, -1), v[1].cpu().numpy().reshape(1, -1))
            self.scaler._on_change = self._sync_device

    def _weight_decays(self):
        """Per-layer rates for the C trainer: hidden layers, mean head, log-variance head (fc.py:155-156)."""
        d = [l.weight_decay or 0.0 for l in self.layers]
        d.append((self.var_layers[0].weight_decay or 0.0) if self.separate_mean_var else d[-1])
        return (C.c_float * len(d))(*d)

    def _pull_params(self):
        """Device parameters -> the host FC layers (so save() and later scaler pushes see the trained weights)."""
        outs_w = [np.empty((self.num_nets, l.input_dim, l.output_dim), np.float32) for l in self.layers]
        outs_b = [np.empty((self.num_nets, 1, l.output_dim), np.float32) for l in self.layers]
        n = len(self.layers)
        W = (C.c_void_p * n)(*[a.ctypes.data for a in outs_w])
        b = (C.c_void_p * n)(*[a.ctypes.data for a in outs_b])
        vw = vb = None
        if self.separate_mean_var:
            vl = self.var_layers[0]
            vw = np.empty((self.num_nets, vl.input_dim, vl.output_dim), np.float32)
            vb = np.empty((self.num_nets, 1, vl.output_dim), np.float32)
        mx, mn = np.empty((1, self.out_dim), np.float32), np.empty((1, self.out_dim), np.float32)
        _lib.check(_lib.load().mopo_ens_get_params(self._handle, W, b, _lib.ptr(vw), _lib.ptr(vb), mx.ctypes.data,
                                                   mn.ctypes.data, _lib.stream_ptr(self.device)))
        for l, w, bb in zip(self.layers, outs_w, outs_b):
            l.weights, l.biases = w, bb
        if self.separate_mean_var:
            self.var_layers[0].weights, self.var_layers[0].biases = vw, vb
        self.max_logvar, self.min_logvar = mx, mn

    def train(self, inputs, targets, batch_size=32, max_epochs=None, max_epochs_since_update=5, hide_progress=False,
              holdout_ratio=0.0, max_logging=1000, max_grad_updates=None, timer=None, max_t=None, idxs=None,
              permutation=None):
        """BNN.train (bnn.py:369-503) with the optimiser loop on the device: the data set, the bootstrap index matrix
        and every Adam step stay in HBM; per epoch the host sees only the E holdout losses that drive _save_best.
        `permutation` / `idxs` (optional) pin the reference's np.random draws (bnn.py:393, 403) for parity tests;
        otherwise they come from a device generator seeded with the model's seed."""
        import itertools, time
        from collections import OrderedDict
        if len(self.layers) < 2:
            raise NotImplementedError("training the 'linear' model on the device is not built (predict / FakeEnv.step are)")
        self._max_epochs_since_update = max_epochs_since_update
        self._start_train()
        break_train = False
        dev = self.device
        lib, st = _lib.load(), _lib.stream_ptr(dev)
        gen = torch.Generator(device=dev); gen.manual_seed(self.seed + 7919 * (1 + self._rng_calls)); self._rng_calls += 1
        inputs = torch.as_tensor(inputs).to(device=dev, dtype=torch.float32)
        targets = torch.as_tensor(targets).to(device=dev, dtype=torch.float32)
        if inputs.dim() != 2 or inputs.shape[1] != self.in_dim or targets.shape != (inputs.shape[0], self.out_dim):
            raise ValueError("Invalid input dimension.")

        def shuffle_rows(arr):                                             # bnn.py:388-390
            order = torch.argsort(torch.rand(arr.shape, device=dev, generator=gen), dim=-1)
            return torch.gather(arr, 1, order.to(torch.int64)).contiguous()

        # Split into training and holdout sets (bnn.py:392-398)
        num_holdout = min(int(inputs.shape[0] * holdout_ratio), max_logging)
        if permutation is None:
            permutation = torch.randperm(inputs.shape[0], device=dev, generator=gen)
        permutation = torch.as_tensor(permutation).to(device=dev, dtype=torch.int64)
        inputs, holdout_inputs = inputs[permutation[num_holdout:]].contiguous(), inputs[permutation[:num_holdout]].contiguous()
        targets, holdout_targets = targets[permutation[num_holdout:]].contiguous(), targets[permutation[:num_holdout]].contiguous()
        N = inputs.shape[0]
        print('[ BNN ] Training {} | Holdout: {}'.format(tuple(inputs.shape), (self.num_nets,) + tuple(holdout_inputs.shape)))
        self.scaler.fit(inputs)                                            # bnn.py:401-402 (pushes mu / sigma to the device)

        if idxs is None:                                                   # bnn.py:403
            idxs = torch.randint(N, (self.num_nets, N), device=dev, generator=gen, dtype=torch.int32)
        idxs = torch.as_tensor(idxs).to(device=dev, dtype=torch.int32).contiguous()
        if idxs.shape[0] != self.num_nets or int(idxs.max()) >= N or int(idxs.min()) < 0:
            raise ValueError("idxs must be [num_nets, n] row indices into the training split")
        epoch_iter = range(max_epochs) if max_epochs is not None else itertools.count()

        # The optimiser loop runs on a side stream: its 35 small launches per batch are captured once into a CUDA graph
        # and replayed (stream capture is not possible on the legacy default stream).
        cur = torch.cuda.current_stream(dev)
        if getattr(self, '_train_stream', None) is None:
            self._train_stream = torch.cuda.Stream(dev)
        self._train_stream.wait_stream(cur)
        with torch.cuda.stream(self._train_stream):
            st = _lib.stream_ptr(dev)
            self._ensure_handle()
            for t_ in (inputs, targets, holdout_inputs, holdout_targets, idxs):    # allocated on the caller's stream, used on this one
                t_.record_stream(self._train_stream)
            idxs = idxs.clone()                                            # a buffer this call owns (shuffled in place every epoch)
            _lib.check(lib.mopo_ens_train_begin(self._handle, 1e-3 if not hasattr(self, '_lr') else self._lr, self._weight_decays()))
            _lib.check(lib.mopo_ens_train_set_deterministic(self._handle, int(bool(self.deterministic))))     # bnn.py:242-247
            mse_dev = torch.empty(2 * self.num_nets, dtype=torch.float32, device=dev)
            n_cols = idxs.shape[1]
            t0 = time.time()
            grad_updates = 0
            try:
                for epoch in epoch_iter:
                    _lib.check(lib.mopo_ens_train_batches(self._handle, inputs.data_ptr(), targets.data_ptr(), N, idxs.data_ptr(),
                                                          n_cols, n_cols, batch_size, None, st))        # bnn.py:421-427
                    grad_updates += int(np.ceil(n_cols / batch_size))
                    idxs.copy_(shuffle_rows(idxs))                             # bnn.py:429; in place: the captured graph is keyed on this buffer
                    if not hide_progress:
                        n_log = min(max_logging, n_cols)
                        _lib.check(lib.mopo_ens_train_mse(self._handle, inputs.data_ptr(), targets.data_ptr(), idxs.data_ptr(),
                                                          n_cols, n_log, mse_dev.data_ptr(), st))       # bnn.py:431-447
                        if holdout_ratio >= 1e-12 and num_holdout > 0:
                            _lib.check(lib.mopo_ens_train_mse(self._handle, holdout_inputs.data_ptr(), holdout_targets.data_ptr(),
                                                              None, 0, num_holdout, mse_dev[self.num_nets:].data_ptr(), st))
                            both = mse_dev.cpu().numpy()
                            losses, holdout_losses = both[:self.num_nets], both[self.num_nets:]
                            break_train = self._save_best(epoch, holdout_losses)                        # bnn.py:467
                        else:
                            losses = mse_dev[:self.num_nets].cpu().numpy()
                        self._last_train_losses = losses
                    t = time.time() - t0
                    if break_train or (max_grad_updates and grad_updates > max_grad_updates):
                        break
                    if max_t and t > max_t:
                        print('Breaking because of timeout: {}! (max: {})'.format(t, max_t))
                        break
                if timer: timer.stamp('bnn_train')
                _lib.check(lib.mopo_ens_train_restore(self._handle, st))       # _set_state, bnn.py:483 (KeyError -> RuntimeError)
                if timer: timer.stamp('bnn_set_state')
                if num_holdout > 0:                                            # bnn.py:486-492
                    _lib.check(lib.mopo_ens_train_mse(self._handle, holdout_inputs.data_ptr(), holdout_targets.data_ptr(),
                                                      None, 0, num_holdout, mse_dev.data_ptr(), st))
                    holdout_losses = mse_dev[:self.num_nets].cpu().numpy()
                else:
                    holdout_losses = np.full(self.num_nets, np.nan, np.float32)   # TF evaluates a mean over zero rows: nan
            finally:
                _lib.check(lib.mopo_ens_train_end(self._handle, st))
                self._pull_params()

        cur.wait_stream(self._train_stream)
        self._grad_updates = grad_updates
        if timer: timer.stamp('bnn_holdout')
        self._end_train(holdout_losses)
        if timer: timer.stamp('bnn_end')
        val_loss = (np.sort(holdout_losses)[:self.num_elites]).mean()
        model_metrics = {'val_loss': val_loss}
        print('[ BNN ] Holdout', np.sort(holdout_losses), model_metrics)
        return OrderedDict(model_metrics)

    # ---- .mat weight I/O (bnn.py:276-281, 559-592) --------------------------------------------
    def _all_vars(self):
        """nonoptvars + optvars order (bnn.py:224-225): [mu, sigma] + per layer [W, b] (+ var layers) + [max_lv, min_lv]."""
        v = [self.scaler.mu, self.scaler.sigma]
        for l in self.layers + self.var_layers:
            v += [l.weights, l.biases]
        return v + [self.max_logvar, self.min_logvar]

    def save(self, savedir, timestep):
        from scipy.io import savemat
        if not self.finalized:
            raise RuntimeError()
        model_dir = self.model_dir if savedir is None else savedir
        if not self.separate_mean_var:
            with open(os.path.join(model_dir, '{}_{}.nns'.format(self.name, timestep)), "w+") as f:
                for layer in self.layers[:-1]:
                    f.write("%s\n" % repr(layer))
                last = self.layers[-1].copy()
                last.set_activation(self.end_act_name)
                last.set_output_dim(last.get_output_dim() // 2)
                f.write("%s\n" % repr(last))
        else:
            with open(os.path.join(model_dir, '{}_{}.nns'.format(self.name, timestep)), "w+") as f:
                for layer in self.layers:
                    f.write("%s\n" % repr(layer))
            with open(os.path.join(model_dir, '{}_{}_var.nns'.format(self.name, timestep)), "w+") as f:
                for layer in self.var_layers:
                    f.write("%s\n" % repr(layer))
        savemat(os.path.join(model_dir, '{}_{}.mat'.format(self.name, timestep)),
                {str(i): v for i, v in enumerate(self._all_vars())})

    def load_params(self, path=None):
        from scipy.io import loadmat
        d = loadmat(path or os.path.join(self.model_dir, "%s.mat" % self.name))
        n_l = len(self.layers)
        vals = [np.asarray(d[str(i)], np.float32) for i in range(2 + 2 * (n_l + len(self.var_layers)) + 2)]
        mu, sigma = vals[0], vals[1]
        Ws = vals[2:2 + 2 * n_l:2]; bs = vals[3:3 + 2 * n_l:2]
        k = 2 + 2 * n_l
        vw = vb = None
        if self.separate_mean_var:
            vw, vb = vals[k], vals[k + 1]; k += 2
        self.set_params(Ws, bs, vw, vb, mu, sigma, vals[k], vals[k + 1])

    def __del__(self):
        try:
            if self._handle is not None:
                _lib.load().mopo_ens_destroy(self._handle)
                self._handle = None
        except Exception:
            pass
